#!/bin/bash
# bench lines of the other shapes on one GPU (compute-sanitizer is closed on this pool, so no memcheck leg)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
for wl in config2 config5_4; do
  timeout 300 python bench.py --workload $wl --steps 10 --no-e2e --no-cpu-baseline > $O/r2_final_bench_$wl.json 2> $O/r2_final_bench_$wl.err; echo "bench $wl rc=$?"
  python -c "import json; d=json.loads(open('$O/r2_final_bench_$wl.json').read().strip().splitlines()[-1]); print('$wl', round(d['value'],1), round(d['ms_per_step'],3))"
done
