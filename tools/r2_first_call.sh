#!/bin/bash
# First GPU call of round 2 (about 3 minutes of box time): what round 1 left unmeasured.
#  1. whole -m gpu suite with the v8 default (empty-bucket slots on)
#  2. k-mer order 13 against the rule's 12 at config 3, now that empty buckets are resolved from the slot
#  3. bench line, ncu launch list and one full capture of k_ms_fast for profiles/traffic.json
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 400 python -m pytest tests -x -q -m gpu --durations=5 > $O/r2_gpu_suite.log 2>&1; echo "suite rc=$?" >> $O/r2_gpu_suite.log; tail -3 $O/r2_gpu_suite.log
for k in 13 12; do
  CMSBWT_PARAMS="kmer_k=$k" timeout 60 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null > $O/r2_bench_k$k.json
  python -c "import json; d=json.loads(open('$O/r2_bench_k$k.json').read().strip().splitlines()[-1]); print($k, d['value'], d['ms_per_step'], d['config']['stage_ms'])"
done
timeout 200 python bench.py > $O/r2_bench_config3.json 2> $O/r2_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_config3.csv \
  python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > $O/r2_ncu_list.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ms_fast -c 1 -o $O/r2_kmsfast_full \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/r2_ncu_full.log 2>&1
ls -la $O | tail -8
