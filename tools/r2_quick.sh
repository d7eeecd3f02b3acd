#!/bin/bash
# quick GPU check: parity file + small full-size pins + a short bench line (group times)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; T=${1:-q}
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2_${T}_parity.log 2>&1; prc=$?; tail -4 $O/r2_${T}_parity.log
if [ $prc -ne 0 ]; then tail -60 $O/r2_${T}_parity.log; exit 1; fi
timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu -k "config1 or config2 or config4_at or config5_at_1_400" > $O/r2_${T}_full_small.log 2>&1; echo "fullsize small rc=$?"; tail -3 $O/r2_${T}_full_small.log
for wl in config3 ${EXTRA_WL}; do
timeout 300 python bench.py --workload $wl --steps 10 --no-e2e --no-cpu-baseline > $O/r2_${T}_bench_$wl.json 2> $O/r2_${T}_bench_$wl.err; echo "bench $wl rc=$?"
python -c "import json; d=json.loads(open('$O/r2_${T}_bench_$wl.json').read().strip().splitlines()[-1]); print('$wl', round(d['value'],1), round(d['ms_per_step'],3), {k:{g:round(x,3) for g,x in v['groups_ms'].items()} for k,v in d['roofline']['stages'].items()})"
done
