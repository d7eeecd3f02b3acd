#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_escapes.py -x -q -m gpu > $O/r2_q5_parity.log 2>&1; prc=$?; tail -4 $O/r2_q5_parity.log
if [ $prc -ne 0 ]; then tail -60 $O/r2_q5_parity.log; exit 1; fi
for v in 1 0; do
CMSBWT_PACK_TMA=$v timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_q5_bench_packtma$v.json 2>/dev/null
python -c "import json; d=json.loads(open('$O/r2_q5_bench_packtma$v.json').read().strip().splitlines()[-1]); print('pack_tma=$v', round(d['ms_per_step'],3), {k:{g:round(x,3) for g,x in v['groups_ms'].items()} for k,v in d['roofline']['stages'].items()})"
done
for wl in config2 config5_4; do
timeout 300 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu-baseline > $O/r2_q5_bench_$wl.json 2>/dev/null
python -c "import json; d=json.loads(open('$O/r2_q5_bench_$wl.json').read().strip().splitlines()[-1]); print('$wl', round(d['value'],1), round(d['ms_per_step'],3), {k:{g:round(x,3) for g,x in v['groups_ms'].items()} for k,v in d['roofline']['stages'].items()})"
done
