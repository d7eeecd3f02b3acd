#!/bin/bash
# A/B of k_msf_gather lanes per window, then the parity file on the default
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
for v in 8 16 32; do
  CMSBWT_GATHER_LANES=$v timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_gab_$v.json 2>/dev/null
  python -c "import json; d=json.loads(open('$O/r2_gab_$v.json').read().strip().splitlines()[-1]); print('lanes $v', round(d['ms_per_step'],3), {g:round(x,3) for g,x in d['roofline']['stages']['stage1_ms']['groups_ms'].items()})"
done
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_escapes.py -x -q -m gpu 2>&1 | tail -2
