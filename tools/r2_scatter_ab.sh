#!/bin/bash
# A/B of k_tile_scatter variants (heads per thread, resident CTAs): group times of the config-3 step
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
for v in 0 1 2 3 4; do
  CMSBWT_SCATTER_VARIANT=$v timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_scab_$v.json 2>/dev/null
  python -c "import json; d=json.loads(open('$O/r2_scab_$v.json').read().strip().splitlines()[-1]); print('variant $v', round(d['ms_per_step'],3), {g:round(x,3) for g,x in d['roofline']['stages']['stage2_head_order']['groups_ms'].items()})"
done
