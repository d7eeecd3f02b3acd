#!/bin/bash
# A/B of one environment switch on the config-3 step: tools/r2_env_ab.sh NAME v1 v2 ...
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; NAME=$1; shift
for v in "$@"; do
  env $NAME=$v timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_env_${NAME}_$v.json 2>/dev/null
  python -c "import json; d=json.loads(open('$O/r2_env_${NAME}_$v.json').read().strip().splitlines()[-1]); print('$NAME=$v', round(d['ms_per_step'],3), {k:{g:round(x,3) for g,x in s['groups_ms'].items()} for k,s in d['roofline']['stages'].items()})"
done
