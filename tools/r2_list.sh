#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; T=${1:-l}
timeout 300 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_${T}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/r2_${T}_launches.csv \
  env CMSBWT_PROFILE=1 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_${T}_ncu_list.log 2>&1
echo rc=$?
