#!/bin/bash
# ncu --set full capture of selected kernels of one bench step: tools/r2_ncu.sh <tag> <regex> <count>
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; T=$1
timeout 300 python bench.py --steps 1 --no-e2e --no-cpu-baseline > $O/r2_${T}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k "regex:$2" -c $3 -o $O/r2_${T}_full \
  env CMSBWT_PROFILE=1 python bench.py --steps 1 --no-e2e --no-cpu-baseline > $O/r2_${T}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la $O/r2_${T}_full.ncu-rep
