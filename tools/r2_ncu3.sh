#!/bin/bash
# full ncu capture of three kernels of one bench step (after the same command ran without ncu)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 300 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_ncu3_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k 'regex:k_tile_scatter|k_fill_pull|k_tile_rank|k_ms_fast' -c 4 -o $O/r2_ncu3 \
  env CMSBWT_PROFILE=1 python bench.py --steps 1 --no-e2e --no-cpu-baseline > $O/r2_ncu3.log 2>&1
echo "ncu rc=$?"; ls -la $O | grep ncu3
