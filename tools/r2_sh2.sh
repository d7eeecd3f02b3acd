#!/bin/bash
# 2-GPU check of the multi-GPU path: sharded tests (native host layer, NCCL, CLI --gpus) + the N=2 bench line
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; T=${1:-sh2}
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu --durations=5 > $O/r2_${T}_sharded.log 2>&1; rc=$?; echo "sharded rc=$rc"; tail -12 $O/r2_${T}_sharded.log
if [ $rc -ne 0 ]; then tail -60 $O/r2_${T}_sharded.log; exit 1; fi
bash tools/r2_scale.sh 2 $T
