#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu --durations=5 -k "nccl or native or cli_gpus or cli_multi" > $O/r2_sh_sharded.log 2>&1; echo "sharded rc=$?"; tail -25 $O/r2_sh_sharded.log
