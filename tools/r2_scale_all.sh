#!/bin/bash
# sharded tests (when 2 GPUs are there) + the strong-scaling bench line at N GPUs: tools/r2_scale_all.sh <N>
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; N=$1
if [ "$N" = 2 ]; then
  timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu > $O/r2_final_sharded_2gpu.log 2>&1; echo "sharded rc=$?"; tail -3 $O/r2_final_sharded_2gpu.log
fi
bash tools/r2_scale.sh $N final_bench
