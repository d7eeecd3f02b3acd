#!/bin/bash
# A/B of NCCL point-to-point channel counts on the N-GPU bench: tools/r2_nccl_ab.sh <N>
cd "${GRAFT_REPO_ROOT:-/root/repo}"
N=${1:-2}
for v in default 8 16 32; do
  if [ $v = default ]; then unset NCCL_MIN_P2P_NCHANNELS; else export NCCL_MIN_P2P_NCHANNELS=$v; fi
  echo "== NCCL_MIN_P2P_NCHANNELS=$v"
  bash tools/r2_scale.sh $N ab_$v 2>&1 | grep "^N="
done
