#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_escapes.py tests/test_gpu_consumers.py -x -q -m gpu > $O/r2_c3_new.log 2>&1; echo "new tests rc=$?"; tail -30 $O/r2_c3_new.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2_c3_parity.log 2>&1; echo "parity rc=$?"; tail -5 $O/r2_c3_parity.log
bash tools/r2_ncu.sh c3 "k_tile_rank|k_tile_scatter|k_ms_fast|k_vrank" 4
