#!/bin/bash
# Round 2, GPU call 1: the whole -m gpu suite on the shipped kernels (oracle-compared tests first, the
# full-size pins last), a bench line, the launch list, and full ncu captures of the main kernels.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > $O/r2_c1_host.txt; nproc >> $O/r2_c1_host.txt; free -g | head -2 >> $O/r2_c1_host.txt
grep -m1 "model name" /proc/cpuinfo >> $O/r2_c1_host.txt
git rev-parse HEAD 2>/dev/null >> $O/r2_c1_host.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2_c1_parity_first.log 2>&1; prc=$?; tail -5 $O/r2_c1_parity_first.log
if [ $prc -ne 0 ]; then echo "parity failed (rc=$prc): stopping before bench/ncu"; tail -60 $O/r2_c1_parity_first.log; exit 1; fi
timeout 1500 python -m pytest tests -x -q -m gpu --durations=12 > $O/r2_c1_gpu_suite.log 2>&1; src=$?; echo "suite rc=$src" >> $O/r2_c1_gpu_suite.log; tail -25 $O/r2_c1_gpu_suite.log
if [ $src -ne 0 ]; then tail -80 $O/r2_c1_gpu_suite.log; fi
timeout 600 python bench.py > $O/r2_c1_bench_config3.json 2> $O/r2_c1_bench.err; echo "bench rc=$?"; tail -c 3000 $O/r2_c1_bench_config3.json
timeout 300 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_c1_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/r2_c1_launches_config3.csv \
  env CMSBWT_PROFILE=1 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_c1_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k 'regex:k_ms_fast|k_tile_place|k_tile_scatter|k_fill|k_pack|k_vrank|k_tail_label|k_tr_rank0|k_msf_gather|k_tile_cov|k_tile_hist|k_kappa|k_link' \
  -c 18 -o $O/r2_c1_kernels_full env CMSBWT_PROFILE=1 python bench.py --steps 1 --no-e2e --no-cpu-baseline > $O/r2_c1_ncu_full.log 2>&1
echo "ncu rc=$?"; ls -la $O | tail -12
