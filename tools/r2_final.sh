#!/bin/bash
# Round 2 final evidence on one B200: the whole -m gpu suite on the shipped commit, smoke(), the bench line
# (with e2e and the same-work CLI comparison), the launch list and full ncu captures of the main kernels.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
echo "commit $(cat .commit_id 2>/dev/null)" > $O/r2_final_gpu_suite.log
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader >> $O/r2_final_gpu_suite.log
timeout 1500 python -m pytest tests -x -q -m gpu --durations=15 >> $O/r2_final_gpu_suite.log 2>&1; src=$?; echo "suite rc=$src" >> $O/r2_final_gpu_suite.log; tail -22 $O/r2_final_gpu_suite.log
if [ $src -ne 0 ]; then tail -80 $O/r2_final_gpu_suite.log; exit 1; fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r2_final_smoke.log
timeout 600 python bench.py > $O/r2_final_bench_config3.json 2> $O/r2_final_bench.err; echo "bench rc=$?"; tail -c 1500 $O/r2_final_bench_config3.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_final_reference_arm.json 2>/dev/null; echo "reference arm rc=$?"
timeout 300 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_final_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/r2_final_launches_config3.csv \
  env CMSBWT_PROFILE=1 python bench.py --steps 2 --no-e2e --no-cpu-baseline > $O/r2_final_ncu_list.log 2>&1
ncu --set full --clock-control none --profile-from-start off \
  -k 'regex:k_ms_fast|k_tile_rank|k_tile_scatter|k_fill_pull|k_pack|k_vrank|k_tail_label|k_tr_rank0|k_msf_gather|k_tile_hist|k_link|k_rs_pass<unsigned int|k_tr_group_cls' \
  -c 24 -o $O/r2_final_kernels_full env CMSBWT_PROFILE=1 python bench.py --steps 1 --no-e2e --no-cpu-baseline > $O/r2_final_ncu_full.log 2>&1
echo "ncu rc=$?"
# the table is made here too: gpurun brings back at most 64 MiB
ncu -i $O/r2_final_kernels_full.ncu-rep --page raw --csv > $O/r2_final_kernels_raw.csv 2>/dev/null
sz=$(du -sm $O | cut -f1); echo "gpurun_out: $sz MiB"
if [ "$sz" -gt 58 ]; then rm -f $O/r2_final_kernels_full.ncu-rep; echo "dropped the .ncu-rep (too large to bring back)"; fi
ls -la $O | grep r2_final
