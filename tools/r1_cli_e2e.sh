#!/bin/bash
# Round-1 check of the streamed ingest/egress on the GPU box: new parity tests, then the
# cms_bwt command line on config 2 (5 Mbp x 100, 0.5 Gbp) with and without --no-stream, then the bench.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
[ "${SKIP_TESTS:-0}" = 1 ] || timeout 420 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -x -q -m gpu \
  -k "streamed_api or cli_binary or cli_multi" > $O/r1_v6_tests.log 2>&1
echo "tests rc=$?" >> $O/r1_v6_tests.log
tail -3 $O/r1_v6_tests.log
python tools/gen_synth.py 5000000 100 0.001 11 /dev/shm/c2
nproc > $O/r1_v6_cli_config2.log
for rep in 1 2; do
  for opt in "" "--no-stream" "--no-host-sa" "-r"; do
    echo "=== cms_bwt $opt (rep $rep)" >> $O/r1_v6_cli_config2.log
    rm -f /dev/shm/out$rep.*
    timeout 120 cms_bwt_b200/cms_bwt $opt -o /dev/shm/out$rep /dev/shm/c2.txt | grep "Time elapsed" >> $O/r1_v6_cli_config2.log
    cat /dev/shm/out$rep.log >> $O/r1_v6_cli_config2.log
    [ $rep = 1 ] && md5sum /dev/shm/out$rep.*bwt >> $O/r1_v6_cli_config2.log
  done
done
rm -f /dev/shm/out* /dev/shm/c2.*
if [ "${FULL_SUITE:-0}" = 1 ]; then
  timeout 330 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_parity.py::test_cli_binary_matches_golden \
    --durations=8 > $O/r1_v6_gpu_suite.log 2>&1
  echo "suite rc=$?" >> $O/r1_v6_gpu_suite.log
  tail -4 $O/r1_v6_gpu_suite.log
fi
if [ "${BENCH:-0}" = 1 ]; then
  timeout 300 python bench.py > $O/r1_v6_bench_config3.json 2> $O/r1_v6_bench.err
  echo "bench rc=$?"
  tail -c 600 $O/r1_v6_bench_config3.json
fi
