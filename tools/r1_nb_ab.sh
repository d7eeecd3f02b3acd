#!/bin/bash
# A/B of the empty-bucket slots (ms_variant bit 16): bench with and without, then parity with it on.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
for v in 26 10; do
  CMSBWT_MS_VARIANT=$v timeout 40 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null > $O/r1_v8_bench_variant$v.json
  python -c "import json; d=json.loads(open('$O/r1_v8_bench_variant$v.json').read().strip().splitlines()[-1]); print($v, d['value'], d['ms_per_step'], d['config']['stage_ms']['ms_kernel_ms'])"
done
CMSBWT_MS_VARIANT=26 timeout 60 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu \
  -k "golden_fixtures or fuzz_vs_oracle or config3_full or config2_full" > $O/r1_v8_nb_tests.log 2>&1
echo "tests rc=$?" >> $O/r1_v8_nb_tests.log
tail -3 $O/r1_v8_nb_tests.log
