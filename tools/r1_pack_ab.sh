#!/bin/bash
# A/B of the two 2-bit packing kernels (CMSBWT_PACK=1: k_pack16, 0: k_pack): parity with the new one, then the bench with each.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu \
  -k "golden_fixtures or fuzz_vs_oracle or config1_scaled or many_short or bad_symbol or config3_full or config2_full" > $O/r1_v7_pack_tests.log 2>&1
echo "tests rc=$?" >> $O/r1_v7_pack_tests.log
tail -3 $O/r1_v7_pack_tests.log
CMSBWT_PACK=1 timeout 120 python bench.py --steps 10 --warmup 3 > $O/r1_v7_bench_pack16.json 2> $O/r1_v7_bench_pack16.err
CMSBWT_PACK=0 timeout 120 python bench.py --steps 10 --warmup 3 > $O/r1_v7_bench_pack32.json 2> $O/r1_v7_bench_pack32.err
python - <<'PY'
import json
for n in ("pack16", "pack32"):
    try:
        d = json.load(open("gpurun_out/r1_v7_bench_%s.json" % n))
        print(n, round(d["ms_per_step"], 3), d["config"]["stage_ms"])
    except Exception as e:
        print(n, "failed", e)
PY
