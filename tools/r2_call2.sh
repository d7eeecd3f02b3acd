#!/bin/bash
# Round 2, GPU call 2 (2 GPUs): C++ NCCL host layer (tests + N=2 bench), k_fill_pull v2 at N=1
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_ordering.py -x -q -m gpu --durations=5 > $O/r2_c2_sharded.log 2>&1; echo "sharded rc=$?"; tail -15 $O/r2_c2_sharded.log
timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu -k "config1 or config2 or config4_at or config5_at_1_400" > $O/r2_c2_full_small.log 2>&1; echo "fullsize small rc=$?"; tail -5 $O/r2_c2_full_small.log
timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_c2_bench_n1.json 2> $O/r2_c2_bench_n1.err; echo "bench n1 rc=$?"
python -c "import json; d=json.loads(open('$O/r2_c2_bench_n1.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:v['groups_ms'] for k,v in d['roofline']['stages'].items()})"
CMSBWT_FILL_STORE=0 timeout 300 python bench.py --steps 10 --no-e2e --no-cpu-baseline > $O/r2_c2_bench_n1_nobulk.json 2>/dev/null
python -c "import json; d=json.loads(open('$O/r2_c2_bench_n1_nobulk.json').read().strip().splitlines()[-1]); print('no bulk store:', d['ms_per_step'], d['roofline']['stages']['stage3_emit']['groups_ms'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r2_c2_bench_n2.json 2> $O/r2_c2_bench_n2.err; echo "bench n2 rc=$?"; tail -c 2500 $O/r2_c2_bench_n2.json; tail -5 $O/r2_c2_bench_n2.err
