#!/bin/bash
# strong scaling of config 3 on N GPUs of one box: tools/r2_scale.sh <N> [tag]
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; N=$1; T=${2:-scale}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 > $O/r2_${T}_n$N.json 2> $O/r2_${T}_n$N.err; echo "bench n$N rc=$?"
python -c "
import json
d=json.loads(open('$O/r2_${T}_n$N.json').read().strip().splitlines()[-1])
print('N=$N', round(d['value'],1), 'Gbp/s', round(d['ms_per_step'],3), 'ms  e2e', round(d['e2e']['value'],1), d['config']['phase_ms_rank0'], 'collectives', d['config']['collectives_per_step'], 'stage', d['config']['stage_ms_rank0'])"
tail -3 $O/r2_${T}_n$N.err
