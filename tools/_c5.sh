run() { CMSBWT_PARAMS="$2" timeout 300 python bench.py --workload $1 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['config']['stage_ms']; print('$1 $2', round(d['ms_per_step'],2), 'sort', round(s['sort_ms'],2))"; }
for mb in 64 128 256 512 4096; do run config5_16 "scatter_mb=$mb"; done
for mb in 48 64 96 128; do run config3 "scatter_mb=$mb"; done
