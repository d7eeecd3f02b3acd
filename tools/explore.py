"""Stage timings of the CUDA path on a synthetic workload (exploration helper)."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import cms_bwt_b200 as P
import gen_device

ref_bp = int(float(sys.argv[1])); copies = int(sys.argv[2]); div = float(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
params = dict(a.split("=") for a in sys.argv[5:])
dev = torch.device("cuda:0")
t = time.time()
ref, text, N, nd = gen_device.make_text(ref_bp, copies, div, 13, dev)
torch.cuda.synchronize()
print("gen %.2fs N=%d docs=%d" % (time.time() - t, N, nd), flush=True)
x = P.augment(ref.cpu().numpy())
ctx = P.Context(0)
for k, v in params.items():
    ctx.set_param(k, int(v))
t = time.time(); ctx.index_upload(x); print("index_upload %.2fs" % (time.time() - t), flush=True)
for r in range(reps):
    t = time.time()
    ctx.ms_run_device(text.data_ptr(), N)
    ctx.heads_sort()
    ctx.emit(P.MODE_PLAIN)
    wall = time.time() - t
    s = ctx.stats()
    tot = s["ms_ms"] + s["sort_ms"] + s["emit_ms"]
    print(json.dumps(dict(wall_ms=wall * 1e3, gbps=(N - nd) / 1e9 / (tot / 1e3), **s)), flush=True)
out = torch.empty(0)
