#!/usr/bin/env python3
"""Stage 1 on a reference with N blocks: the 2-bit kernel with the reference's run list against the general 8-bit
kernel (set_param fast = 0), same collection, same bytes.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE); sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tests"))
import cms_bwt_b200 as P          # noqa: E402
import gen_synth                  # noqa: E402
import oracle_lib as O            # noqa: E402

ref, copies = gen_synth.make_collection(2000000, 20, 0.001, 7)
ref = ref.copy()
rng = np.random.default_rng(1)
for a in rng.integers(0, len(ref) - 20000, 10):
    ref[a:a + 10000] = ord("N")
ref[rng.integers(0, len(ref), 200)] = ord("N")
x = P.augment(ref)
text = O.docs_to_text(copies)
ctx = P.Context(0)
ctx.index_upload(x)
out = {}
for fast in (1, 0):
    ctx.set_param("fast", fast)
    best = None
    for _ in range(3):
        bwt = ctx.run(text, P.MODE_PLAIN).copy()
        st = ctx.stats()
        best = st["ms_ms"] if best is None else min(best, st["ms_ms"])
    out["fast=%d" % fast] = {"fast_path": st["fast_path"], "ms_stage_ms": round(best, 3), "heads": st["n_heads"]}
    if fast:
        keep = bwt
    else:
        out["identical_bytes"] = bool(np.array_equal(keep, bwt))
out["workload"] = "2 Mbp reference with 10 N blocks of 10 kbp + 200 single N, 20 copies at 0.1 %% (%d symbols)" % len(text)
print(json.dumps(out))
