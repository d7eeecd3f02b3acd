"""Stream-ordering contract of the boundary (include/cmsbwt_b200.h, "STREAM ORDERING"): a device text
that the caller is still producing on ITS stream when cmsbwt_ms_run_device[_ordered] / run_sharded is
called must be complete before the library (which runs on its own non-blocking stream) reads it.
Round 1 lost this race on the driver's box (What's weak #2).  Bit-exact against the oracle."""
import numpy as np
import pytest
import torch

import cms_bwt_b200 as P
from cms_bwt_b200 import multi
import oracle_lib as O
import gen_synth

pytestmark = pytest.mark.gpu


def _case():
    ref, copies = gen_synth.make_collection(30000, 6, 0.002, 21)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    return x, text, O.Oracle(x).bwt(text)


def _busy(dev):
    """~20-50 ms of work queued on torch's current stream"""
    a = torch.empty(1 << 28, dtype=torch.float32, device=dev)
    for _ in range(6):
        a.normal_()
    return a


@pytest.mark.parametrize("ordered", [False, True])
def test_text_written_late_on_the_callers_stream(ordered):
    dev = torch.device("cuda:0")
    x, text, want = _case()
    N = len(text)
    host = torch.from_numpy(np.concatenate([text, np.full(64, 2, np.uint8)])).pin_memory()
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        for it in range(20):
            d = torch.zeros(N + 64, dtype=torch.uint8, device=dev)       # stale contents: all zero
            keep = _busy(dev)                                            # the copy below queues behind this
            d.copy_(host, non_blocking=True)
            if ordered:
                ctx.ms_run_device_ordered(d.data_ptr(), N, torch.cuda.current_stream(dev).cuda_stream)
            else:
                ctx.ms_run_device(d.data_ptr(), N)
            ctx.heads_sort()
            nb = ctx.emit(P.MODE_PLAIN)
            got = ctx.result_download().copy()
            assert nb == N and np.array_equal(got, want), "iteration %d (ordered=%s)" % (it, ordered)
            del keep, d
    finally:
        ctx.close()


def test_run_sharded_right_after_producing_the_shard():
    """the failing leg of round 1: torch ops build the shard, run_sharded follows immediately"""
    dev = torch.device("cuda:0")
    x, text, want = _case()
    N = len(text)
    nd = int((text == 2).sum())
    host = torch.from_numpy(np.concatenate([text, np.full(64, 2, np.uint8)])).pin_memory()
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        shared = multi.ThreadComm.Shared(1)
        for it in range(20):
            d = torch.zeros(N + 64, dtype=torch.uint8, device=dev)
            keep = _busy(dev)
            d.copy_(host, non_blocking=True)
            sp, out = multi.run_sharded(ctx, dev, d.data_ptr(), N, nd, 0, 1, comm=multi.ThreadComm(shared, 0))
            assert np.array_equal(out.cpu().numpy(), want), "iteration %d" % it
            del keep, d
    finally:
        ctx.close()
