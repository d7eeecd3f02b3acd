"""cmsbwt_rle_encode_static: run-length records of a device-resident BWT with no allocation in the library
(scratch in a file-scope __device__ array, include/cmsbwt_b200.h) against the existing entry point
(cmsbwt_run / cmsbwt_emit in CMSBWT_MODE_RLE), bit for bit, on calls in a row that share the scratch.
Collected last (tests/conftest.py)."""
import numpy as np
import pytest

import cms_bwt_b200 as P
import oracle_lib as O
import gen_synth

pytestmark = pytest.mark.gpu


def _rle_numpy(b):
    """maximal-run records (LE u64 length, u8 symbol) of a byte array"""
    if len(b) == 0:
        return np.zeros(0, np.uint8)
    st = np.flatnonzero(np.concatenate([[True], b[1:] != b[:-1]]))
    ln = np.diff(np.concatenate([st, [len(b)]])).astype("<u8")
    rec = np.zeros((len(st), 9), np.uint8)
    rec[:, :8] = ln.view(np.uint8).reshape(-1, 8)
    rec[:, 8] = b[st]
    return rec.reshape(-1)


def _case(ref_bp, copies, seed):
    ref, cps = gen_synth.make_collection(ref_bp, copies, 0.003, seed)
    return P.augment(ref), O.docs_to_text([np.zeros(0, np.uint8)] + cps)


def test_static_scratch_rle_matches_the_existing_entry_point():
    import torch
    dev = torch.device("cuda:0")
    ctx = P.Context(0)
    try:
        # two inputs of different size (the second leaves stale entries of the first in the scratch), then the first again
        cases = [_case(60000, 9, 5), _case(9000, 3, 6)]
        want = []
        for x, text in cases:
            ctx.index_upload(x)
            bwt = ctx.run(text, P.MODE_PLAIN).copy()
            rl = ctx.run(text, P.MODE_RLE).copy()            # the existing entry point
            assert np.array_equal(rl, _rle_numpy(bwt))
            want.append((bwt, rl))
        cap = max(len(rl) for _, rl in want) // 9 + 7
        d_rec = torch.empty(9 * cap, dtype=torch.uint8, device=dev)
        d_n = torch.zeros(1, dtype=torch.int64, device=dev)
        st = torch.cuda.current_stream(dev).cuda_stream
        free_after = []
        for k in (0, 1, 0, 1):
            bwt, rl = want[k]
            d_bwt = torch.from_numpy(bwt).to(dev)
            d_rec.fill_(0xEE)
            torch.cuda.synchronize(dev)
            free0 = torch.cuda.mem_get_info(dev)[0]
            n = ctx.rle_encode_static(d_bwt.data_ptr(), len(bwt), d_rec.data_ptr(), cap, d_n.data_ptr(), st)
            torch.cuda.synchronize(dev)
            free_after.append(torch.cuda.mem_get_info(dev)[0] - free0)
            assert n == len(rl) // 9 == int(d_n.item())
            got = d_rec[:9 * n].cpu().numpy()
            assert np.array_equal(got, rl), np.flatnonzero(got != rl)[:10]
            assert bool((d_rec[9 * n:] == 0xEE).all())       # nothing written past the last record
        # once the kernels are loaded (first call), a call changes nothing in the device's free memory
        assert free_after[1:] == [0, 0, 0], free_after

        # an unaligned input, enqueue-only form (no host wait inside the call), count read from the device word
        bwt, _ = want[0]
        d_pad = torch.from_numpy(np.concatenate([np.zeros(3, np.uint8), bwt])).to(dev)
        d_rec.fill_(0)
        torch.cuda.synchronize(dev)
        assert ctx.rle_encode_static(d_pad.data_ptr() + 3, len(bwt), d_rec.data_ptr(), cap, d_n.data_ptr(), st,
                                     wait=False) is None
        torch.cuda.synchronize(dev)
        n = int(d_n.item())
        assert np.array_equal(d_rec[:9 * n].cpu().numpy(), want[0][1])

        # the BWT where the emit stage left it (never on the host), against cmsbwt_emit's own records
        x, text = cases[1]
        ctx.index_upload(x)
        ctx.ms_run(text); ctx.heads_sort(); ctx.emit(P.MODE_PLAIN)
        n = ctx.rle_encode_static(ctx.result_device(), len(text), d_rec.data_ptr(), cap, None, None)
        assert np.array_equal(d_rec[:9 * n].cpu().numpy(), want[1][1])

        # small and degenerate inputs
        for arr in (np.zeros(0, np.uint8), np.array([7], np.uint8), np.full(5000, 65, np.uint8),
                    np.arange(4097, dtype=np.uint32).astype(np.uint8), np.repeat(np.arange(40, dtype=np.uint8), 4096)):
            d_in = torch.from_numpy(arr).to(dev) if len(arr) else torch.zeros(16, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize(dev)
            n = ctx.rle_encode_static(d_in.data_ptr(), len(arr), d_rec.data_ptr(), cap, None, st)
            ref = _rle_numpy(arr)
            assert n == len(ref) // 9
            assert np.array_equal(d_rec[:9 * n].cpu().numpy(), ref)
    finally:
        ctx.close()


def test_static_scratch_rle_limits():
    import torch
    dev = torch.device("cuda:0")
    ctx = P.Context(0)
    try:
        d = torch.zeros(4096, dtype=torch.uint8, device=dev)
        d_rec = torch.zeros(9 * 4, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize(dev)
        with pytest.raises(P.CmsBwtError) as e:               # beyond the static scratch: rejected before any launch
            ctx.rle_encode_static(d.data_ptr(), P.RLE_STATIC_MAX + 1, d_rec.data_ptr(), 4, None, None)
        assert e.value.code == -3 and "CMSBWT_RLE_STATIC_MAX" in str(e.value)
        arr = np.arange(64, dtype=np.uint8)                   # 64 runs into room for 4: the first 4, and an error
        d_in = torch.from_numpy(arr).to(dev)
        torch.cuda.synchronize(dev)
        with pytest.raises(P.CmsBwtError) as e:
            ctx.rle_encode_static(d_in.data_ptr(), 64, d_rec.data_ptr(), 4, None, None)
        assert e.value.code == -3 and "cap_records" in str(e.value)
        assert np.array_equal(d_rec.cpu().numpy(), _rle_numpy(arr)[:36])
    finally:
        ctx.close()
