"""Full-size parity checks on the GPU (BASELINE.json configs 2 and 4-scaled), through properties that
do not need the oracle at that size: the same collection pushed through INDEPENDENT code paths of
the product (2-bit lane kernel + tiles, 8-bit warp kernel + sort-based sweep, the multi-rank path
with its exchanges) must give identical bytes; the BWT's symbol histogram equals the text's; the
run-length records decode to the plain BWT; the window length does not change the result.  (The
small and medium cases of test_gpu_parity.py / test_gpu_sharded.py pin those paths to the oracle
and the reference binary's golden outputs.)"""
import threading

import numpy as np
import pytest
import torch

import cms_bwt_b200 as P
from cms_bwt_b200 import multi
import gen_device

pytestmark = pytest.mark.gpu


def _dev_u8(ptr, n, dev):
    return torch.as_tensor(multi._DevArray(int(ptr), int(n), "|u1"), device=dev)


def _run_single(ctx, text, N, mode=P.MODE_PLAIN, **params):
    for k, v in params.items():
        ctx.set_param(k, v)
    ctx.ms_run_device(text.data_ptr(), N)
    ctx.heads_sort()
    nbytes = ctx.emit(mode)
    out = _dev_u8(ctx.result_device(), nbytes, text.device).clone()
    st = ctx.stats()
    ctx.set_param("fast", 1); ctx.set_param("tile_budget", 1 << 36); ctx.set_param("chunk_len", 0)
    return out, st


def _rle_decode_dev(rec):
    rec = rec.view(-1, 9)
    runs = rec[:, :8].contiguous().view(torch.int64).view(-1)
    return torch.repeat_interleave(rec[:, 8].contiguous(), runs)


def _run_two_ranks(x, dev, ref_bp, copies, div, seed, svs=0, many=False):
    world = 2
    plan = multi.shard_documents([ref_bp] * copies, world)
    shared = multi.ThreadComm.Shared(world)
    res, err = [None] * world, [None] * world

    def work(r):
        ctx = None
        try:
            torch.cuda.set_device(dev)
            a, b = plan[r]
            if many:
                _, text, N, nd = gen_device.make_text_many(ref_bp, copies, div, seed, dev, first_copy=a, ncopies=b - a)
            else:
                _, text, N, nd = gen_device.make_text(ref_bp, copies, div, seed, dev, first_copy=a, ncopies=b - a, svs=svs)
            ctx = P.Context(dev.index)
            ctx.index_upload(x)
            sp, out = multi.run_sharded(ctx, dev, text.data_ptr(), N, nd, r, world, comm=multi.ThreadComm(shared, r))
            res[r] = (sp.out_begin, out.clone())
        except BaseException as e:      # noqa: BLE001
            err[r] = e
            shared.barrier.abort()
        finally:
            if ctx is not None:
                ctx.close()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    assert all(e is None for e in err)
    assert res[1][0] == res[0][1].numel()
    return torch.cat([res[0][1], res[1][1]])


def test_config2_full_size_three_paths_agree():
    """config 2: 5 Mbp reference + 100 copies at 0.1 % (0.5 Gbp), plain and -r"""
    dev = torch.device("cuda:0")
    ref_bp, copies, div, seed = 5_000_000, 100, 0.001, 11
    ref, text, N, nd = gen_device.make_text(ref_bp, copies, div, seed, dev)
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["fast_path"] == 1 and st["light_tiles"] == 1
        assert a.numel() == N
        assert torch.equal(torch.bincount(a.to(torch.int64), minlength=256),
                           torch.bincount(text[:N].to(torch.int64), minlength=256))
        b, st = _run_single(ctx, text, N, chunk_len=3000)
        assert torch.equal(a, b), "window length changed the BWT"
        c, st = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert st["fast_path"] == 0 and st["light_tiles"] == 0
        assert torch.equal(a, c), "general kernel + sweep differ from the lane kernel + tiles"
        rl, st = _run_single(ctx, text, N, mode=P.MODE_RLE)
        assert rl.numel() == 9 * st["n_runs"]
        assert torch.equal(_rle_decode_dev(rl), a), "RLE records do not decode to the plain BWT"
        del b, c, rl
    finally:
        ctx.close()
    d = _run_two_ranks(x, dev, ref_bp, copies, div, seed)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"


def test_config4_scaled_many_short_documents():
    """config 4 at 1/10: 30 kbp reference + 30,000 copies (0.9 Gbp, 30,001 documents); every
    reference position is covered 30,000 deep, so the counting falls back to the sort-based sweep"""
    dev = torch.device("cuda:0")
    ref, text, N, nd = gen_device.make_text(30_000, 30_000, 0.001, 9, dev)
    assert nd == 30_001
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["n_docs"] == nd and st["fast_path"] == 1
        assert torch.equal(torch.bincount(a.to(torch.int64), minlength=256),
                           torch.bincount(text[:N].to(torch.int64), minlength=256))
        assert bool((a[:nd] != 2).sum() <= nd)            # the first nd symbols are the terminators' predecessors
        c, st2 = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert torch.equal(a, c)
        rl, st3 = _run_single(ctx, text, N, mode=P.MODE_RLE)
        assert torch.equal(_rle_decode_dev(rl), a)
        del c, rl
    finally:
        ctx.close()
    del text
    # two ranks: every tile of the 118 holds ~10^5 entries, so the owners count with the sweep
    d = _run_two_ranks(x, dev, 30_000, 30_000, 0.001, 9)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"


def test_config3_full_size_paths_agree():
    """config 3, the configuration the metric is quoted on: 59 Mbp reference + 100 haplotypes at
    0.1 % (5.9 Gbp).  Lane kernel + tiles == warp kernel + sweep == two-rank path, byte for byte."""
    dev = torch.device("cuda:0")
    if torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    ref_bp, copies, div, seed = 59_000_000, 100, 0.001, 13
    ref, text, N, nd = gen_device.make_text(ref_bp, copies, div, seed, dev)
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["fast_path"] == 1 and st["light_tiles"] == 1 and a.numel() == N
        hist_t = torch.zeros(256, dtype=torch.int64, device=dev)
        hist_a = torch.zeros(256, dtype=torch.int64, device=dev)
        for o in range(0, N, 1 << 29):
            hist_t += torch.bincount(text[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
            hist_a += torch.bincount(a[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
        assert torch.equal(hist_t, hist_a)
        c, st = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert st["fast_path"] == 0 and st["light_tiles"] == 0
        assert torch.equal(a, c), "general kernel + sweep differ from the lane kernel + tiles"
        del c
    finally:
        ctx.close()
    del text
    d = _run_two_ranks(x, dev, ref_bp, copies, div, seed)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"


def test_config5_scaled_chr1_reference_one_percent_with_svs():
    """config 5 at 1/16 of the collection: the full 248 Mbp reference (n > 2^27: 28-bit SA ranks, the
    widest keys of the tail sort) + 4 haplotypes at 1 % divergence with 20 block SVs each (deletion,
    tandem duplication, reversal, translocation of 1-50 kbp): ~10 % of the positions are heads.
    Lane kernel + tiles == warp kernel + sweep == two-rank path; histogram; RLE round trip."""
    dev = torch.device("cuda:0")
    if torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    ref_bp, copies, div, seed, svs = 248_000_000, 4, 0.01, 17, 20
    ref, text, N, nd = gen_device.make_text(ref_bp, copies, div, seed, dev, svs=svs)
    x = P.augment(ref.cpu().numpy())
    del ref
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["fast_path"] == 1 and a.numel() == N
        assert st["n_heads"] > 0.05 * N
        hist_t = torch.zeros(256, dtype=torch.int64, device=dev)
        hist_a = torch.zeros(256, dtype=torch.int64, device=dev)
        for o in range(0, N, 1 << 29):
            hist_t += torch.bincount(text[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
            hist_a += torch.bincount(a[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
        assert torch.equal(hist_t, hist_a)
        c, st = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert st["fast_path"] == 0 and st["light_tiles"] == 0
        assert torch.equal(a, c), "general kernel + sweep differ from the lane kernel + tiles"
        del c
        rl, st = _run_single(ctx, text, N, mode=P.MODE_RLE)
        assert torch.equal(_rle_decode_dev(rl), a), "RLE records do not decode to the plain BWT"
        del rl
    finally:
        ctx.close()
    del text
    d = _run_two_ranks(x, dev, ref_bp, copies, div, seed, svs=svs)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"


def test_config4_full_size_300k_documents():
    """config 4 at full size: 30 kbp reference + 300,000 copies at 0.1 % (9.0 Gbp, 300,001 documents,
    N > 2^33).  Lane kernel + sweep == warp kernel + sweep == two-rank path (owners count their deep
    tile ranges with the sweep); histogram; the first 300,001 symbols are the terminators' rows."""
    dev = torch.device("cuda:0")
    if torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    ref_bp, copies, div, seed = 30_000, 300_000, 0.001, 9
    ref, text, N, nd = gen_device.make_text_many(ref_bp, copies, div, seed, dev)
    assert nd == copies + 1 and N > 2 ** 33
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["n_docs"] == nd and st["fast_path"] == 1 and st["light_tiles"] == 0 and a.numel() == N
        hist_t = torch.zeros(256, dtype=torch.int64, device=dev)
        hist_a = torch.zeros(256, dtype=torch.int64, device=dev)
        for o in range(0, N, 1 << 29):
            hist_t += torch.bincount(text[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
            hist_a += torch.bincount(a[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
        assert torch.equal(hist_t, hist_a)
        # rows of the nd terminators come first, in document order: the symbol before each terminator
        ends = torch.nonzero(text[:N] == 2).view(-1)
        assert ends.numel() == nd
        last = torch.where(ends > 0, text[(ends - 1).clamp(min=0)], torch.full_like(text[ends], 2))
        assert torch.equal(a[:nd], last)
        c, st2 = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert st2["fast_path"] == 0
        assert torch.equal(a, c), "general kernel differs from the lane kernel"
        del c
    finally:
        ctx.close()
    del text
    d = _run_two_ranks(x, dev, ref_bp, copies, div, seed, many=True)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"
