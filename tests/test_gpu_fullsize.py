"""Full-size parity on the GPU: every BASELINE.json config pinned to the REFERENCE.

tests/golden/fullsize_pins.json holds, per config, the md5 of the `.bwt` the unmodified reference
binary wrote (or, where the reference cannot run the size, of the independent GSA oracle's output --
SURVEY.md 8c O1/O2), the md5 and count of its maximal-run 9-byte records (O3), all made in the build
container by tests/golden/make_pins.py on inputs from tools/genpin.c.  genpin is deterministic, so
this file regenerates the same collection on the GPU box's host, pushes it through the C ABI and
compares checksums: bit-exact BWT and RLBWT at full size.  The other code paths of the product (8-bit
warp kernel + sort-based sweep, the two-rank path with its exchanges) must then give the same bytes
as the pinned one.  Integer/byte work: the bar is bit-exact."""
import hashlib
import json
import os
import threading

import numpy as np
import pytest
import torch

import cms_bwt_b200 as P
from cms_bwt_b200 import multi
import genpin

pytestmark = pytest.mark.gpu
PINS = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_pins.json")))


def _dev_u8(ptr, n, dev):
    return torch.as_tensor(multi._DevArray(int(ptr), int(n), "|u1"), device=dev)


def _md5_dev(t, chunk=1 << 28):
    m = hashlib.md5()
    for a in range(0, t.numel(), chunk):
        m.update(t[a:a + chunk].cpu().numpy().tobytes())
    return m.hexdigest()


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


def _run_two_ranks(x, dev, g):
    """the same collection as two shards of whole copies, ranks = threads on this GPU"""
    world = 2
    plan = multi.shard_documents([g["ref_bp"]] * g["copies"], world)
    shared = multi.ThreadComm.Shared(world)
    res, err = [None] * world, [None] * world

    def work(r):
        ctx = None
        try:
            torch.cuda.set_device(dev)
            a, b = plan[r]
            _, ht, N, nd = genpin.make_text(g["ref_bp"], g["copies"], g["div"], g["seed"], g["svs"], first=a, count=b - a,
                                            threads=8)
            text = torch.from_numpy(ht[:N + 64]).to(dev)
            torch.cuda.synchronize()
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


def _pinned(name, big=False, two_ranks=True, general=True, expect_light=None):
    if name not in PINS:
        pytest.skip("no pin for %s in tests/golden/fullsize_pins.json" % name)
    dev = torch.device("cuda:0")
    if big and torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    pin = PINS[name]
    g = pin["gen"]
    ref, ht, N, nd = genpin.make_text(g["ref_bp"], g["copies"], g["div"], g["seed"], g["svs"])
    assert N == pin["N"], "generator drifted: N %d != pinned %d" % (N, pin["N"])
    text = torch.from_numpy(ht[:N + 64]).to(dev)
    torch.cuda.synchronize()
    del ht
    x = P.augment(ref)
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["n_docs"] == nd and st["fast_path"] == 1 and a.numel() == N
        if expect_light is not None:
            assert st["light_tiles"] == expect_light
        assert _md5_dev(a) == pin["bwt_md5"], "%s: BWT differs from %s" % (name, pin["source"])
        rl, st2 = _run_single(ctx, text, N, mode=P.MODE_RLE)
        assert st2["n_runs"] == pin["runs"] and rl.numel() == 9 * pin["runs"]
        assert _md5_dev(rl) == pin["rle_md5"], "%s: RLBWT differs from the maximal-run RLE of the reference's BWT" % name
        del rl
        b, _ = _run_single(ctx, text, N, chunk_len=3000 if N < (1 << 31) else 5000)
        assert torch.equal(a, b), "window length changed the BWT"
        del b
        if general:
            c, st3 = _run_single(ctx, text, N, fast=0, tile_budget=0)
            assert st3["fast_path"] == 0 and st3["light_tiles"] == 0
            assert torch.equal(a, c), "general kernel + sweep differ from the pinned path"
            del c
    finally:
        ctx.close()
    del text
    if two_ranks:
        d = _run_two_ranks(x, dev, g)
        assert torch.equal(a, d), "two-rank path differs from the pinned single-GPU path"


def test_config1_pinned_to_the_reference():
    """config 1: 1 Mbp reference + 10 copies at 0.1 %, plain and -r, vs the reference binary"""
    _pinned("config1", expect_light=1)


def test_config2_pinned_to_the_reference():
    """config 2: 5 Mbp reference + 100 copies at 0.1 % (0.5 Gbp), plain and -r, vs the reference binary"""
    _pinned("config2", expect_light=1)


def test_config4_at_1_100_pinned_to_the_reference():
    """config 4 at 1/100: 30 kbp reference + 3,000 copies, vs the reference binary (its small path)"""
    _pinned("config4_100")


def test_config5_at_1_400_pinned_to_the_reference():
    """config 5 shape at 1/400: 5 Mbp + 8 copies at 1 % + 20 block SVs each, vs the reference binary"""
    _pinned("config5_400")


def test_config4_at_1_10_pinned_to_the_gsa_oracle():
    """config 4 at 1/10: 30 kbp reference + 30,000 copies (0.9 Gbp, 30,001 documents); every reference
    position is covered 30,000 deep, so the counting takes the sort-based sweep"""
    _pinned("config4_10", expect_light=0)


def test_config3_pinned_to_the_reference():
    """config 3, the configuration the metric is quoted on: 59 Mbp reference + 100 haplotypes at 0.1 %
    (5.9 Gbp), vs the md5 of the unmodified reference binary's .bwt on the same files"""
    _pinned("config3", big=True, expect_light=1)


def test_config5_at_1_16_pinned_to_the_gsa_oracle():
    """config 5 at 1/16 of the collection: the full 248 Mbp reference (n > 2^27: 28-bit SA ranks) + 4
    haplotypes at 1 % with 20 block SVs each; ~10 % of the positions are heads"""
    _pinned("config5_16", big=True)


def test_config4_full_size_300k_documents():
    """config 4 at full size: 30 kbp reference + 300,000 copies at 0.1 % (9.0 Gbp, 300,001 documents,
    N > 2^33).  Pinned to the reference binary when tests/golden/fullsize_pins.json has the entry;
    in any case lane kernel + sweep == warp kernel + sweep == two-rank path, histogram, and the first
    300,001 symbols are the terminators' rows."""
    dev = torch.device("cuda:0")
    if torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    g = dict(ref_bp=30_000, copies=300_000, div=0.001, seed=9, svs=0)
    ref, ht, N, nd = genpin.make_text(g["ref_bp"], g["copies"], g["div"], g["seed"], g["svs"])
    assert nd == g["copies"] + 1 and N > 2 ** 33
    text = torch.from_numpy(ht[:N + 64]).to(dev)
    torch.cuda.synchronize()
    del ht
    x = P.augment(ref)
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        a, st = _run_single(ctx, text, N)
        assert st["n_docs"] == nd and st["fast_path"] == 1 and st["light_tiles"] == 0 and a.numel() == N
        if "config4" in PINS:
            assert N == PINS["config4"]["N"] and _md5_dev(a) == PINS["config4"]["bwt_md5"]
        hist_t = torch.zeros(256, dtype=torch.int64, device=dev)
        hist_a = torch.zeros(256, dtype=torch.int64, device=dev)
        for o in range(0, N, 1 << 29):
            hist_t += torch.bincount(text[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
            hist_a += torch.bincount(a[o:min(N, o + (1 << 29))].to(torch.int32), minlength=256)
        assert torch.equal(hist_t, hist_a)
        ends = torch.nonzero(text[:N] == 2).view(-1)
        assert ends.numel() == nd
        last = torch.where(ends > 0, text[(ends - 1).clamp(min=0)], torch.full_like(text[ends], 2))
        assert torch.equal(a[:nd], last)
        del ends, last
        c, st2 = _run_single(ctx, text, N, fast=0, tile_budget=0)
        assert st2["fast_path"] == 0
        assert torch.equal(a, c), "general kernel differs from the lane kernel"
        del c
    finally:
        ctx.close()
    del text
    d = _run_two_ranks(x, dev, g)
    assert torch.equal(a, d), "two-rank path differs from the single-GPU path"
