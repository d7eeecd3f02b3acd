"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the
oracle (oracle/cms_oracle.c) on seeded inputs and against the committed golden fixtures made
from the unmodified reference binary.  Integer/byte work: the bar is bit-exact."""
import glob
import os
import numpy as np
import pytest

import cms_bwt_b200 as P
import oracle_lib as O
import fuzz_cases
import gen_synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def ctx():
    c = P.Context(0)
    yield c
    c.close()


def _inputs(ref_file, coll_file, prefix=P.UINT64_MAX):
    x = P.augment(P.read_reference(ref_file))
    text, nd = P.parse_collection(coll_file, prefix)
    return x, text


MODES = {                       # which kernels run (DESIGN.md sections 3-5)
    "fast+tiles": dict(fast=1, tile_budget=1 << 36, tile_margin=48, tile_short=1, tie_local=1),     # 2-bit lane kernel + tile kernels
    "fast+tiles/noshort": dict(fast=1, tile_budget=1 << 36, tile_margin=48, tile_short=0, tie_local=0),   # short ranges walked with the near lists
    "fast+tiles/far": dict(fast=1, tile_budget=1 << 36, tile_margin=0),  # same, nearly every head takes the "far" path
    "fast+tiles/near": dict(fast=1, tile_budget=1 << 36, tile_margin=1 << 30),   # same, every head and range is "near"
    "general+sweep": dict(fast=0, tile_budget=0),        # 8-bit warp kernel + sort-based sweep
    "fast+sweep": dict(fast=1, tile_budget=0),
}


def _check_case(ctx, x, text, want_bwt=None, desc="", chunk=None, check_ms=True, mode="fast+tiles",
                expect_fast=None):
    o = O.Oracle(x)
    for k, v in MODES[mode].items():
        ctx.set_param(k, v)
    desc = "%s [%s]" % (desc, mode)
    if chunk:
        ctx.set_param("chunk_len", chunk)
    ctx.index_upload(x)                       # SA built on the GPU
    ix = ctx.index_download()
    assert np.array_equal(ix["sa"], o.sa), "SA " + desc
    assert np.array_equal(ix["isa"], o.isa), "ISA " + desc
    assert np.array_equal(ix["plcp_prime"].astype(np.int64), o.plcp_prime.astype(np.int64)), "PLCP' " + desc
    assert np.array_equal(ix["ref_bwt"], o.ref_bwt), "refBWT " + desc
    ctx.set_ms_dump(check_ms)
    ctx.ms_run(text)
    if check_ms:
        pos, ln, sm, hd = ctx.ms_download()
        opos, oln, osm, ohd = o.ms(text)
        assert np.array_equal(ln, oln), "MS len " + desc
        assert np.array_equal(pos, opos), "MS pos " + desc
        assert np.array_equal(sm, osm), "MS smaller " + desc
        assert np.array_equal(hd, ohd), "MS heads " + desc
    if expect_fast is not None:
        assert ctx.stats()["fast_path"] == int(expect_fast), "fast path selection " + desc
    ctx.heads_sort()
    if mode.endswith("sweep"):
        assert ctx.stats()["light_tiles"] == 0
    ctx.emit(P.MODE_PLAIN)
    got = ctx.result_download().copy()
    want = o.bwt(text) if want_bwt is None else want_bwt
    if not np.array_equal(got, want):
        bad = np.flatnonzero(got != want) if len(got) == len(want) else []
        raise AssertionError("BWT %s: %d mismatches, first %s" % (desc, len(bad), bad[:8]))
    # RLE mode
    ctx.ms_run(text)
    ctx.heads_sort()
    ctx.emit(P.MODE_RLE)
    rl = ctx.result_download().copy()
    assert np.array_equal(rl, O.rle(want)), "RLBWT " + desc
    ctx.set_param("chunk_len", 0)
    st = ctx.stats()
    for k, v in MODES["fast+tiles"].items():
        ctx.set_param(k, v)
    return st


def test_golden_fixtures(ctx):
    files = sorted(glob.glob(os.path.join(GOLDEN, "*.npz")))
    assert files
    for f in files:
        z = np.load(f)
        x, text = _inputs(z["ref_file"].tobytes(), z["coll_file"].tobytes(), int(z["prefix"]))
        for mode in MODES:
            _check_case(ctx, x, text, z["bwt"], os.path.basename(f), mode=mode)
            # also with tiny chunks (exercises chunk borders inside phrases and documents)
            _check_case(ctx, x, text, z["bwt"], os.path.basename(f) + " chunk=64", chunk=64, mode=mode)


@pytest.mark.parametrize("block", range(4))
def test_fuzz_vs_oracle(ctx, block):
    for seed in range(block * 50, block * 50 + 50):
        c = fuzz_cases.make_case(seed)
        x, text = _inputs(c["ref_file"], c["coll_file"])
        if len(text) == 0:
            continue
        for mode in ("fast+tiles", "general+sweep") + (("fast+tiles/far", "fast+tiles/near", "fast+tiles/noshort") if seed % 5 == 0 else ()):
            _check_case(ctx, x, text, None, c["desc"], chunk=[64, 100, 8192][seed % 3], mode=mode)


def test_config1_scaled_vs_oracle(ctx):
    ref, copies = gen_synth.make_collection(200000, 6, 0.001, 7)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)      # leading empty document
    st = _check_case(ctx, x, text, None, "200k x6", expect_fast=True)
    assert st["n_docs"] == 7 and st["light_tiles"] == 1
    _check_case(ctx, x, text, None, "200k x6", mode="general+sweep", expect_fast=False)


def test_sv_and_dense_divergence(ctx):
    ref, copies = gen_synth.make_collection(100000, 5, 0.01, 5, svs=4)
    x = P.augment(ref)
    text = O.docs_to_text(copies)
    for mode in MODES:
        _check_case(ctx, x, text, None, "100k x5 1% + SVs", mode=mode)


def test_many_short_documents(ctx):
    ref, copies = gen_synth.make_collection(3000, 2000, 0.002, 9)
    x = P.augment(ref)
    text = O.docs_to_text(copies)
    for mode in MODES:
        _check_case(ctx, x, text, None, "3k x2000", check_ms=True, mode=mode)
    # deep enough that a small pair-test budget sends stage 2 to the sort-based sweep
    ctx.set_param("tile_budget", 1000)
    ctx.index_upload(x)
    got = ctx.run(text, P.MODE_PLAIN).copy()
    assert ctx.stats()["light_tiles"] == 0
    ctx.set_param("tile_budget", 1 << 36)
    assert np.array_equal(got, O.Oracle(x).bwt(text))


def test_duplicate_documents_and_identity(ctx):
    rng = np.random.default_rng(1)
    ref = gen_synth.ACGT[rng.integers(0, 4, 5000)]
    x = P.augment(ref)
    docs = [ref.copy() for _ in range(40)] + [ref[100:900].copy() for _ in range(10)]
    text = O.docs_to_text(docs)
    for mode in MODES:
        st = _check_case(ctx, x, text, None, "40 exact copies", mode=mode)
    assert st["rank_rounds"] >= 0


def _repetitive(kind, rng):
    acgt = np.frombuffer(b"ACGT", np.uint8)
    if kind == "units":          # 1000 noisy copies of one 500-bp unit: k-mers with hundreds of suffixes
        unit = acgt[rng.integers(0, 4, 500)]
        ref = np.concatenate([gen_synth.mutate(unit, 0.01, rng) for _ in range(1000)])
        nc = 8
    elif kind == "period7":      # tandem repeat of period 7 with 0.2 % noise
        unit = acgt[rng.integers(0, 4, 7)]
        ref = np.concatenate([gen_synth.mutate(unit, 0.002, rng) for _ in range(60000)])
        nc = 6
    else:                        # homopolymer: one k-mer, one slot, every suffix in it
        ref = np.full(200000, 65, np.uint8)
        nc = 5
    copies = [gen_synth.mutate(ref, 0.005, rng) for _ in range(nc)]
    return ref, copies


@pytest.mark.parametrize("kind", ["units", "period7", "homopolymer"])
def test_repetitive_references_medium(ctx, kind):
    """0.2-0.5 Mbp references that are all repeats (overflowing k-mer slots, exact searches, long tie
    chains in the tail ranking, skewed tiles), 1-4 Mbp of copies, against the oracle: MS per position,
    BWT and RLBWT in all three kernel modes"""
    ref, copies = _repetitive(kind, np.random.default_rng(5))
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    want = O.Oracle(x).bwt(text)
    for mode in MODES:
        st = _check_case(ctx, x, text, want, "repetitive " + kind, mode=mode)
    # a homopolymer reference gets C, G and T in its augmentation: several runs of other symbols, still the 2-bit kernel
    assert st["fast_path"] == 1


def test_host_sa_path_matches_gpu_sa(ctx):
    """index_upload with a host SA (the reference's libsais route) gives the same output."""
    ref, copies = gen_synth.make_collection(50000, 4, 0.002, 3)
    x = P.augment(ref)
    text = O.docs_to_text(copies)
    o = O.Oracle(x)
    ctx.index_upload(x, o.sa)
    got = ctx.run(text, P.MODE_PLAIN).copy()
    assert np.array_equal(got, o.bwt(text))


def test_bad_symbol_is_reported(ctx):
    x = P.augment(np.frombuffer(b"ACGTACGTAGCTAGCTAGGATC", dtype=np.uint8))
    ctx.index_upload(x)
    text = np.frombuffer(b"ACGT\x90CGT\x02", dtype=np.uint8)
    with pytest.raises(P.CmsBwtError) as e:
        ctx.run(text)
    assert e.value.code == -4


def test_round_trip_properties_medium(ctx):
    """Size-independent properties on an input too large for the oracle's O(N) python checks:
    symbol histogram of the BWT equals that of the text; RLE decodes to the plain BWT; output
    is identical for different chunk lengths."""
    ref, copies = gen_synth.make_collection(1000000, 10, 0.001, 42)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    ctx.index_upload(x)
    a = ctx.run(text, P.MODE_PLAIN).copy()
    assert np.array_equal(np.bincount(a, minlength=256), np.bincount(text, minlength=256))
    ctx.set_param("chunk_len", 1000)
    b = ctx.run(text, P.MODE_PLAIN).copy()
    ctx.set_param("chunk_len", 0)
    assert np.array_equal(a, b)
    rl = ctx.run(text, P.MODE_RLE).copy()
    assert np.array_equal(O.rle_decode(rl), a)
    want = O.Oracle(x, ctx.index_download()["sa"]).bwt(text)
    assert np.array_equal(a, want)


def test_streamed_api_matches_run(ctx):
    """cmsbwt_text_begin/append/ms_run_staged + cmsbwt_result_read give the bytes of
    cmsbwt_run, with pieces from cmsbwt_parse_piece and staging slots smaller than the text."""
    for name in ("synth_20k_x8.npz", "fuzz_042.npz", "edge_headerless.npz", "synth_600_x200.npz"):
        z = np.load(os.path.join(GOLDEN, name))
        coll = z["coll_file"].tobytes()
        ctx.index_upload(P.augment(P.read_reference(z["ref_file"].tobytes())))
        for piece, chunk in ((1 << 20, 1 << 20), (4096, 1000), (37, 64)):
            if piece < 64 and len(coll) > 30000:
                continue
            got = ctx.run_streamed(P.parse_pieces(coll, piece), len(coll) + 1, P.MODE_PLAIN, chunk)
            assert np.array_equal(got, z["bwt"]), (name, piece, chunk)
        got = ctx.run_streamed(P.parse_pieces(coll, 5000), len(coll) + 1, P.MODE_RLE, 999)
        assert np.array_equal(got, O.rle(z["bwt"])), name


@pytest.mark.parametrize("stream", [True, False])
def test_cli_binary_matches_golden(workdir, stream):
    """The cms_bwt command line (cms_bwt_b200/host/cms_bwt.cpp; process contract of main.cpp:14-132)
    on the golden fixtures of the unmodified reference binary: .bwt, .rl_bwt (-r), -p, with the
    streamed ingest/egress cut into many small pieces and with --no-stream."""
    import subprocess
    exe = os.path.join(os.path.dirname(GOLDEN), "..", "cms_bwt_b200", "cms_bwt")
    env = dict(os.environ, CMSBWT_PIECE_KB="3", CMSBWT_OUT_KB="2")
    for f in sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))):
        z = np.load(f)
        base = os.path.basename(f)[:-4]
        d = os.path.join(workdir, "cli_%s_%d" % (base, stream))
        os.makedirs(d, exist_ok=True)
        z["ref_file"].tofile(os.path.join(d, "ref.fa"))
        z["coll_file"].tofile(os.path.join(d, "coll.fa"))
        with open(os.path.join(d, "list.txt"), "w") as fh:
            fh.write("%s\n%s  \n" % (os.path.join(d, "ref.fa"), os.path.join(d, "coll.fa")))
        for rle in (False, True):
            for host_sa in ((True, False) if not rle else (True,)):
                cmd = [exe, "-o", os.path.join(d, "out")]
                if int(z["prefix"]) != P.UINT64_MAX:
                    cmd += ["-p", str(int(z["prefix"]))]
                if rle:
                    cmd.append("-r")
                if not stream:
                    cmd.append("--no-stream")
                if not host_sa:
                    cmd.append("--no-host-sa")
                env["CMSBWT_OUT_MMAP"] = "0" if rle else "1"      # pwrite / shared-mapping writers
                r = subprocess.run(cmd + [os.path.join(d, "list.txt")], env=env, stdout=subprocess.PIPE,
                                   stderr=subprocess.STDOUT, timeout=300)
                assert r.returncode == 0, (base, r.stdout.decode()[-600:])
                assert b"==== Time elapsed:" in r.stdout
                got = np.fromfile(os.path.join(d, "out.rl_bwt" if rle else "out.bwt"), np.uint8)
                want = z["rl_bwt"] if rle else z["bwt"]
                assert np.array_equal(got, want), (base, rle, stream, host_sa)
                assert os.path.exists(os.path.join(d, "out.log"))
