"""Pins the oracle (oracle/cms_oracle.c): against the unmodified reference binary and the
independent GSA oracle where oracle/_ref has been built (this container), and against the
committed golden fixtures everywhere."""
import glob
import os
import numpy as np
import pytest

import oracle_lib as O

HERE = os.path.dirname(os.path.abspath(__file__))
import fuzz_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def oracle_bwt_from_files(ref_file, coll_file, prefix=2 ** 64 - 1):
    x = O.augment(O.read_reference(ref_file))
    text, ds = O.parse_collection(coll_file, prefix)
    o = O.Oracle(x)
    return o.bwt(text), text, x


def _write(workdir, name, case):
    rp = os.path.join(workdir, name + ".ref")
    cp = os.path.join(workdir, name + ".coll")
    open(rp, "wb").write(case["ref_file"])
    open(cp, "wb").write(case["coll_file"])
    return rp, cp


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("block", range(8))
def test_oracle_matches_reference_binary_fuzz(block, workdir):
    for seed in range(block * 25, block * 25 + 25):
        case = fuzz_cases.make_case(seed)
        rp, cp = _write(workdir, "fz%d" % seed, case)
        text, ds = O.parse_collection(case["coll_file"])
        if len(text) == 0:
            continue
        want = O.run_reference(rp, cp, workdir, name="fz%d" % seed)
        got, _, _ = oracle_bwt_from_files(case["ref_file"], case["coll_file"])
        assert np.array_equal(got, want), case["desc"]
        if O.have_gsabwt():
            gsa = O.run_gsabwt(cp, workdir, name="fz%d" % seed)
            assert np.array_equal(gsa, want), case["desc"]


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
def test_oracle_matches_reference_binary_config1_scaled(workdir):
    import gen_synth
    ref, copies = gen_synth.make_collection(200000, 6, 0.001, 7)
    pre = os.path.join(workdir, "c1s")
    gen_synth.write_files(pre, ref, copies)
    want = O.run_reference(pre + ".ref.fa", pre + ".coll.fa", workdir, name="c1s")
    got, _, _ = oracle_bwt_from_files(open(pre + ".ref.fa", "rb").read(),
                                      open(pre + ".coll.fa", "rb").read())
    assert np.array_equal(got, want)


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("P", [5000, 2007, 2006, 2008, 2001, 2002, 2003, 100, 4003, 9999, 10005, 10006, 20000])
def test_oracle_prefix_matches_reference(P, workdir):
    """-p semantics (:1259-1262): P symbols, or P-1 when the cut falls on a boundary."""
    rng = np.random.default_rng(3)
    ref = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 3000)]
    docs = [ref[:2000].copy() for _ in range(5)]
    for d in docs:
        d[rng.integers(0, 2000, 5)] = ord("A")
    case = dict(ref_file=b">r\n" + bytes(ref) + b"\n",
                coll_file=b"".join(b">s\n" + bytes(d) + b"\n" for d in docs))
    rp, cp = _write(workdir, "pfx%d" % P, case)
    want = O.run_reference(rp, cp, workdir, prefix=P, name="pfx%d" % P)
    got, _, _ = oracle_bwt_from_files(case["ref_file"], case["coll_file"], prefix=P)
    assert np.array_equal(got, want)


def test_oracle_ms_matches_bruteforce_definition():
    """O4: MS length = longest prefix of the suffix occurring in the augmented reference; pos is
    the lexicographic neighbour chosen by the rule of :53-76."""
    for seed in (1, 2, 3, 4, 5, 6):
        case = fuzz_cases.make_case(1000 + seed)
        x = O.augment(O.read_reference(case["ref_file"]))
        text, ds = O.parse_collection(case["coll_file"])
        if len(text) == 0 or len(text) > 3000:
            continue
        o = O.Oracle(x)
        pos, ln, sm, hd = o.ms(text)
        xb = bytes(x)
        sa = o.sa
        sufs = [xb[s:] for s in sa]
        tb = bytes(text)
        for q in range(len(text)):
            if text[q] == 2:
                assert ln[q] == 0
                continue
            e = tb.index(b"\x02", q)
            T = tb[q:e + 1]
            L = 0
            while L < len(T) and xb.find(T[:L + 1]) >= 0:
                L += 1
            assert ln[q] == L, (case["desc"], q)
            # neighbour rule: succ if lcp(T,succ) >= lcp(T,pred) else pred
            import bisect
            k = bisect.bisect_left(sufs, T)
            def lcp(a, b):
                m = 0
                while m < len(a) and m < len(b) and a[m] == b[m]:
                    m += 1
                return m
            ls = lcp(T, sufs[k]) if k < len(sufs) else -1
            lp = lcp(T, sufs[k - 1]) if k > 0 else -1
            if ls >= lp:
                assert pos[q] == sa[k] and sm[q] == 1, (case["desc"], q)
            else:
                assert pos[q] == sa[k - 1] and sm[q] == 0, (case["desc"], q)


def test_oracle_rle_roundtrip():
    rng = np.random.default_rng(0)
    b = np.repeat(rng.integers(65, 70, 200).astype(np.uint8), rng.integers(1, 9, 200))
    rec = O.rle(b)
    assert len(rec) % 9 == 0
    assert np.array_equal(O.rle_decode(rec), b)
    r = rec.reshape(-1, 9)
    assert np.all(r[1:, 8] != r[:-1, 8])          # maximal runs


def test_oracle_matches_golden():
    files = sorted(glob.glob(os.path.join(GOLDEN, "*.npz")))
    assert files, "no golden fixtures committed"
    for f in files:
        z = np.load(f)
        got, text, x = oracle_bwt_from_files(z["ref_file"].tobytes(), z["coll_file"].tobytes(),
                                             int(z["prefix"]))
        assert np.array_equal(got, z["bwt"]), f
        assert np.array_equal(O.rle(got), z["rl_bwt"]), f


# ---- full-size pins (tests/golden/fullsize_pins.json, made by tests/golden/make_pins.py) -------------
def test_genpin_text_is_the_parse_of_its_fasta(tmp_path):
    """tools/genpin.c: the text it hands to the GPU tests equals the reference's document list of the
    FASTA files it writes for the reference binary; shards of copies concatenate to the whole."""
    import genpin
    import cms_bwt_b200 as P
    pre = str(tmp_path / "gp")
    genpin.write_files(20000, 5, 0.002, 3, 2, pre)
    ref, t, N, nd = genpin.make_text(20000, 5, 0.002, 3, 2)
    text, nd2 = P.parse_collection(open(pre + ".coll.fa", "rb").read())
    assert nd == nd2 == 6 and np.array_equal(t[:N], text)
    assert np.array_equal(ref, P.read_reference(open(pre + ".ref.fa", "rb").read()))
    _, t0, N0, d0 = genpin.make_text(20000, 5, 0.002, 3, 2, first=0, count=2)
    _, t1, N1, d1 = genpin.make_text(20000, 5, 0.002, 3, 2, first=2, count=3)
    assert np.array_equal(np.concatenate([t0[:N0], t1[:N1]]), t[:N]) and d0 + d1 == nd


def test_pins_cover_every_baseline_config():
    import json
    pins = json.load(open(os.path.join(HERE, "golden", "fullsize_pins.json")))
    for name in ("config1", "config2", "config3", "config4_100", "config4_10", "config5_400", "config5_16"):
        assert name in pins, name
        p = pins[name]
        assert len(p["bwt_md5"]) == 32 and len(p["rle_md5"]) == 32 and p["runs"] > 0 and p["N"] > 0


def test_oracle_restatement_matches_the_config1_pin():
    """the C restatement (oracle/cms_oracle.c) on the genpin config-1 collection gives the bytes whose
    md5 the unmodified reference binary produced (pin) -- ties the oracle to the full-size pins"""
    import hashlib
    import json
    import genpin
    pins = json.load(open(os.path.join(HERE, "golden", "fullsize_pins.json")))
    p = pins["config1"]
    g = p["gen"]
    ref, t, N, nd = genpin.make_text(g["ref_bp"], g["copies"], g["div"], g["seed"], g["svs"])
    assert N == p["N"]
    x = O.augment(ref)
    bwt = O.Oracle(x).bwt(t[:N].copy())
    assert hashlib.md5(bwt.tobytes()).hexdigest() == p["bwt_md5"]
    rl = O.rle(bwt)
    assert len(rl) == 9 * p["runs"] and hashlib.md5(rl.tobytes()).hexdigest() == p["rle_md5"]
