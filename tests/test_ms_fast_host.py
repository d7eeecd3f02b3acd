"""CPU check of the lane state machine of the fast matching-statistics kernel.

cms_bwt_b200/csrc/ms_fast.cuh is plain C++ when compiled without nvcc; tests/harness/msf_host.cpp
runs it lane by lane (round-robin, like the threads of a warp) and this test compares every
position's (pos, len, smaller, head) and the compacted head list with the oracle
(oracle/cms_oracle.c) on the golden fixtures, the fuzz cases and synthetic collections, for several
window lengths (so window borders fall inside phrases, hard zones and documents).
The product never runs this host build: it is test infrastructure for "host logic".
"""
import ctypes as C
import glob
import os
import subprocess
import numpy as np
import pytest

import oracle_lib as O
import fuzz_cases
import gen_synth

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "harness", "msf_host.cpp")
HDR = os.path.join(ROOT, "cms_bwt_b200", "csrc", "ms_fast.cuh")
WORK = os.path.join(HERE, "_work")
SO = os.path.join(WORK, "libmsf_host.so")


@pytest.fixture(scope="module")
def harness():
    os.makedirs(WORK, exist_ok=True)
    if (not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(HDR))):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", SRC, "-o", SO])
    L = C.CDLL(SO)
    L.msf_host_run.restype = C.c_longlong
    L.msf_host_run.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                               C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_int] + [C.c_void_p] * 8
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pick_k(n):
    k = 1
    while k < 12 and (1 << (2 * (k + 1))) <= n:
        k += 1
    return k


def run_case(L, x, text, WL, k=None, skip_min=20, nlanes=7, desc="", slots=None):
    if slots is None:                      # both search paths: k-mer slot table, and ktab + sak only
        r = run_case(L, x, text, WL, k, skip_min, nlanes, desc + " [slots]", slots=1)
        if r is not None:                  # + neighbours of empty buckets read from the slot
            run_case(L, x, text, WL, k, skip_min, nlanes, desc + " [slots+nb]", slots=2)
        return r if r is None else run_case(L, x, text, WL, k, skip_min, nlanes, desc + " [ktab]", slots=0)
    o = O.Oracle(x)
    sa = o.sa
    pp = o.plcp_prime.astype(np.int32)
    N = len(text)
    k = k or pick_k(len(x))
    dpos = np.zeros(N, np.uint32); dlen = np.zeros(N, np.uint32)
    dsm = np.zeros(N, np.uint8); dhd = np.zeros(N, np.uint8)
    hg = np.zeros(N, np.uint64); hpos = np.zeros(N, np.uint32); hlen = np.zeros(N, np.uint32)
    hcs = np.zeros(N, np.uint16)
    xc = np.ascontiguousarray(x); tc = np.ascontiguousarray(text)
    H = L.msf_host_run(_p(xc), len(xc), _p(sa), _p(pp), _p(tc), N, WL, k, skip_min, nlanes, slots,
                       _p(dpos), _p(dlen), _p(dsm), _p(dhd), _p(hg), _p(hpos), _p(hlen), _p(hcs))
    if H == -1:
        return None
    assert H >= 0, "pool overflow " + desc
    opos, olen, osm, ohd = o.ms(text)
    tag = "%s WL=%d k=%d" % (desc, WL, k)
    assert np.array_equal(dlen, olen), "MS len " + tag + " first bad %s" % np.flatnonzero(dlen != olen)[:5]
    assert np.array_equal(dpos, opos), "MS pos " + tag + " first bad %s" % np.flatnonzero(dpos != opos)[:5]
    assert np.array_equal(dsm, osm), "MS smaller " + tag
    assert np.array_equal(dhd, ohd), "MS heads " + tag + " first bad %s" % np.flatnonzero(dhd != ohd)[:5]
    idx = np.flatnonzero(ohd)
    assert H == len(idx), "head count " + tag
    assert np.array_equal(hg[:H], idx.astype(np.uint64)), "head positions " + tag
    assert np.array_equal(hpos[:H], opos[idx]) and np.array_equal(hlen[:H], olen[idx]), "head records " + tag
    prev = np.where((idx == 0) | (text[np.maximum(idx, 1) - 1] == 2), 2, text[np.maximum(idx, 1) - 1])
    assert np.array_equal(hcs[:H] & 0xFF, prev.astype(np.uint16)), "head symbols " + tag
    assert np.array_equal(hcs[:H] >> 8, osm[idx].astype(np.uint16)), "head smaller flags " + tag
    return H


def test_golden_and_fuzz(harness):
    ran = 0
    for f in sorted(glob.glob(os.path.join(HERE, "golden", "*.npz"))):
        z = np.load(f)
        x = O.augment(O.read_reference(z["ref_file"].tobytes()))
        text, _ = O.parse_collection(z["coll_file"].tobytes(), int(z["prefix"]))
        for WL in (64, 100, 4096):
            if run_case(harness, x, text, WL, desc=os.path.basename(f)) is not None:
                ran += 1
    for seed in range(200):
        c = fuzz_cases.make_case(seed)
        x = O.augment(O.read_reference(c["ref_file"]))
        text, _ = O.parse_collection(c["coll_file"])
        if len(text) == 0:
            continue
        if run_case(harness, x, text, [64, 100, 4096][seed % 3], desc=c["desc"]) is not None:
            ran += 1
    assert ran >= 180, "too few cases exercised the fast path (%d)" % ran      # references with runs of other symbols included


@pytest.mark.parametrize("WL", [37, 256, 4096])
def test_synthetic(harness, WL):
    ref, copies = gen_synth.make_collection(60000, 6, 0.002, 3)
    x = O.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    assert run_case(harness, x, text, WL, desc="60k x6") is not None
    ref, copies = gen_synth.make_collection(20000, 5, 0.01, 5, svs=4)
    assert run_case(harness, O.augment(ref), O.docs_to_text(copies), WL, desc="20k x5 sv") is not None


def test_small_k_and_repeats(harness):
    rng = np.random.default_rng(2)
    unit = gen_synth.ACGT[rng.integers(0, 4, 23)]
    ref = np.concatenate([np.tile(unit, 40), gen_synth.ACGT[rng.integers(0, 4, 500)], np.tile(unit, 25)])
    docs = [ref[100:1200].copy(), ref[5:900].copy(), np.tile(unit, 60), ref.copy()]
    docs[1][300] = ord("A") if docs[1][300] != ord("A") else ord("C")
    x = O.augment(ref)
    text = O.docs_to_text(docs)
    for k in (2, 5, 8):
        for WL in (50, 1000):
            assert run_case(harness, x, text, WL, k=k, skip_min=2, desc="tandem") is not None
    # many short documents (separators closer than k + 16)
    ref, copies = gen_synth.make_collection(3000, 300, 0.003, 9)
    short = [c[a:a + ln] for c in copies for a, ln in ((5, 11), (40, 30), (100, 3))]
    assert run_case(harness, O.augment(ref), O.docs_to_text(short), 64, desc="short docs") is not None


def _sprinkle(text, rng, rate, symbols=b"NnRYacgt\x03\x7f@B"):
    """replaces symbols of the documents (never a separator) by symbols outside A,C,G,T"""
    t = text.copy()
    idx = np.flatnonzero(t != 2)
    pick = idx[rng.random(len(idx)) < rate]
    t[pick] = np.frombuffer(symbols, np.uint8)[rng.integers(0, len(symbols), len(pick))]
    return t


@pytest.mark.parametrize("WL", [64, 1000])
def test_text_escapes(harness, WL):
    """symbols outside A,C,G,T in the TEXT (N, IUPAC, lower case, the first and last augmentation symbols)
    stay on the 2-bit path through the break list: per-position MS against the oracle"""
    rng = np.random.default_rng(17)
    ref, copies = gen_synth.make_collection(30000, 5, 0.003, 4)
    x = O.augment(ref)
    base = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    for rate in (0.0005, 0.01, 0.2):
        assert run_case(harness, x, _sprinkle(base, rng, rate), WL, desc="escapes %g" % rate) is not None
    # runs of escapes, escapes next to separators and at document starts, a run that copies the augmentation
    t = base.copy()
    ends = np.flatnonzero(t == 2)
    for e in ends[1:4]:
        t[e - 1] = ord("N"); t[e + 1] = ord("n")
    t[100:140] = ord("N")
    aug = x[len(ref):len(x) - 2]
    t[2000:2000 + 30] = aug[:30]                      # matches the augmentation itself for 30 symbols
    t[3000:3005] = aug[-5:]
    t[5000 - 1] = ref[-1]; t[5000:5000 + 10] = aug[:10]  # reference end + augmentation: a match through an escape
    k = len(ref) - 40
    t[7000:7040] = ref[k:k + 40]; t[7040:7045] = aug[:5]
    assert run_case(harness, x, t, WL, desc="escape runs") is not None
    # short documents made of escapes only
    docs = [np.frombuffer(b"NNNN", np.uint8), np.frombuffer(b"N", np.uint8), ref[50:90].copy(), np.frombuffer(b"ACNGT", np.uint8)]
    assert run_case(harness, x, O.docs_to_text(docs), 64, desc="escape documents") is not None


@pytest.mark.parametrize("WL", [64, 1000])
def test_reference_runs(harness, WL):
    """symbols outside A,C,G,T in the REFERENCE (N blocks, IUPAC codes, lower case, symbols that sort before 'A',
    between the bases and after 'T') stay on the 2-bit path through the run list (ms_fast.cuh: msf_room_at):
    per-position MS against the oracle, text with and without the same symbols"""
    rng = np.random.default_rng(23)
    ref, copies = gen_synth.make_collection(30000, 5, 0.003, 6)
    ref = ref.copy()
    ref[0:7] = ord("N")                                   # the reference starts inside a run
    ref[5000:5400] = ord("N")                             # a long block
    for q in (9000, 9001, 9040, 12000, 12017, 12018, 12019, 20000):
        ref[q] = ord("N")                                 # single symbols closer than k + 16 to each other
    ref[15000] = ord("R"); ref[15003] = ord("y"); ref[15100] = ord("#"); ref[15200] = ord("B")
    ref[15300] = ord("D"); ref[15400] = ord("H"); ref[15500] = ord("a"); ref[15600] = ord("|")
    ref[-3:] = ord("N")                                   # a run right in front of the augmentation
    x = O.augment(ref)
    base = O.docs_to_text([np.zeros(0, np.uint8)] + copies)   # copies of the clean reference: matches end at the runs
    assert run_case(harness, x, base, WL, desc="reference runs") is not None
    t = base.copy()
    t[5100:5200] = ord("N")                               # the text has the block too: matches run through it
    t[9000] = ord("N"); t[12017:12020] = ord("N"); t[15000] = ord("R"); t[15100] = ord("#")
    assert run_case(harness, x, t, WL, desc="reference runs + text escapes") is not None
    assert run_case(harness, x, _sprinkle(base, rng, 0.01), WL, desc="reference runs + sprinkled escapes") is not None
    # documents cut from the impure reference itself (exact copies through every run)
    docs = [ref[4990:5410].copy(), ref[8990:9050].copy(), ref[14990:15610].copy(), ref[-40:].copy(), ref[:30].copy()]
    assert run_case(harness, x, O.docs_to_text(docs), 64, desc="documents with the reference's runs") is not None
    # tiny k (runs closer than k) and a reference that is mostly runs
    ref2 = gen_synth.ACGT[rng.integers(0, 4, 3000)].copy()
    ref2[rng.random(3000) < 0.01] = ord("N")
    docs2 = [ref2[a:a + 200].copy() for a in range(0, 2800, 150)]
    for k in (2, 5):
        assert run_case(harness, O.augment(ref2), O.docs_to_text(docs2), 64, k=k, skip_min=2, desc="dense runs") is not None


def test_reference_runs_fuzz(harness):
    """random references over A,C,G,T with symbols from the whole legal range 3..127 sprinkled in at several densities
    (single symbols and blocks), documents cut from the reference (so matches run through the runs) and mutated
    copies; small k so that runs fall inside k-mers and inside the 16 key bases; per-position MS against the oracle"""
    ran = 0
    for seed in range(48):
        rng = np.random.default_rng(1000 + seed)
        n = int(rng.integers(300, 2500))
        ref = gen_synth.ACGT[rng.integers(0, 4, n)].copy()
        pool = np.array([c for c in range(3, 128) if chr(c) not in "ACGT"], np.uint8)
        if seed % 3 == 0:
            pool = np.frombuffer(b"N", np.uint8)
        rate = [0.002, 0.01, 0.04][seed % 3]
        hit = rng.random(n) < rate
        ref[hit] = pool[rng.integers(0, len(pool), int(hit.sum()))]
        for _ in range(int(rng.integers(0, 3))):                       # blocks
            a = int(rng.integers(0, n - 40)); ref[a:a + int(rng.integers(2, 40))] = pool[int(rng.integers(0, len(pool)))]
        if not np.isin(ref, gen_synth.ACGT).any():
            continue
        docs = []
        for _ in range(6):
            a = int(rng.integers(0, n - 30)); b = min(n, a + int(rng.integers(20, 400)))
            d = ref[a:b].copy()
            if rng.random() < 0.6:
                m = rng.random(len(d)) < 0.02
                d[m] = gen_synth.ACGT[rng.integers(0, 4, int(m.sum()))]
            docs.append(d)
        x = O.augment(ref)
        text = O.docs_to_text(docs)
        k = [2, 3, 4, 5][seed % 4]
        r = run_case(harness, x, text, [37, 64, 500][seed % 3], k=k, skip_min=[2, 8, 20][seed % 3], desc="runs fuzz %d" % seed)
        if r is not None:
            ran += 1
    assert ran >= 30, ran


def test_long_blocks_in_reference_and_text(harness):
    """N blocks of tens of thousands of symbols that the reference AND the copies carry (haplotypes built from a
    reference keep its gaps): a match crosses a block byte-wise once and its non-head run covers the block,
    also when a document or a window starts inside one -- never one exact search per symbol of the block."""
    import time
    rng = np.random.default_rng(41)
    ref = gen_synth.ACGT[rng.integers(0, 4, 90000)].copy()
    ref[0:12000] = ord("N")                       # the reference starts with a block (a telomere gap)
    ref[40000:60000] = ord("N")
    ref[70000:70300] = ord("n")
    copies = []
    for c in range(3):
        d = ref.copy()
        snp = np.flatnonzero(np.isin(d, gen_synth.ACGT) & (rng.random(len(d)) < 0.002))
        d[snp] = gen_synth.ACGT[rng.integers(0, 4, len(snp))]
        copies.append(d)
    copies.append(ref[5000:45000].copy())         # starts and ends inside blocks
    copies.append(ref[41000:41050].copy())        # nothing but a piece of a block
    shorter = ref.copy(); shorter = np.concatenate([shorter[:45000], shorter[52000:]])   # a block 7000 symbols shorter
    copies.append(shorter)
    x = O.augment(ref)
    text = O.docs_to_text(copies)
    t0 = time.time()
    for WL in (1000, 8192):
        assert run_case(harness, x, text, WL, desc="long blocks") is not None
    assert time.time() - t0 < 120, "the block symbols are being searched one by one"
