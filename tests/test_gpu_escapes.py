"""Symbols outside A,C,G,T in the collection stay on the 2-bit matching-statistics kernel (SURVEY.md 8f-3:
the reference takes any byte 3..127 through its augmentation, CMS-BWT-functions.cpp:231-237; round 1
sent the whole job to the 8-bit kernel for a single `N`).  Per-position MS, BWT and RLBWT against the
oracle, with `fast_path == 1` asserted; the golden fixtures of the reference binary whose reference is
pure ACGT and whose collection carries other symbols must run on the fast path too."""
import glob
import os

import numpy as np
import pytest

import cms_bwt_b200 as P
import oracle_lib as O
import gen_synth
from test_gpu_parity import _check_case, _inputs, GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = P.Context(0)
    yield c
    c.close()


def _sprinkle(text, rng, rate, symbols=b"NnRYacgt\x03\x7f@B"):
    t = text.copy()
    idx = np.flatnonzero(t != 2)
    pick = idx[rng.random(len(idx)) < rate]
    t[pick] = np.frombuffer(symbols, np.uint8)[rng.integers(0, len(symbols), len(pick))]
    return t, len(pick)


@pytest.mark.parametrize("rate", [0.0002, 0.005])
def test_sprinkled_escapes_stay_on_the_fast_path(ctx, rate):
    rng = np.random.default_rng(23)
    ref, copies = gen_synth.make_collection(200000, 6, 0.002, 12)
    x = P.augment(ref)
    text, nesc = _sprinkle(O.docs_to_text([np.zeros(0, np.uint8)] + copies), rng, rate)
    for chunk in (None, 100):
        st = _check_case(ctx, x, text, None, "escapes %g" % rate, chunk=chunk, expect_fast=True)
        assert st["n_escapes"] == nesc and st["light_tiles"] == 1


def test_escape_runs_and_matches_through_the_augmentation(ctx):
    ref, copies = gen_synth.make_collection(30000, 5, 0.003, 4)
    x = P.augment(ref)
    t = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    ends = np.flatnonzero(t == 2)
    for e in ends[1:4]:
        t[e - 1] = ord("N"); t[e + 1] = ord("n")
    t[100:140] = ord("N")
    aug = x[len(ref):len(x) - 2]
    t[2000:2030] = aug[:30]
    t[3000:3005] = aug[-5:]
    t[4999] = ref[-1]; t[5000:5010] = aug[:10]
    k = len(ref) - 40
    t[7000:7040] = ref[k:k + 40]; t[7040:7045] = aug[:5]
    for chunk in (64, 1000, None):
        _check_case(ctx, x, t, None, "escape runs", chunk=chunk, expect_fast=True)
    docs = [np.frombuffer(b"NNNN", np.uint8), np.frombuffer(b"N", np.uint8), ref[50:90].copy(),
            np.frombuffer(b"ACNGT", np.uint8)]
    _check_case(ctx, x, O.docs_to_text(docs), None, "escape documents", chunk=64, expect_fast=True)


def test_dense_escapes_fall_back_to_the_general_kernel(ctx):
    """a lower-case collection (every symbol an escape) would make the break list as long as the text:
    the 8-bit kernel takes it, same bytes"""
    ref, copies = gen_synth.make_collection(20000, 4, 0.003, 6)
    x = P.augment(ref)
    text = O.docs_to_text(copies)
    low = np.where(text == 2, 2, text | 0x20).astype(np.uint8)
    _check_case(ctx, x, low, None, "lower case", expect_fast=False)


def test_golden_fixtures_with_other_symbols_run_fast(ctx):
    ran = 0
    for f in sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))):
        z = np.load(f)
        x, text = _inputs(z["ref_file"].tobytes(), z["coll_file"].tobytes(), int(z["prefix"]))
        body = x[:-2]
        acgt = np.isin(body, np.frombuffer(b"ACGT", np.uint8))
        m = int(np.argmin(acgt)) if not acgt.all() else len(body)
        pure = m > 0 and not acgt[m:].any() and body[m] < ord("A") if m < len(body) else False
        other = ~np.isin(text, np.frombuffer(b"ACGT\x02", np.uint8))
        if not pure or not other.any() or other.sum() > len(text) // 32 + 4096:
            continue
        st = _check_case(ctx, x, text, z["bwt"], os.path.basename(f), expect_fast=True)
        assert st["n_escapes"] == int(other.sum())
        ran += 1
    print("golden fixtures with escapes on the fast path:", ran)


def test_reference_with_other_symbols_stays_on_the_fast_path(ctx):
    """N blocks, IUPAC codes, lower case and symbols below 'A' / above 'T' in the REFERENCE (second
    half): run list beside the 2-bit reference (ms_fast.cuh: msf_room_at), matches end at the runs or run
    through them where the text carries the same symbols"""
    rng = np.random.default_rng(29)
    ref, copies = gen_synth.make_collection(200000, 5, 0.002, 14)
    ref = ref.copy()
    ref[0:7] = ord("N")
    ref[50000:53000] = ord("N")
    for q in (90000, 90001, 90040, 120000, 120017, 120018, 120019):
        ref[q] = ord("N")
    ref[150000] = ord("R"); ref[150003] = ord("y"); ref[150100] = ord("#"); ref[150200] = ord("B")
    ref[150300] = ord("a"); ref[150400] = ord("|")
    ref[-3:] = ord("N")
    x = P.augment(ref)
    base = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    for chunk in (None, 100):
        _check_case(ctx, x, base, None, "reference runs", chunk=chunk, expect_fast=True)
    t = base.copy()
    t[51000:51500] = ord("N"); t[90000] = ord("N"); t[120017:120020] = ord("N"); t[150000] = ord("R")
    t, _ = _sprinkle(t, rng, 0.001)
    _check_case(ctx, x, t, None, "reference runs + text escapes", expect_fast=True)
    docs = [ref[49990:53010].copy(), ref[89990:90050].copy(), ref[149990:150410].copy(), ref[-40:].copy(), ref[:30].copy()]
    _check_case(ctx, x, O.docs_to_text(docs), None, "documents with the reference's runs", chunk=64, expect_fast=True)


def test_golden_fixtures_with_impure_references_run_fast(ctx):
    """the fixtures of the reference binary whose REFERENCE carries N / lower case (VERDICT r1 item 10 names
    fuzz_003, fuzz_077, fuzz_150, fuzz_199): identical bytes with fast_path == 1"""
    ran = []
    for f in sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))):
        z = np.load(f)
        x, text = _inputs(z["ref_file"].tobytes(), z["coll_file"].tobytes(), int(z["prefix"]))
        acgt = np.isin(x, np.frombuffer(b"ACGT", np.uint8))
        na = ~acgt
        runs = int(np.sum(na[1:] & ~na[:-1])) + int(na[0])
        other = ~np.isin(text, np.frombuffer(b"ACGT\x02", np.uint8))
        if runs < 2 or runs > len(x) // 8 + 16 or not acgt.any() or other.sum() > len(text) // 32 + 4096 or len(text) == 0:
            continue
        _check_case(ctx, x, text, z["bwt"], os.path.basename(f), expect_fast=True)
        ran.append(os.path.basename(f))
    print("golden fixtures with impure references on the fast path:", ran)
    assert len(ran) >= 4
