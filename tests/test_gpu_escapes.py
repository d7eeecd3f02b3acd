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
