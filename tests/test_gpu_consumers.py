"""Consumer of the path's output fed from device memory: the FM-index built over the BWT
the emit stage left in HBM (cmsbwt_fm_build) and its batched backward search (cmsbwt_fm_count), against
brute-force occurrence counts in the collection."""
import numpy as np
import pytest

import cms_bwt_b200 as P
import oracle_lib as O
import gen_synth

pytestmark = pytest.mark.gpu


def _count(hay, needle):
    n, k = 0, hay.find(needle)
    while k >= 0:
        n += 1
        k = hay.find(needle, k + 1)
    return n


@pytest.mark.parametrize("with_escapes", [False, True])
def test_fm_index_counts_match_brute_force(with_escapes):
    rng = np.random.default_rng(31)
    ref, copies = gen_synth.make_collection(40000, 7, 0.004, 8)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    if with_escapes:
        idx = np.flatnonzero(text != 2)
        text[idx[rng.random(len(idx)) < 0.002]] = ord("N")
    hay = text.tobytes()
    ctx = P.Context(0)
    try:
        ctx.index_upload(x)
        bwt = ctx.run(text, P.MODE_PLAIN).copy()
        assert np.array_equal(bwt, O.Oracle(x).bwt(text))
        ctx.ms_run(text); ctx.heads_sort(); ctx.emit(P.MODE_PLAIN)
        sig = ctx.fm_build()
        assert sig == len(np.unique(text))
        pats = []
        for L in (1, 2, 5, 12, 20, 40, 200):
            for _ in range(30):
                a = int(rng.integers(0, len(text) - L))
                p = text[a:a + L]
                if (p == 2).any():
                    continue
                pats.append(p.tobytes())
                q = bytearray(p.tobytes()); q[L // 2] = ord("ACGT"[int(rng.integers(0, 4))])
                pats.append(bytes(q))
        pats += [b"", b"Z", b"ACGTZ", b"N", b"NN"]
        got = ctx.fm_count(pats)
        want = np.array([len(text) if len(p) == 0 else _count(hay, p) for p in pats], np.uint64)
        assert np.array_equal(got, want), np.flatnonzero(got != want)[:10]
    finally:
        ctx.close()
