"""CPU run of the lane state machine (tests/harness/msf_host.cpp, TEST INFRASTRUCTURE) on the all-repeat
reference families of tests/test_gpu_parity.py::test_repetitive_references_medium at reduced size, and
on many short documents: k-mer buckets that overflow, buckets that are empty (resolved from the
neighbours kept in the unused slot, ms_fast.cuh msf_slot_fill), suffixes that touch the augmentation."""
import numpy as np
import pytest

import oracle_lib as O
import gen_synth
from test_ms_fast_host import harness, run_case    # noqa: F401  (fixture + checker)

ACGT = np.frombuffer(b"ACGT", np.uint8)


def _family(kind, rng):
    if kind == "units":
        unit = ACGT[rng.integers(0, 4, 120)]
        ref = np.concatenate([gen_synth.mutate(unit, 0.01, rng) for _ in range(60)])
    elif kind == "period7":
        unit = ACGT[rng.integers(0, 4, 7)]
        ref = np.concatenate([gen_synth.mutate(unit, 0.002, rng) for _ in range(1500)])
    elif kind == "two_letters":          # most k-mers absent: nearly every bucket is empty
        ref = np.frombuffer(b"AT", np.uint8)[rng.integers(0, 2, 6000)].copy()
        ref[[1000, 1001, 4000]] = np.frombuffer(b"CGC", np.uint8)      # all four letters occur (fast path)
    else:                                # short reference, k-mer table much larger than n
        ref = ACGT[rng.integers(0, 4, 700)]
    return ref


@pytest.mark.parametrize("kind", ["units", "period7", "two_letters", "short"])
def test_repeat_families(harness, kind):
    rng = np.random.default_rng(11)
    ref = _family(kind, rng)
    x = O.augment(ref)
    copies = [gen_synth.mutate(ref, 0.01, rng) for _ in range(4)]
    # novel sequence too: random ACGT that shares no long k-mer with a two-letter reference
    copies.append(ACGT[rng.integers(0, 4, 800)])
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    for k in ((4, 6) if kind != "short" else (5, 7)):
        for WL in (64, 1024):
            assert run_case(harness, x, text, WL, k=k, skip_min=4, desc="%s k=%d WL=%d" % (kind, k, WL)) is not None
