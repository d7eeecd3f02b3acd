#!/usr/bin/env python3
"""Synthetic highly-similar DNA collections.

reference = i.i.d. uniform ACGT from numpy.random.default_rng(seed); every copy is
derived independently from the reference: each position draws u~U(0,1);
u < d/3 -> substitute by one of the other three bases, d/3 <= u < 2d/3 -> delete,
2d/3 <= u < d -> keep and insert one uniform base after it.  Config 5 style block
events (deletion / tandem duplication / reversal / translocation) are applied
before the point mutations when `svs` > 0.

Files are FASTA, 60 columns, every line newline-terminated, one record per copy
(`>s<k>`); the reference is `>ref`.  `<prefix>.txt` is the two-line list file the
cms_bwt command line takes (reference main.cpp:97-115).

usage: gen_synth.py <ref_bp> <copies> <divergence> <seed> <out_prefix> [svs]
"""
import sys
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def make_reference(n, rng):
    return ACGT[rng.integers(0, 4, n)]


def apply_svs(s, rng, count, lo=1000, hi=50000):
    """count block events, length uniform in [lo, hi]: 30% deletion, 30% tandem
    duplication, 20% reversal, 20% translocation."""
    for _ in range(count):
        n = len(s)
        L = int(rng.integers(lo, hi + 1))
        if L >= n // 2:
            L = max(1, n // 4)
        a = int(rng.integers(0, n - L))
        u = rng.random()
        blk = s[a:a + L]
        if u < 0.3:
            s = np.concatenate([s[:a], s[a + L:]])
        elif u < 0.6:
            s = np.concatenate([s[:a + L], blk, s[a + L:]])
        elif u < 0.8:
            s = np.concatenate([s[:a], blk[::-1], s[a + L:]])
        else:
            rest = np.concatenate([s[:a], s[a + L:]])
            b = int(rng.integers(0, len(rest) + 1))
            s = np.concatenate([rest[:b], blk, rest[b:]])
    return s


def mutate(ref, div, rng):
    n = len(ref)
    r = rng.random(n)
    snp = r < div / 3
    dele = (r >= div / 3) & (r < 2 * div / 3)
    ins = (r >= 2 * div / 3) & (r < div)
    s = ref.copy()
    s[snp] = ACGT[(np.searchsorted(ACGT, s[snp]) + rng.integers(1, 4, int(snp.sum()))) % 4]
    reps = np.ones(n, dtype=np.int64)
    reps[dele] = 0
    reps[ins] = 2
    t = np.repeat(s, reps)
    idx = np.cumsum(reps) - 1
    ip = idx[ins]
    t[ip] = ACGT[rng.integers(0, 4, len(ip))]
    return t


def make_collection(ref_bp, copies, div, seed, svs=0):
    """Returns (ref uint8[n], [copy uint8[...]] * copies)."""
    rng = np.random.default_rng(seed)
    ref = make_reference(ref_bp, rng)
    out = []
    for _ in range(copies):
        base = apply_svs(ref, rng, svs) if svs else ref
        out.append(mutate(base, div, rng))
    return ref, out


def fasta_bytes(name, s, w=60):
    parts = [b">" + name.encode() + b"\n"]
    m = len(s) // w * w
    if m:
        body = np.concatenate(
            [s[:m].reshape(-1, w), np.full((m // w, 1), 10, np.uint8)], axis=1)
        parts.append(body.tobytes())
    if len(s) > m:
        parts.append(s[m:].tobytes() + b"\n")
    return b"".join(parts)


def write_files(prefix, ref, copies):
    with open(prefix + ".ref.fa", "wb") as f:
        f.write(fasta_bytes("ref", ref))
    with open(prefix + ".coll.fa", "wb") as f:
        for k, t in enumerate(copies):
            f.write(fasta_bytes("s%d" % k, t))
    with open(prefix + ".txt", "w") as f:
        f.write(prefix + ".ref.fa\n" + prefix + ".coll.fa\n")


def main():
    n = int(sys.argv[1]); K = int(sys.argv[2]); div = float(sys.argv[3])
    seed = int(sys.argv[4]); out = sys.argv[5]
    svs = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    ref, copies = make_collection(n, K, div, seed, svs)
    write_files(out, ref, copies)


if __name__ == "__main__":
    main()
