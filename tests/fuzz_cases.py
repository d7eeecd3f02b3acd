"""Adversarial small inputs for the parity tests (families of SURVEY.md §8c O2): tandem
references of period 1/2/3/7, 2-5 symbol alphabets, block deletions / tandem duplications /
reversals / translocations, exact duplicate documents, empty documents, blank-line splits,
unterminated last lines, header-less collections, raw references ending in "$\\n", non-ACGT
symbols (N, lower case)."""
import numpy as np


def _fasta(name, s, w):
    out = [b">" + name + b"\n"]
    if w is None:
        out.append(bytes(s) + b"\n")
    else:
        for i in range(0, len(s), w):
            out.append(bytes(s[i:i + w]) + b"\n")
    return b"".join(out)


def _mutate(rng, s, rate, alpha):
    s = list(s)
    out = []
    for c in s:
        u = rng.random()
        if u < rate / 3:
            out.append(int(alpha[rng.integers(0, len(alpha))]))
        elif u < 2 * rate / 3:
            continue
        elif u < rate:
            out.append(c)
            out.append(int(alpha[rng.integers(0, len(alpha))]))
        else:
            out.append(c)
    return np.array(out, dtype=np.uint8)


def _sv(rng, s):
    n = len(s)
    if n < 8:
        return s
    L = int(rng.integers(1, max(2, n // 3)))
    a = int(rng.integers(0, n - L))
    blk = s[a:a + L]
    u = rng.random()
    if u < 0.3:
        return np.concatenate([s[:a], s[a + L:]])
    if u < 0.6:
        return np.concatenate([s[:a + L], blk, s[a + L:]])
    if u < 0.8:
        return np.concatenate([s[:a], blk[::-1], s[a + L:]])
    rest = np.concatenate([s[:a], s[a + L:]])
    b = int(rng.integers(0, len(rest) + 1))
    return np.concatenate([rest[:b], blk, rest[b:]])


def make_case(seed):
    """-> dict(ref_file=bytes, coll_file=bytes, desc=str)."""
    rng = np.random.default_rng(seed)
    sigma = int(rng.integers(2, 6))
    alpha = np.frombuffer(b"ACGTN"[:sigma] if rng.random() < 0.8 else b"acgtn"[:sigma], dtype=np.uint8)
    n = int(rng.integers(1, 400)) if rng.random() < 0.7 else int(rng.integers(400, 3000))
    kind = rng.integers(0, 4)
    if kind == 0:
        period = int(rng.choice([1, 2, 3, 7]))
        unit = alpha[rng.integers(0, sigma, period)]
        ref = np.tile(unit, n // period + 1)[:n]
    elif kind == 1:
        half = alpha[rng.integers(0, sigma, max(1, n // 2))]
        ref = np.concatenate([half, half])[:max(1, n)]
    else:
        ref = alpha[rng.integers(0, sigma, n)]
    ndocs = int(rng.integers(1, 9))
    rate = float(rng.choice([0.0, 0.01, 0.05, 0.2]))
    docs = []
    for _ in range(ndocs):
        u = rng.random()
        if u < 0.15 and docs:
            docs.append(docs[int(rng.integers(0, len(docs)))].copy())     # exact duplicate
            continue
        if u < 0.2:
            docs.append(alpha[rng.integers(0, sigma, int(rng.integers(1, 50)))])  # unrelated
            continue
        s = ref.copy()
        for _ in range(int(rng.integers(0, 3))):
            s = _sv(rng, s)
        s = _mutate(rng, s, rate, alpha)
        if len(s) == 0:
            s = alpha[rng.integers(0, sigma, 1)]
        if rng.random() < 0.1:
            s = np.concatenate([s, np.frombuffer(b"NnX", dtype=np.uint8)[rng.integers(0, 3, 2)]])
        docs.append(s)
    w = [10, 60, None][int(rng.integers(0, 3))]
    # reference file: FASTA (possibly multi-record) or raw (possibly ending in "$\n")
    u = rng.random()
    if u < 0.6:
        ref_file = _fasta(b"ref", ref, w)
    elif u < 0.75 and len(ref) > 4:
        k = len(ref) // 2
        ref_file = _fasta(b"r1", ref[:k], w) + _fasta(b"r2 more", ref[k:], w)
    elif u < 0.9:
        ref_file = bytes(ref) + b"$\n"
    else:
        ref_file = bytes(ref) + b"\n"
    if ref_file[:1] == b">" and False:
        pass
    if ref_file[:1] != b">" and ref[0] == ord(">"):
        ref_file = _fasta(b"ref", ref, w)
    # collection file
    parts = []
    headerless = rng.random() < 0.1
    for k, d in enumerate(docs):
        if headerless:
            parts.append(bytes(d) + b"\n\n")          # blank line closes the document
        else:
            parts.append(_fasta(b"s%d" % k, d, w))
            if rng.random() < 0.1:
                parts.append(b"\n")                    # blank line -> an empty document
    coll = b"".join(parts)
    if rng.random() < 0.15 and coll.endswith(b"\n"):
        coll = coll[:-1]                               # unterminated last line (dropped)
    return dict(ref_file=ref_file, coll_file=coll,
                desc="seed=%d sigma=%d n=%d docs=%d rate=%g kind=%d w=%s" %
                     (seed, sigma, len(ref), ndocs, rate, kind, w))
