"""Host-side MODEL of the data-parallel formulation the CUDA path implements (TEST CODE).

It takes per-position matching statistics (from the oracle) and restates stages 2-3 the way
cms_bwt_b200/csrc does them -- tails, label doubling, tile items with interval masks, offsets --
in numpy/pure Python, so the *algorithm* can be checked against the golden vectors on the CPU
box, and so the CUDA intermediate arrays can be diffed against it on the GPU box.

Formulation:
  head i (text order): gstart, pos, len, sm, ch; sentinel heads (len = 0, pos = n-1) at every
  document terminator.  g_end = gstart + len.  Heads with equal g_end share a TAIL (the text
  suffix at g_end).  Tail label = (ISA[pos_e], cls_j) where j is the phrase containing g_end and
  pos_e its reference position; terminal tails (g_end is a terminator) get (0, doc index).
  Tail rank = rank of the label sequence along next-tail pointers (prefix doubling).
  kappa_i = (cls_i, tailrank[tail(i)]),  cls = sm ? E : 0xFFFFFFFF - E,  E = pos + len.
  within(h) = #{phrases p : pos_p <= pos_h <= pos_p + toNext_p, kappa_p < kappa_h}
  offset(h) = base[ISA[pos_h]] + within(h);  out = fill(refBWT by coverage) then scatter heads.
"""
import numpy as np


def heads_from_ms(text, pos, ln, sm, hd, n):
    """-> dict of text-order head arrays incl. sentinels (from oracle per-position MS)."""
    text = np.asarray(text)
    is_term = text == 2
    flag = (hd.astype(bool)) | is_term
    idx = np.flatnonzero(flag)
    h = dict(gstart=idx.astype(np.int64), pos=pos[idx].astype(np.int64),
             len=ln[idx].astype(np.int64), sm=sm[idx].astype(np.int64))
    prev = np.where(idx > 0, text[np.maximum(idx - 1, 0)], 2).astype(np.uint8)
    h["ch"] = prev
    h["pos"][is_term[idx]] = n - 1
    h["len"][is_term[idx]] = 0
    h["sm"][is_term[idx]] = 0
    return h


def pipeline(text, h, x, sa, W=64):
    n = len(x)
    isa = np.empty(n, np.int64)
    isa[sa] = np.arange(n)
    H = len(h["gstart"])
    gstart, pos, ln, sm, ch = h["gstart"], h["pos"], h["len"], h["sm"], h["ch"]
    sent = ln == 0
    nxt_start = np.concatenate([gstart[1:], [len(text)]])
    toNext = np.where(sent, 0, nxt_start - gstart - 1)
    g_end = gstart + ln
    E = pos + ln
    cls = np.where(sm == 1, E, 0xFFFFFFFF - E)
    # coverage
    diff = np.zeros(n + 1, np.int64)
    np.add.at(diff, pos, 1)
    np.add.at(diff, pos + toNext + 1, -1)
    cov = np.cumsum(diff)[:n]
    # tails
    tflag = np.concatenate([[True], g_end[1:] != g_end[:-1]])
    tail_id = np.cumsum(tflag) - 1
    first = np.flatnonzero(tflag)
    Tn = len(first)
    e = g_end[first]
    j = np.searchsorted(gstart, e, side="right") - 1
    docidx = np.cumsum(sent) - 1
    term = ln[j] == 0
    pos_e = np.where(term, 0, pos[j] + (e - gstart[j]))
    lab_hi = np.where(term, 0, isa[np.minimum(pos_e, n - 1)])
    lab_lo = np.where(term, docidx[j], cls[j])
    label = (lab_hi.astype(np.uint64) << np.uint64(32)) | lab_lo.astype(np.uint64)
    nxt = np.where(term, np.arange(Tn), tail_id[j])
    # doubling
    def dense(keys):
        order = np.argsort(keys, kind="stable")
        sk = keys[order]
        newflag = np.concatenate([[0], (sk[1:] != sk[:-1]).astype(np.int64)])
        r = np.empty(len(keys), np.int64)
        r[order] = np.cumsum(newflag)
        return r, int(newflag.sum()) + 1
    rank, distinct = dense(label)
    rounds = 0
    while distinct < Tn:
        key = (rank.astype(np.uint64) << np.uint64(32)) | rank[nxt].astype(np.uint64)
        rank, d2 = dense(key)
        nxt = nxt[nxt]
        rounds += 1
        assert d2 > distinct or d2 == Tn or rounds < 64
        distinct = d2
    kappa = (cls.astype(np.uint64) << np.uint64(32)) | rank[tail_id].astype(np.uint64)
    # tile items
    end = pos + toNext
    t0 = pos // W
    t1 = end // W
    within = np.zeros(H, np.int64)
    items = {}
    for i in range(H):
        for t in range(t0[i], t1[i] + 1):
            a = t * W
            lo = max(pos[i], a) - a
            hi = min(end[i], a + W - 1) - a
            items.setdefault(t, []).append((int(kappa[i]), lo, hi, i if t == t0[i] else -1))
    for t, lst in items.items():
        lst.sort(key=lambda z: z[0])
        run = np.zeros(W, np.int64)
        for (k, lo, hi, hid) in lst:
            if hid >= 0:
                within[hid] = run[pos[hid] - t * W]
            run[lo:hi + 1] += 1
    # offsets
    cnt_sa = cov[sa]
    base = np.concatenate([[0], np.cumsum(cnt_sa)])
    N = int(base[-1])
    ref_bwt = np.where(sa > 0, x[np.maximum(sa - 1, 0)], 0).astype(np.uint8)
    out = np.repeat(ref_bwt, cnt_sa)
    off = base[isa[pos]] + within
    out[off] = ch
    return dict(out=out, N=N, within=within, off=off, tailrank=rank, tail_id=tail_id,
                kappa=kappa, cov=cov, toNext=toNext, rounds=rounds)
