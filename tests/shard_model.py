"""numpy MODEL of the device steps of the multi-GPU shard API (TEST INFRASTRUCTURE, CPU only).

`NumpyShardOps` offers the same methods as cms_bwt_b200.multi.CudaShardOps, so that the host layer
(cms_bwt_b200.multi.ShardedPath: shard plan, split sizes, collectives, refinement loop) can be run
with world_size 2 on the gloo backend and checked against the oracle.  Every step restates what the
CUDA kernels of cms_bwt_b200/csrc do (file:function named per method); stage 1 comes from the
oracle's per-position matching statistics.  Nothing here is ever used by the product.
"""
import numpy as np
import torch

import model_pipeline as MP

LT_SHIFT, LT_W = 8, 256                 # tiles.cu
FILL_BYTES = 256 * 4 * 16               # emit.cu
U32 = 0xFFFFFFFF


def _i32(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32).copy())


def _i64(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64).copy())


def _u32(t):
    return t.numpy().view(np.uint32).astype(np.int64)


def _u64(t):
    return t.numpy().view(np.uint64)


class NumpyShardOps:
    def __init__(self, oracle, text):
        self.x, self.n = oracle.x, oracle.n
        self.sa = oracle.sa.astype(np.int64)
        self.isa = np.empty(self.n, np.int64)
        self.isa[self.sa] = np.arange(self.n)
        self.ref_bwt = oracle.ref_bwt
        self.text = np.asarray(text, dtype=np.uint8)
        pos, ln, sm, hd = oracle.ms(self.text)                     # stage 1 (ms_fast.cu / ms.cu)
        self.h = MP.heads_from_ms(self.text, pos.astype(np.int64), ln.astype(np.int64), sm.astype(np.int64), hd, self.n)
        self.buf = {}

    # ---- plumbing ------------------------------------------------------------------------------
    def sync(self):
        pass

    def empty(self, n, dtype):
        return torch.zeros(max(int(n), 0), dtype=dtype)

    def from_list(self, vals, dtype):
        return torch.tensor(vals, dtype=dtype)

    def buffer(self, name):
        return self.buf[name]

    def tail_keys(self):
        return self.buf["tail_keys"]

    def tail_cls(self):
        return self.buf["tail_cls"]

    # ---- sharded.cu: the span of the 32-bit tail keys (SA rank << qb | quantised cls order)
    def key_span(self):
        isa_bits = 1
        while isa_bits < 32 and (1 << isa_bits) < self.n:
            isa_bits += 1
        kb = isa_bits + min(6, 32 - isa_bits)
        return min(1 << kb, (self.n + 1) << (kb - isa_bits))

    def lower_bound(self, name, values):
        a = _u32(self.buf[name]) if name == "tail_keys" else _u64(self.buf[name])
        return [int(np.searchsorted(a, np.array(v, dtype=a.dtype), side="left")) for v in values]

    # ---- heads.cu: sort_prepare (k_link, k_tail_first, k_tail_label) + tailrank.cu: shard_tails_sort
    def tails(self, doc_base):
        h, n = self.h, self.n
        g, pos, ln, sm = h["gstart"], h["pos"], h["len"], h["sm"]
        H = len(g)
        sent = ln == 0
        nxt_start = np.concatenate([g[1:], [len(self.text)]])
        self.toNext = np.where(sent, 0, nxt_start - g - 1)
        g_end = g + ln
        self.cls = np.where(sm == 1, pos + ln, U32 - (pos + ln))
        tflag = np.concatenate([[True], g_end[1:] != g_end[:-1]]) if H else np.zeros(0, bool)
        self.tail_id = np.cumsum(tflag) - 1
        first = np.flatnonzero(tflag)
        Tn = len(first)
        e = g_end[first]
        j = np.searchsorted(g, e, side="right") - 1
        self.sexcl = np.cumsum(sent) - sent
        term = ln[j] == 0
        pe = np.where(term, 0, pos[j] + (e - g[j]))
        r = self.isa[np.minimum(pe, n - 1)]
        isa_bits = 1
        while isa_bits < 32 and (1 << isa_bits) < n:
            isa_bits += 1
        qb = min(6, 32 - isa_bits)
        Q, half = 1 << qb, (1 << qb) >> 1
        d = doc_base + self.sexcl[j]
        rest = (pos[j] + ln[j] - pe) >> 5
        f = np.minimum(rest, half - 1) if half else np.zeros(Tn, np.int64)
        key = np.where(term, np.minimum(d, Q - 1), (r << qb) | np.where(sm[j] == 1, f, Q - 1 - f))
        self.label_lo = np.where(term, d, self.cls[j])
        self.nexth = np.where(term, np.arange(Tn), self.tail_id[j])
        self.Tn = Tn
        self.perm = np.argsort(key, kind="stable")
        self.buf["tail_keys"] = _i32(key[self.perm])
        self.buf["tail_cls"] = _i32(self.label_lo[self.perm])
        self.trank = np.zeros(Tn, np.int64)
        return Tn

    # ---- tailrank.cu: shard_rank_owner (tail_rank_ex, cls_only)
    def rank_owner(self, keys, cls, M, rank_base):
        k, c = _u32(keys), _u32(cls)
        full = (k.astype(np.uint64) << np.uint64(32)) | c.astype(np.uint64)
        order = np.argsort(full, kind="stable")
        sf = full[order]
        idx = np.arange(M)
        newf = np.concatenate([[True], sf[1:] != sf[:-1]]) if M else np.zeros(0, bool)
        first = np.maximum.accumulate(np.where(newf, idx, 0)) if M else idx
        nxt_same = np.concatenate([sf[1:] == sf[:-1], [False]]) if M else np.zeros(0, bool)
        tie_sorted = (~newf) | nxt_same
        rank = np.empty(M, np.int64)
        tie = np.zeros(M, np.uint8)
        rank[order] = rank_base + first
        tie[order] = tie_sorted
        return _i32(rank), torch.from_numpy(tie)

    # ---- tailrank.cu: shard_rank_apply
    def rank_apply(self, rs, ts):
        self.trank[self.perm] = _u32(rs)
        self.sub = self.perm[ts.numpy() != 0]
        return len(self.sub)

    # ---- tailrank.cu: shard_tie_keys (k_tr_key_next)
    def tie_keys(self):
        t = self.sub
        key = (self.trank[t].astype(np.uint64) << np.uint64(32)) | self.trank[self.nexth[t]].astype(np.uint64)
        self.buf["tie_keys"] = _i64(key)
        return self.buf["tie_keys"]

    # ---- tailrank.cu: shard_tie_refine
    def tie_refine(self, allkeys, m_total, my_off):
        key = _u64(allkeys)
        order = np.argsort(key, kind="stable")
        sk = key[order]
        idx = np.arange(m_total)
        hi = sk >> np.uint64(32)
        newg = np.concatenate([[True], hi[1:] != hi[:-1]])
        news = np.concatenate([[True], sk[1:] != sk[:-1]])
        grp = np.maximum.accumulate(np.where(newg, idx, 0))
        sg = np.maximum.accumulate(np.where(news, idx, 0))
        nr = np.empty(m_total, np.int64)
        nt = np.zeros(m_total, bool)
        nr[order] = hi.astype(np.int64) + (sg - grp)
        nt[order] = (~news) | np.concatenate([sk[1:] == sk[:-1], [False]])
        m = len(self.sub)
        self.trank[self.sub] = nr[my_off:my_off + m]
        sub2 = self.sub[nt[my_off:my_off + m]]
        jump = self.nexth[self.nexth[sub2]]
        self.nexth[sub2] = jump
        self.sub = sub2
        return len(sub2)

    # ---- heads.cu: sort_kappa + tiles.cu: shard_tiles_local (k_tile_hist, k_tile_scatter)
    def tiles_local(self, world):
        h = self.h
        pos, ln, ch = h["pos"], h["len"], h["ch"].astype(np.int64)
        kappa = (self.cls.astype(np.uint64) << np.uint64(32)) | self.trank[self.tail_id].astype(np.uint64)
        ntiles = (self.n >> LT_SHIFT) + 1
        tp = (ntiles + world - 1) // world
        nt = tp * world
        real = np.flatnonzero(ln != 0)
        ht = pos[real] >> LT_SHIFT
        o = np.argsort(ht, kind="stable")
        hi_ = real[o]
        he = np.empty((len(hi_), 2), np.uint64)
        he[:, 0] = hi_.astype(np.uint64) | (((pos[hi_] & (LT_W - 1)) | (ch[hi_] << 8)).astype(np.uint64) << np.uint64(32))
        he[:, 1] = kappa[hi_]
        hoff = np.concatenate([[0], np.cumsum(np.bincount(ht, minlength=nt))])
        rt, ra, rk = [], [], []
        for i in real[self.toNext[real] > 0]:
            p, tn = int(pos[i]), int(self.toNext[i])
            for t in range((p + 1) >> LT_SHIFT, ((p + tn) >> LT_SHIFT) + 1):
                a = t << LT_SHIFT
                lo, hi = max(p + 1, a) - a, min(p + tn, a + LT_W - 1) - a
                rt.append(t); ra.append(lo | (hi << 16)); rk.append(kappa[i])
        rt = np.array(rt, np.int64)
        o = np.argsort(rt, kind="stable")
        re = np.empty((len(rt), 2), np.uint64)
        re[:, 0] = np.array(ra, np.uint64)[o]
        re[:, 1] = np.array(rk, np.uint64)[o]
        roff = np.concatenate([[0], np.cumsum(np.bincount(rt, minlength=nt))])
        self.tp = tp
        self.buf["tile_heads"] = _i64(he.reshape(-1))
        self.buf["tile_ranges"] = _i64(re.reshape(-1))
        self.buf["tile_hoff"] = _i32(hoff)
        self.buf["tile_roff"] = _i32(roff)
        return tp

    # ---- tiles.cu: shard_tiles_merge (k_sh_tile_counts, k_sh_tile_merge, k_tile_cov)
    def tiles_merge(self, world, rank, he, re, hoffs, roffs, he_slab, re_slab, ndocs_total):
        tp, n = self.tp, self.n
        he = _u64(he).reshape(-1, 2)
        re = _u64(re).reshape(-1, 2)
        ho = _u32(hoffs).reshape(world, tp + 1)
        ro = _u32(roffs).reshape(world, tp + 1)
        self.rank, self.tiles = rank, []
        cov = np.zeros(tp * LT_W, np.int64)
        maxent = 0
        for t in range(tp):
            hs = [he[he_slab[s] + ho[s, t] - ho[s, 0]: he_slab[s] + ho[s, t + 1] - ho[s, 0]] for s in range(world)]
            rs = [re[re_slab[s] + ro[s, t] - ro[s, 0]: re_slab[s] + ro[s, t + 1] - ro[s, 0]] for s in range(world)]
            th, tr = np.concatenate(hs), np.concatenate(rs)
            self.tiles.append((th, tr))
            maxent = max(maxent, len(th) + len(tr))
            q = ((th[:, 0] >> np.uint64(32)) & np.uint64(LT_W - 1)).astype(np.int64)
            np.add.at(cov, t * LT_W + q, 1)
            for a in tr[:, 0].astype(np.int64):
                cov[t * LT_W + (a & 0xFFFF): t * LT_W + (a >> 16) + 1] += 1
        p_last = n - 1 - rank * tp * LT_W
        if 0 <= p_last < tp * LT_W:
            cov[p_last] += ndocs_total
        valid = (np.arange(tp * LT_W) + rank * tp * LT_W) < n
        cov[~valid] = 0
        self.buf["cov_slice"] = _i32(cov)
        return maxent

    # ---- emit.cu: shard_emit (k_cnt_sa, k_fill) + tiles.cu: shard_place (k_tile_place)
    def emit(self, cov_full, N_total, world, rank, doc_base):
        n = self.n
        cov = _u32(cov_full)
        cnt = cov[self.sa]
        base = np.concatenate([[0], np.cumsum(cnt)])
        assert int(base[-1]) == N_total, (int(base[-1]), N_total)
        nblocks = (N_total + FILL_BYTES - 1) // FILL_BYTES
        bp = (nblocks + world - 1) // world
        b0 = min(nblocks, rank * bp)
        b1 = min(nblocks, b0 + bp)
        chunk, obeg = bp * FILL_BYTES, b0 * FILL_BYTES
        olen = min(N_total, b1 * FILL_BYTES) - min(N_total, obeg)
        self.obeg, self.olen = obeg, olen
        self.out = np.repeat(self.ref_bwt, cnt)[obeg:obeg + olen].copy()
        plc = []
        for t, (th, tr) in enumerate(self.tiles):
            hq = ((th[:, 0] >> np.uint64(32)) & np.uint64(LT_W - 1)).astype(np.int64)
            hc = ((th[:, 0] >> np.uint64(40)) & np.uint64(0xFF)).astype(np.int64)
            lo = (tr[:, 0].astype(np.int64)) & 0xFFFF
            hi = (tr[:, 0].astype(np.int64)) >> 16
            for k in range(len(th)):
                q, kap = hq[k], th[k, 1]
                w = int(np.sum((hq == q) & (th[:, 1] < kap))) + int(np.sum((tr[:, 1] < kap) & (lo <= q) & (q <= hi)))
                p = (self.rank * self.tp + t) * LT_W + q
                plc.append(((int(base[self.isa[p]]) + w) << 8) | int(hc[k]))
        h = self.h
        for i in np.flatnonzero(h["len"] == 0):
            plc.append(((doc_base + int(self.sexcl[i])) << 8) | int(h["ch"][i]))
        plc = np.sort(np.array(plc, np.uint64))
        self.buf["placements"] = _i64(plc)
        dest = (plc >> np.uint64(8)).astype(np.int64) // max(chunk, 1)
        return obeg, olen, np.bincount(dest, minlength=world)[:world].astype(np.int64).tolist()

    # ---- tiles.cu: shard_rank_pull (k_tile_rank with the routing table, k_sh_hw_route): bucket records and
    # head lists of this rank's reference positions, grouped by the owner of the position's SA rank, in
    # SA-rank order.  Record = two 64-bit words: size | first head (relative to the destination's
    # slab) << 32, heads | (SA rank - 1) << 32.
    def rank_pull(self, world, rank):
        n, tp = self.n, self.tp
        rper = (n - 1 + world - 1) // world
        self.rper = rper
        recs = []                                   # (sa rank - 1, size, [(within, sym)...])
        for t, (th, tr) in enumerate(self.tiles):
            hq = ((th[:, 0] >> np.uint64(32)) & np.uint64(LT_W - 1)).astype(np.int64)
            hc = ((th[:, 0] >> np.uint64(40)) & np.uint64(0xFF)).astype(np.int64)
            lo = (tr[:, 0].astype(np.int64)) & 0xFFFF
            hi = (tr[:, 0].astype(np.int64)) >> 16
            for q in range(LT_W):
                p = (rank * tp + t) * LT_W + q
                if p >= n or self.isa[p] == 0:
                    continue                        # SA rank 0 is the terminators' bucket
                at = np.flatnonzero(hq == q)
                covq = int(np.sum((lo <= q) & (q <= hi)))
                heads = []
                for k in at:
                    kap = th[k, 1]
                    w = int(np.sum(th[at, 1] < kap)) + int(np.sum((tr[:, 1] < kap) & (lo <= q) & (q <= hi)))
                    heads.append((w, int(hc[k])))
                recs.append((int(self.isa[p]) - 1, covq + len(at), sorted(heads)))
        recs.sort()
        dest = [r // rper for r, _, _ in recs]
        rec_cnt = np.bincount(dest, minlength=world).astype(np.int64)
        hw_cnt = np.zeros(world, np.int64)
        byte_cnt = np.zeros(world, np.int64)
        words, hws = [], []
        for (r, size, heads), d in zip(recs, dest):
            words.append(size | (int(hw_cnt[d]) << 32))
            words.append(len(heads) | (r << 32))
            hws += [(w << 8) | sym for w, sym in heads]
            hw_cnt[d] += len(heads)
            byte_cnt[d] += size
        self.buf["vrec_send"] = _i64(np.array(words, np.uint64))
        self.buf["hw_send"] = _i64(np.array(hws, np.uint64))
        h = self.h
        term = np.zeros(int(np.sum(h["len"] == 0)), np.uint8)
        for i in np.flatnonzero(h["len"] == 0):
            term[int(self.sexcl[i])] = h["ch"][i]
        self.buf["term_send"] = torch.from_numpy(term)
        return rec_cnt.tolist(), hw_cnt.tolist(), byte_cnt.tolist()

    # ---- emit.cu: shard_emit_pull (k_sh_vscatter, then the single-GPU emit on this rank's SA ranks)
    def emit_pull(self, world, rank, ndocs_total, recs, rec_counts, hws, hw_counts, terms, expect_bytes):
        n, rper = self.n, self.rper
        r0 = rank * rper
        nranks = 0 if r0 >= n - 1 else min(rper, n - 1 - r0)
        assert sum(rec_counts) == nranks, (sum(rec_counts), nranks)
        w = _u64(recs).reshape(-1, 2)
        hw = _u64(hws)
        hw_base = np.concatenate([[0], np.cumsum(hw_counts)])
        src = np.repeat(np.arange(world), rec_counts)
        size = np.zeros(nranks, np.int64); first = np.zeros(nranks, np.int64); nh = np.zeros(nranks, np.int64)
        seen = np.zeros(nranks, bool)
        for j in range(len(w)):
            r = int(w[j, 1] >> np.uint64(32)) - r0
            assert 0 <= r < nranks and not seen[r]
            seen[r] = True
            size[r] = int(w[j, 0] & np.uint64(U32))
            first[r] = int(w[j, 0] >> np.uint64(32)) + int(hw_base[src[j]])
            nh[r] = int(w[j, 1] & np.uint64(U32))
        assert seen.all()
        parts = []
        if rank == 0:
            t = terms.numpy()
            assert len(t) == ndocs_total
            parts.append(t.copy())
        for r in range(nranks):
            b = np.full(size[r], self.ref_bwt[r0 + r + 1], np.uint8)
            for x in hw[first[r]:first[r] + nh[r]]:
                b[int(x >> np.uint64(8))] = int(x & np.uint64(0xFF))
            parts.append(b)
        self.out = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
        assert len(self.out) == expect_bytes, (len(self.out), expect_bytes)
        self.buf["result"] = torch.from_numpy(self.out)

    # ---- emit.cu: shard_apply
    def apply(self, plc, count):
        p = _u64(plc)[:count]
        off = (p >> np.uint64(8)).astype(np.int64) - self.obeg
        assert ((off >= 0) & (off < self.olen)).all()
        self.out[off] = (p & np.uint64(0xFF)).astype(np.uint8)
        self.buf["result"] = torch.from_numpy(self.out)

    # ---- emit.cu: rle_encode on the local slice
    def rle(self):
        import oracle_lib as O
        rec = O.rle(self.out) if len(self.out) else np.zeros(0, np.uint8)
        self.buf["rle"] = torch.from_numpy(np.ascontiguousarray(rec))
        return len(rec) // 9
