"""Multi-GPU host layer of the CMS-BWT path: one process per GPU, documents sharded
whole, three exchange phases between the device steps of include/cmsbwt_b200.h ("shard API").

The reference is single threaded and has no counterpart of this file; what it mirrors is the order
of the reference's stages (CMS-BWT-functions.cpp:1138-1213 MS per document, :1359-1603 head order,
:1618-1777 emit) with the collectives where data has to change owner:

    stage 1   matching statistics of this rank's documents            (no communication)
    stage 2a  tails -> owners by 32-bit key range      all-to-all     (sampled splitters)
              ranks + tie flags back                   all-to-all
              refinement of the still tied tails       all-gather per round (small)
    stage 2b  tile entries -> owners by reference-position range      all-to-all
              bucket sizes of every position range     all-gather
    stage 3   head placements -> owners by output byte range          all-to-all

`ShardedPath` is written against an `ops` object (the device steps) and `torch.distributed` (the
collectives) only, so the same orchestration runs on NCCL with `CudaShardOps` (the product) and,
in tests/, on gloo with a numpy model of the device steps (test infrastructure, CPU).
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import lib, CmsBwtError, MODE_PLAIN

SAMPLES_PER_RANK = 1024
PACK_FROM_WORLD = int(__import__("os").environ.get("CMSBWT_PACK_FROM_WORLD", 4))   # pack paired exchanges into one collective
# from this many ranks on (measured: the packing passes cost more than a collective at 2 ranks)
TIE_CAPACITY = int(__import__("os").environ.get("CMSBWT_TIE_CAPACITY", 1 << 17))             # tied tails per rank that ride along with the count of a refinement round


def shard_documents(doc_lengths, world):
    """Contiguous ranges of whole documents with balanced symbol counts (incl. terminators).
    doc_lengths: symbols per document without its terminator.  -> list of (first, last+1)."""
    total = sum(l + 1 for l in doc_lengths)
    bounds, acc, d = [0], 0, 0
    for r in range(1, world):
        target = total * r // world
        while d < len(doc_lengths) and acc + (doc_lengths[d] + 1) // 2 < target:
            acc += doc_lengths[d] + 1
            d += 1
        bounds.append(d)
    bounds.append(len(doc_lengths))
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def file_shard_cuts(buf, world):
    """Byte ranges of the collection FILE, one per rank, for a sharded ingest: the split points
    size*r/world are moved forward to the next header line (a '>' at a line start), so every range
    but the first starts with a header and every range but the last ends with a newline.
    buf: bytes / mmap.  -> world + 1 offsets (a range may come out empty: the caller falls back)."""
    size = len(buf)
    cuts = [0]
    for r in range(1, world):
        p = max(size * r // world, cuts[-1])
        if p <= 0:
            q = 0
        else:
            k = buf.find(b"\n>", p - 1)
            q = k + 1 if k >= 0 else size
        cuts.append(q)
    cuts.append(size)
    return cuts


def file_shards_usable(buf, cuts):
    """True when every rank's range holds at least one complete line (an unterminated last line is
    dropped by the reference's getline loop, which would leave the last rank without text)."""
    if not all(x < y for x, y in zip(cuts, cuts[1:])):
        return False
    if len(cuts) == 2:
        return True
    k = buf.find(b"\n", cuts[-2])          # the last range: its header line and one more complete line
    return k >= 0 and buf.find(b"\n", k + 1) >= 0


def file_shard_text(buf, cuts, rank, threads=1):
    """The text of rank's byte range under the reference's document rules (:1138-1140, :1257-1275),
    such that the concatenation over the ranks is exactly parse_collection(whole file): the 0x02 a
    header line emits closes the document BEFORE it, so the header that starts a range hands its
    separator to the previous rank.  -> (uint8 text, documents)."""
    import numpy as np
    from . import parse_collection_mt
    world = len(cuts) - 1
    a, b = cuts[rank], cuts[rank + 1]
    sub = np.frombuffer(buf, np.uint8, b - a, a)
    text, nd = parse_collection_mt(sub, threads=threads)
    if rank > 0:                       # the first line is a header: its separator belongs to rank - 1
        if len(text) == 0 or text[0] != 2:
            raise ValueError("range %d does not start with a complete header line" % rank)
        text, nd = text[1:], nd - 1
    if rank < world - 1:
        # an open document at the end was closed by the stand-alone parse (= the next header's 0x02);
        # after a header or blank line the next header still emits one (an empty document)
        j = buf.rfind(b"\n", a, b - 1)
        ls = j + 1 if j >= 0 else a      # start of the range's last line (it ends at b - 1)
        first = buf[ls] if ls < b - 1 else None
        if first is None or first == ord(">"):
            text, nd = np.concatenate([text, np.array([2], np.uint8)]), nd + 1
    return text, nd


def merge_rle_borders(parts):
    """Concatenates per-rank lists of (run, symbol) records, merging a run that straddles a border
    (the -r writer emits maximal runs, CMS-BWT-functions.cpp:1695-1777 / BASELINE.md section 2)."""
    out = []
    for recs in parts:
        for run, sym in recs:
            if out and out[-1][1] == sym:
                out[-1] = (out[-1][0] + run, sym)
            else:
                out.append((run, sym))
    return out


def plan_rle_borders(info):
    """info[r] = (records, first_run, first_sym, last_run, last_sym) of rank r's slice.  A run that
    straddles borders belongs to the leftmost rank it touches.  -> per rank (drop_first, add_to_last):
    whether its first record is absorbed by a rank to the left, and what its own last record gains."""
    W = len(info)
    drop, add = [False] * W, [0] * W
    owner, sym = None, None
    for r, (R, frun, fsym, lrun, lsym) in enumerate(info):
        if R == 0:
            continue
        if owner is not None and fsym == sym:
            drop[r] = True
            add[owner] += frun
            if R == 1:
                continue                    # the whole slice went into the open run; it stays open
        owner, sym = r, lsym
    return drop, add


class DistComm:
    """The two collectives the path needs, on torch.distributed (NCCL on GPUs, gloo in the CPU tests)."""

    def __init__(self, group=None):
        self.group = group

    def all_gather(self, out, inp):
        dist.all_gather_into_tensor(out, inp, group=self.group)

    def all_to_all(self, out, inp, out_splits, in_splits):
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=self.group)


class ThreadComm:
    """Same collectives between `world` threads of ONE process (one context per thread, all on the
    same device): used to run the multi-rank path on a single-GPU box."""

    class Shared:
        def __init__(self, world):
            import threading
            self.world = world
            self.barrier = threading.Barrier(world)
            self.slots = [None] * world

    def __init__(self, shared, rank):
        self.sh, self.rank = shared, rank

    def _exchange(self, item):
        self.sh.slots[self.rank] = item
        self.sh.barrier.wait()
        items = list(self.sh.slots)
        self.sh.barrier.wait()
        return items

    def all_gather(self, out, inp):
        items = self._exchange(inp)
        n = inp.numel()
        for r, t in enumerate(items):
            out[r * n:(r + 1) * n].copy_(t)
        if out.is_cuda:
            torch.cuda.synchronize(out.device)
        self.sh.barrier.wait()                   # nobody may free its input before all have copied

    def all_to_all(self, out, inp, out_splits, in_splits):
        items = self._exchange((inp, list(in_splits)))
        o = 0
        for r, (t, splits) in enumerate(items):
            a = sum(splits[:self.rank])
            out[o:o + out_splits[r]].copy_(t[a:a + splits[self.rank]])
            o += out_splits[r]
        if out.is_cuda:
            torch.cuda.synchronize(out.device)
        self.sh.barrier.wait()


class ShardedPath:
    """Stages 2-3 across ranks.  `ops` has already run stage 1 on this rank's documents."""

    def __init__(self, ops, rank, world, group=None, comm=None):
        self.ops, self.rank, self.world = ops, rank, world
        self.comm = comm if comm is not None else DistComm(group)
        self.comm_bytes = 0
        self.collectives = 0
        self.phase_ms = {}
        self._t = None
        self.pull = True           # stage 3 by SA-rank range where the ops offer it (the numpy model); False: byte ranges

    def _mark(self, name):
        """wall time of the phase that just ended (every device step and collective is synchronous)"""
        import time
        now = time.perf_counter()
        if self._t is not None and name:
            self.phase_ms[name] = self.phase_ms.get(name, 0.0) + (now - self._t) * 1e3
        self._t = now

    # ---- collectives ---------------------------------------------------------------------------
    def _sync(self):
        self.ops.sync()

    def _allgather_ints(self, vals):
        t = self.ops.from_list(vals, torch.int64)
        out = self.ops.empty(len(vals) * self.world, torch.int64)
        self.comm.all_gather(out, t)
        self._sync()
        self.collectives += 1
        return out.cpu().view(self.world, len(vals)).tolist()

    def _a2a_counts(self, send_counts):
        t = self.ops.from_list(send_counts, torch.int64)
        out = self.ops.empty(self.world, torch.int64)
        self.comm.all_to_all(out, t, [1] * self.world, [1] * self.world)
        self._sync()
        self.collectives += 1
        return out.cpu().tolist()

    def _a2a(self, send, send_counts, recv_counts, width=1):
        """all-to-all-v of a flat tensor; counts are in entries of `width` elements."""
        out = self.ops.empty(sum(recv_counts) * width, send.dtype)
        self.comm.all_to_all(out, send, [c * width for c in recv_counts], [c * width for c in send_counts])
        self._sync()
        self.collectives += 1
        self.comm_bytes += (send.numel() + out.numel()) * send.element_size()
        return out

    def _allgather_v(self, t, counts):
        """all-gather of tensors of different lengths (padded to the longest)."""
        mx = max(counts)
        if mx == 0:
            return self.ops.empty(0, t.dtype)
        pad = self.ops.empty(mx, t.dtype)
        pad[:t.numel()] = t
        out = self.ops.empty(mx * self.world, t.dtype)
        self.comm.all_gather(out, pad)
        self._sync()
        self.collectives += 1
        self.comm_bytes += out.numel() * out.element_size()
        return torch.cat([out[r * mx:r * mx + counts[r]] for r in range(self.world)])

    # ---- the path ------------------------------------------------------------------------------
    def run(self, N_local, ndocs_local, plan=None):
        """plan: [(N, ndocs)] of every rank when the caller already knows the shard plan (saves the
        first collective)."""
        ops, W, R = self.ops, self.world, self.rank
        self._mark(None)
        g = [list(x) for x in plan] if plan is not None else self._allgather_ints([N_local, ndocs_local])
        N_total = sum(x[0] for x in g)
        ndocs_total = sum(x[1] for x in g)
        doc_base = sum(x[1] for x in g[:R])
        self.N_total, self.ndocs_total, self.doc_base = N_total, ndocs_total, doc_base

        # -- stage 2a: global order of the tails
        self._mark("x_sizes")
        Tn = ops.tails(doc_base)
        self._mark("d_tails")
        ks, cs = ops.tail_keys(), ops.tail_cls()
        if hasattr(ops, "key_span") and self.pull:
            # as csrc/sharded.cu: FIXED ranges of the key space (keys are SA ranks of the reference << qb: equal
            # ranges hold about equal shares); destination = min(W - 1, key // kper)
            kper = max(1, (ops.key_span() + W - 1) // W)
            splitters = [min(d * kper, 0xFFFFFFFF) for d in range(1, W)]
        else:
            splitters = self._splitters(ks, Tn)
        b = [0] + ops.lower_bound("tail_keys", splitters) + [Tn]
        send = [b[d + 1] - b[d] for d in range(W)]
        cm = self._allgather_ints(send)             # the whole count matrix: cm[src][dst]
        recv = [cm[s_][R] for s_ in range(W)]
        pack = W >= PACK_FROM_WORLD
        if pack:    # key and cls travel as one 64-bit word per tail (one collective instead of two)
            kc = ((ks.to(torch.int64) & 0xFFFFFFFF) << 32) | (cs.to(torch.int64) & 0xFFFFFFFF)
            got = self._a2a(kc, send, recv)
            rk = (got >> 32).to(torch.int32)
            rc = got.to(torch.int32)
        else:
            rk = self._a2a(ks, send, recv)
            rc = self._a2a(cs, send, recv)
        M = sum(recv)
        Ms = [[sum(cm[s_][d] for s_ in range(W))] for d in range(W)]
        rank_base = sum(x[0] for x in Ms[:R])
        if sum(x[0] for x in Ms) >= 2 ** 32:
            raise CmsBwtError(-3, "more than 2^32 tails in the collection")
        self._mark("x_tails")
        r_out, t_out = ops.rank_owner(rk, rc, M, rank_base)
        self._mark("d_rank_owner")
        if pack:
            back = self._a2a((r_out.to(torch.int64) & 0xFFFFFFFF) | (t_out.to(torch.int64) << 32), recv, send)
            rs = back.to(torch.int32)
            ts = (back >> 32).to(torch.uint8)
        else:
            rs = self._a2a(r_out, recv, send)
            ts = self._a2a(t_out, recv, send)
        self._mark("x_ranks")
        m = ops.rank_apply(rs, ts)
        self.rank_rounds = 0
        # refinement rounds: every rank sends [m, keys...] padded to a common capacity, so a round
        # is ONE all-gather as long as every tied subset fits (they only shrink from round to round)
        cap = TIE_CAPACITY
        while True:
            keys = ops.tie_keys() if m else self.ops.empty(0, torch.int64)
            buf = self.ops.empty(cap + 1, torch.int64)
            buf[0] = m
            if 0 < m <= cap:
                buf[1:m + 1] = keys
            allb = self.ops.empty((cap + 1) * W, torch.int64)
            self.comm.all_gather(allb, buf)
            self._sync()
            self.collectives += 1
            allb = allb.view(W, cap + 1)
            ms = [int(v) for v in allb[:, 0].cpu().tolist()]
            if sum(ms) == 0:
                break
            if max(ms) <= cap:
                allk = torch.cat([allb[r, 1:ms[r] + 1] for r in range(W)])
            else:
                allk = self._allgather_v(keys, ms)
            m = ops.tie_refine(allk.contiguous(), sum(ms), sum(ms[:R]))
            cap = max(16, min(cap, max(ms)))
            self.rank_rounds += 1

        # -- stage 2b: tile entries to the owners of the reference-position ranges
        self._mark("d_x_refine")
        tp = ops.tiles_local(W)
        self._mark("d_tiles_local")
        hoff, roff = ops.buffer("tile_hoff"), ops.buffer("tile_roff")
        edge = torch.arange(0, W * tp + 1, tp, device=hoff.device)
        hb = [int(v) & 0xFFFFFFFF for v in hoff[edge].cpu().tolist()]
        rb = [int(v) & 0xFFFFFFFF for v in roff[edge].cpu().tolist()]
        hsend = [hb[d + 1] - hb[d] for d in range(W)]
        rsend = [rb[d + 1] - rb[d] for d in range(W)]
        cm = self._allgather_ints(hsend + rsend)
        hrecv = [cm[s_][R] for s_ in range(W)]
        rrecv = [cm[s_][W + R] for s_ in range(W)]
        he = self._a2a(ops.buffer("tile_heads"), hsend, hrecv, width=2)
        re = self._a2a(ops.buffer("tile_ranges"), rsend, rrecv, width=2)
        rows = (torch.arange(W, device=hoff.device)[:, None] * tp +
                torch.arange(tp + 1, device=hoff.device)[None, :]).reshape(-1)
        eq = [2 * (tp + 1)] * W                     # per destination: its hoff slice, then its roff slice
        both = torch.stack([hoff[rows].view(W, tp + 1), roff[rows].view(W, tp + 1)], dim=1).reshape(-1)
        got2 = self._a2a(both.contiguous(), eq, eq).view(W, 2, tp + 1)
        hoffs = got2[:, 0, :].contiguous().view(-1)
        roffs = got2[:, 1, :].contiguous().view(-1)
        he_slab = [sum(hrecv[:s]) for s in range(W)]
        re_slab = [sum(rrecv[:s]) for s in range(W)]
        self._mark("x_tiles")
        # a tile range deeper than shared memory holds (very many documents over a short reference)
        # is counted by its owner with the sort-based sweep instead of the tile kernel (tiles.cu)
        self.tile_max_entries = ops.tiles_merge(W, R, he, re, hoffs, roffs, he_slab, re_slab, ndocs_total)
        self._mark("d_tiles_merge")
        if hasattr(ops, "rank_pull") and self.pull:
            return self._emit_pull(N_total, ndocs_total, g)
        cov = ops.buffer("cov_slice")
        cov_full = ops.empty(cov.numel() * W, cov.dtype)
        self.comm.all_gather(cov_full, cov)
        self._sync()
        self.collectives += 1
        self.comm_bytes += cov_full.numel() * 4

        # -- stage 3: own output range + placements to the owners of the output ranges
        self._mark("x_cov")
        obeg, olen, psend = ops.emit(cov_full, N_total, W, R, doc_base)
        self._mark("d_emit")
        plc = ops.buffer("placements")
        precv = self._a2a_counts(psend)
        got = self._a2a(plc, psend, precv)
        self._mark("x_placements")
        ops.apply(got, sum(precv))
        self._mark("d_apply")
        self.out_begin, self.out_len = obeg, olen
        return ops.buffer("result")

    def _emit_pull(self, N_total, ndocs_total, g):
        """Stage 3 as the C++ host layer does it (csrc/sharded.cu): output owned by SA-RANK range.  The owner of a
        tile range sends every position's bucket record and head list to the owner of the position's
        SA rank; offsets, fill and run-length records are then local, the slices follow each other in
        rank order.  Only the numpy model offers the two steps (the CUDA steps are behind
        cmsbwt_run_sharded, not in the cmsbwt_shard_* API)."""
        ops, W, R = self.ops, self.world, self.rank
        rec_cnt, hw_cnt, byte_cnt = ops.rank_pull(W, R)
        self._mark("d_place")
        cm = self._allgather_ints(rec_cnt + hw_cnt + byte_cnt)
        rrecv = [cm[s_][R] for s_ in range(W)]
        hrecv = [cm[s_][W + R] for s_ in range(W)]
        olen = [sum(cm[s_][2 * W + d] for s_ in range(W)) for d in range(W)]
        olen[0] += ndocs_total                     # the terminators (one byte per document) lead the output
        if sum(olen) != N_total:
            raise CmsBwtError(-5, "coverage total %d != N %d" % (sum(olen), N_total))
        recs = self._a2a(ops.buffer("vrec_send"), rec_cnt, rrecv, width=2)
        hws = self._a2a(ops.buffer("hw_send"), hw_cnt, hrecv)
        nd = [x[1] for x in g]
        tsend = [nd[R] if d == 0 else 0 for d in range(W)]
        trecv = nd if R == 0 else [0] * W
        terms = self._a2a(ops.buffer("term_send"), tsend, trecv)
        self._mark("x_records")
        ops.emit_pull(W, R, ndocs_total, recs, rrecv, hws, hrecv, terms, olen[R])
        self._mark("d_emit")
        self.out_begin, self.out_len = sum(olen[:R]), olen[R]
        return ops.buffer("result")

    def rle(self):
        """-r: this rank's 9-byte records with the runs that straddle rank borders merged.
        -> (uint8 tensor of records, index of its first record in the global record list)"""
        ops, W, R = self.ops, self.world, self.rank
        nrec = ops.rle()
        rec = ops.buffer("rle")
        if nrec:
            edge = torch.cat([rec[:9], rec[-9:]]).cpu().numpy()
            mine = [nrec, int.from_bytes(edge[0:8].tobytes(), "little"), int(edge[8]),
                    int.from_bytes(edge[9:17].tobytes(), "little"), int(edge[17])]
        else:
            mine = [0, 0, 0, 0, 0]
        info = self._allgather_ints(mine)
        drop, add = plan_rle_borders(info)
        out = rec.clone()
        if drop[R]:
            out = out[9:]
        if add[R]:
            n = out.numel()
            run = int.from_bytes(out[n - 9:n - 1].cpu().numpy().tobytes(), "little") + add[R]
            out[n - 9:n - 1] = torch.tensor(list(run.to_bytes(8, "little")), dtype=torch.uint8, device=out.device)
        counts = [info[r][0] - (1 if drop[r] else 0) for r in range(W)]
        return out, sum(counts[:R])

    def _splitters(self, ks, Tn):
        """W-1 key values from SAMPLES_PER_RANK evenly spaced local samples per rank; equal keys
        always land on one owner because the slabs are cut at lower bounds of key VALUES."""
        W = self.world
        S = SAMPLES_PER_RANK
        if Tn > 0:
            idx = (torch.arange(S, dtype=torch.int64, device=ks.device) * (Tn - 1)) // (S - 1)   # exact: Tn can exceed 2^24
            smp = ks[idx].to(torch.int64) & 0xFFFFFFFF
        else:
            smp = torch.full((S,), 0xFFFFFFFF, dtype=torch.int64, device=ks.device)
        allsmp = self.ops.empty(S * W, torch.int64)
        self.comm.all_gather(allsmp, smp.contiguous())
        self._sync()
        self.collectives += 1
        srt = torch.sort(allsmp).values.cpu().tolist()
        return [srt[(d * S * W) // W] for d in range(1, W)]


# ---- the CUDA device steps ------------------------------------------------------------------------
class _DevArray:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = dict(shape=(n,), typestr=typestr, data=(ptr, False), version=2)


_TYPES = {  # shard buffer -> (elements per entry, typestr, torch dtype)
    "tail_keys": (1, "<i4", torch.int32), "tail_cls": (1, "<i4", torch.int32), "tie_keys": (1, "<i8", torch.int64),
    "tile_heads": (2, "<i8", torch.int64), "tile_ranges": (2, "<i8", torch.int64),
    "tile_hoff": (1, "<i4", torch.int32), "tile_roff": (1, "<i4", torch.int32),
    "cov_slice": (1, "<i4", torch.int32), "placements": (1, "<i8", torch.int64), "result": (1, "|u1", torch.uint8), "rle": (1, "|u1", torch.uint8),
}


class CudaShardOps:
    """Device steps of the shard API on one GPU (ctypes over libcmsbwt_b200.so)."""

    def __init__(self, ctx, device):
        self.ctx, self.h, self.device = ctx, ctx.h, torch.device(device)
        L = lib()
        vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
        L.cmsbwt_shard_tails.argtypes = [vp, u64, C.POINTER(u32)]
        L.cmsbwt_shard_buffer.argtypes = [vp, C.c_char_p, C.POINTER(vp), C.POINTER(u64)]
        L.cmsbwt_shard_lower_bound.argtypes = [vp, C.c_char_p, C.POINTER(u64), u32, C.POINTER(u64)]
        L.cmsbwt_shard_rank_owner.argtypes = [vp, vp, vp, u32, u32, vp, vp]
        L.cmsbwt_shard_rank_apply.argtypes = [vp, vp, vp, C.POINTER(u32)]
        L.cmsbwt_shard_tie_keys.argtypes = [vp, C.POINTER(u32)]
        L.cmsbwt_shard_tie_refine.argtypes = [vp, vp, u32, u32, C.POINTER(u32)]
        L.cmsbwt_shard_tiles_local.argtypes = [vp, u32, C.POINTER(u32)]
        L.cmsbwt_shard_tiles_merge.argtypes = [vp, u32, u32, vp, vp, vp, vp, C.POINTER(u64), C.POINTER(u64), u64,
                                               C.POINTER(u32)]
        L.cmsbwt_shard_emit.argtypes = [vp, vp, u64, u32, u32, u64, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
        L.cmsbwt_shard_apply.argtypes = [vp, vp, u64]
        L.cmsbwt_shard_rle.argtypes = [vp, C.POINTER(u64)]
        self.L = L

    def _ck(self, rc):
        self.ctx._ck(rc)

    def sync(self):
        torch.cuda.current_stream(self.device).synchronize()

    def empty(self, n, dtype):
        return torch.empty(max(int(n), 0), dtype=dtype, device=self.device)

    def from_list(self, vals, dtype):
        return torch.tensor(vals, dtype=dtype, device=self.device)

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr() if t.numel() else 0)

    def buffer(self, name):
        p, n = C.c_void_p(), C.c_uint64()
        self._ck(self.L.cmsbwt_shard_buffer(self.h, name.encode(), C.byref(p), C.byref(n)))
        per, ts, dt = _TYPES[name]
        if n.value == 0 or not p.value:
            return torch.empty(0, dtype=dt, device=self.device)
        return torch.as_tensor(_DevArray(p.value, n.value * per, ts), device=self.device)

    def tails(self, doc_base):
        n = C.c_uint32()
        self._ck(self.L.cmsbwt_shard_tails(self.h, doc_base, C.byref(n)))
        return n.value

    def tail_keys(self):
        return self.buffer("tail_keys")

    def tail_cls(self):
        return self.buffer("tail_cls")

    def lower_bound(self, name, values):
        k = len(values)
        if k == 0:
            return []
        v = (C.c_uint64 * k)(*values)
        o = (C.c_uint64 * k)()
        self._ck(self.L.cmsbwt_shard_lower_bound(self.h, name.encode(), v, k, o))
        return list(o)

    def rank_owner(self, keys, cls, M, rank_base):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        r = self.empty(M, torch.int32)
        t = self.empty(M, torch.uint8)
        self._ck(self.L.cmsbwt_shard_rank_owner(self.h, self._ptr(keys), self._ptr(cls), M, rank_base,
                                                self._ptr(r), self._ptr(t)))
        return r, t

    def rank_apply(self, rs, ts):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        m = C.c_uint32()
        self._ck(self.L.cmsbwt_shard_rank_apply(self.h, self._ptr(rs), self._ptr(ts), C.byref(m)))
        return m.value

    def tie_keys(self):
        m = C.c_uint32()
        self._ck(self.L.cmsbwt_shard_tie_keys(self.h, C.byref(m)))
        return self.buffer("tie_keys")

    def tie_refine(self, allkeys, m_total, my_off):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        m = C.c_uint32()
        self._ck(self.L.cmsbwt_shard_tie_refine(self.h, self._ptr(allkeys), m_total, my_off, C.byref(m)))
        return m.value

    def tiles_local(self, world):
        tp = C.c_uint32()
        self._ck(self.L.cmsbwt_shard_tiles_local(self.h, world, C.byref(tp)))
        return tp.value

    def tiles_merge(self, world, rank, he, re, hoffs, roffs, he_slab, re_slab, ndocs_total):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        me = C.c_uint32()
        hs = (C.c_uint64 * world)(*he_slab)
        rs = (C.c_uint64 * world)(*re_slab)
        self._ck(self.L.cmsbwt_shard_tiles_merge(self.h, world, rank, self._ptr(he), self._ptr(re), self._ptr(hoffs),
                                                 self._ptr(roffs), hs, rs, ndocs_total, C.byref(me)))
        return me.value

    def emit(self, cov_full, N_total, world, rank, doc_base):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        a, b = C.c_uint64(), C.c_uint64()
        cnt = (C.c_uint64 * world)()
        self._ck(self.L.cmsbwt_shard_emit(self.h, self._ptr(cov_full), N_total, world, rank, doc_base,
                                          C.byref(a), C.byref(b), cnt))
        return a.value, b.value, list(cnt)

    def rle(self):
        n = C.c_uint64()
        self._ck(self.L.cmsbwt_shard_rle(self.h, C.byref(n)))
        return n.value

    def apply(self, plc, count):
        self.sync()      # inputs may come from torch ops on torch's stream; the library has its own
        self._ck(self.L.cmsbwt_shard_apply(self.h, self._ptr(plc), count))


def run_sharded(ctx, device, d_text_ptr, N_local, ndocs_local, rank, world, group=None, comm=None, plan=None):
    """Whole path for this rank's documents (text resident in HBM): stage 1 on the shard, then the
    exchange phases.  -> (ShardedPath, result tensor = this rank's contiguous slice of the BWT)."""
    ctx.ms_run_device(d_text_ptr, N_local)
    sp = ShardedPath(CudaShardOps(ctx, device), rank, world, group, comm)
    out = sp.run(N_local, ndocs_local, plan)
    return sp, out


# ---- the C++ host layer (cms_bwt_b200/csrc/sharded.cu): the same path behind one call -----------------
PHASES = ("sizes", "tails", "x_tails", "rank_owner", "x_ranks", "refine", "tiles_local", "x_tiles", "tiles_merge",
          "x_cov", "emit", "x_placements", "apply", "rle", "place", "x_records")


def comm_init(ctx, rank, world, device=None, group=None):
    """NCCL communicator of the library for this rank's context: rank 0 makes the ncclUniqueId, the
    process group (any backend) broadcasts it, every rank calls cmsbwt_comm_init (collective)."""
    L = lib()
    buf = (C.c_uint8 * 128)()
    if rank == 0:
        ctx._ck(L.cmsbwt_nccl_unique_id(buf))
    if world > 1:
        dev = device if (device is not None and dist.get_backend(group) == "nccl") else "cpu"
        t = torch.tensor(list(buf), dtype=torch.uint8, device=dev)
        dist.broadcast(t, 0, group=group)
        buf = (C.c_uint8 * 128)(*t.cpu().tolist())
    ctx._ck(L.cmsbwt_comm_init(ctx.h, buf, rank, world))


class NativeRun:
    """result of run_sharded_native: this rank's slice of the BWT (or its records) and the counters"""
    pass


def run_sharded_native(ctx, device, d_text_ptr, N_local, ndocs_local, mode=MODE_PLAIN):
    """cmsbwt_run_sharded: stage 1 on this rank's documents, then stages 2-3 with the library's own
    NCCL exchanges.  comm_init must have been called on ctx."""
    L = lib()
    ob, ol, tot = C.c_uint64(), C.c_uint64(), C.c_uint64()
    ctx.N = int(N_local)
    ctx._ck(L.cmsbwt_run_sharded(ctx.h, C.c_void_p(int(d_text_ptr)), int(N_local), int(ndocs_local), mode,
                                 C.byref(ob), C.byref(ol), C.byref(tot)))
    r = NativeRun()
    r.out_begin, r.out_len, r.N_total = ob.value, ol.value, tot.value
    dev = torch.device(device)
    p = L.cmsbwt_result_device(ctx.h)
    if mode == MODE_PLAIN:
        r.out = (torch.as_tensor(_DevArray(p, ol.value, "|u1"), device=dev) if ol.value
                 else torch.empty(0, dtype=torch.uint8, device=dev))
        r.records, r.first_record = None, 0
    else:
        rp, nrec, first = C.c_void_p(), C.c_uint64(), C.c_uint64()
        ctx._ck(L.cmsbwt_sharded_rle(ctx.h, C.byref(rp), C.byref(nrec), C.byref(first)))
        r.records = (torch.as_tensor(_DevArray(rp.value, nrec.value * 9, "|u1"), device=dev) if nrec.value
                     else torch.empty(0, dtype=torch.uint8, device=dev))
        r.first_record = first.value
        r.out = None
    coll, xb = C.c_uint64(), C.c_uint64()
    ph = (C.c_double * 16)()
    k = L.cmsbwt_sharded_stats(ctx.h, C.byref(coll), C.byref(xb), ph, 16)
    r.collectives, r.comm_bytes = coll.value, xb.value
    r.phase_ms = {PHASES[i]: ph[i] for i in range(min(k, len(PHASES)))}
    r.rank_rounds = ctx.stats()["rank_rounds"]
    return r


# ---- the command line across GPUs -----------------------------------------------------------------
def main(argv=None):
    """`cms_bwt` over all ranks of a torchrun launch: same options and files as the reference's
    main.cpp:37-132 (-p/-b/-r/-o, two-line list file), every rank writes its slice of <out>.bwt /
    <out>.rl_bwt at its offset.  usage: torchrun --nproc-per-node N tools/cms_bwt_multi.py [opts] list"""
    import argparse
    import os
    import time
    import numpy as np
    from . import Context, read_reference, augment, parse_collection_mt, UINT64_MAX
    ap = argparse.ArgumentParser(prog="cms_bwt_multi")
    ap.add_argument("-p", type=int, default=UINT64_MAX)
    ap.add_argument("-b", type=float, default=2.0)            # accepted and ignored
    ap.add_argument("-r", action="store_true")
    ap.add_argument("-o", default=None)
    ap.add_argument("listfile")
    a = ap.parse_args(argv)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    t0 = time.time()
    with open(a.listfile) as f:
        lines = f.read().split("\n")
    ref_path, coll_path = lines[0], lines[1].rstrip()
    outname = a.o or a.listfile
    with open(ref_path, "rb") as f:
        x = augment(read_reference(f.read()))
    threads = max(1, min(16, (os.cpu_count() or 1) // world))
    import mmap
    with open(coll_path, "rb") as f:
        size = os.fstat(f.fileno()).st_size
        buf = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ) if size else b""
    cuts = file_shard_cuts(buf, world)
    if world > 1 and a.p >= size and file_shards_usable(buf, cuts):
        # sharded ingest: every rank parses only its byte range of the file (cut at header lines)
        text_r, nd_r = file_shard_text(buf, cuts, rank, threads)
        lo, hi, d0, d1 = 0, len(text_r), 0, nd_r
        text = text_r
    else:
        # -p, one rank, or fewer header lines than ranks: whole file on every rank, documents
        # balanced by length
        text, ndocs = parse_collection_mt(buf, a.p, threads)
        ends = np.flatnonzero(text == 2)
        starts = np.concatenate([[0], ends[:-1] + 1])
        plan = shard_documents((ends - starts).tolist(), world)
        d0, d1 = plan[rank]
        lo = int(starts[d0]) if d0 < len(ends) else len(text)
        hi = int(ends[d1 - 1]) + 1 if d1 > d0 else lo
    # decided by all ranks together: a rank that left alone would leave the others waiting in a collective
    empty = torch.tensor([1 if hi == lo else 0], dtype=torch.int64, device=dev if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(empty, op=dist.ReduceOp.MAX)
    if int(empty.item()):
        dist.destroy_process_group()
        raise SystemExit("cms_bwt_multi: more ranks than documents")
    shard = torch.from_numpy(np.concatenate([text[lo:hi], np.full(64, 2, np.uint8)])).to(dev)
    ctx = Context(local)
    ctx.index_upload(x)
    sp, out = run_sharded(ctx, dev, shard.data_ptr(), hi - lo, d1 - d0, rank, world)
    if a.r:
        rec, first = sp.rle()
        data, offset, path = rec.cpu().numpy(), 9 * first, outname + ".rl_bwt"
    else:
        data, offset, path = out.cpu().numpy(), sp.out_begin, outname + ".bwt"
    if rank == 0:
        open(path, "wb").close()
    dist.barrier()
    fd = os.open(path, os.O_WRONLY)
    os.pwrite(fd, memoryview(np.ascontiguousarray(data)), offset)
    os.close(fd)
    dist.barrier()
    if rank == 0:
        print("==== Time elapsed: %d ms" % int((time.time() - t0) * 1e3))
    ctx.close()
    dist.destroy_process_group()
