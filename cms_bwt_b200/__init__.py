"""cms_bwt_b200 -- ctypes binding of libcmsbwt_b200.so (include/cmsbwt_b200.h).

The library is the product; this module only loads it and mirrors the reference's interface for
the path: `compute_bwt(args, ref_file, coll_file)` stands where the reference's
`computeBWT(Args, refFile, collFile)` stands (CMS-BWT.h:69, CMS-BWT-functions.cpp:1788-1797), with
the same Args fields (utils.h:8-14: prefixLength, buffer, format, outname).  There is no CPU
fallback: importing works without a GPU (so the symbol table can be checked), every compute call
fails loudly if the shared library or a CUDA device is missing.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcmsbwt_b200.so")

MODE_PLAIN, MODE_RLE = 0, 1
UINT64_MAX = 2 ** 64 - 1
RLE_STATIC_MAX = 1 << 30          # cmsbwt_rle_encode_static: symbols per call (include/cmsbwt_b200.h)

EXPORTS = [
    "cmsbwt_ctx_create", "cmsbwt_ctx_destroy", "cmsbwt_last_error", "cmsbwt_get_stats",
    "cmsbwt_augment", "cmsbwt_read_reference", "cmsbwt_parse_collection", "cmsbwt_build_sa",
    "cmsbwt_index_upload", "cmsbwt_index_download", "cmsbwt_ms_run", "cmsbwt_ms_run_device",
    "cmsbwt_ms_run_device_ordered",
    "cmsbwt_set_ms_dump", "cmsbwt_ms_download", "cmsbwt_heads_sort", "cmsbwt_heads_download",
    "cmsbwt_emit", "cmsbwt_result_device", "cmsbwt_result_download", "cmsbwt_run",
    "cmsbwt_set_param",
    "cmsbwt_shard_tails", "cmsbwt_shard_buffer", "cmsbwt_shard_lower_bound", "cmsbwt_shard_rank_owner",
    "cmsbwt_shard_rank_apply", "cmsbwt_shard_tie_keys", "cmsbwt_shard_tie_refine",
    "cmsbwt_shard_tiles_local", "cmsbwt_shard_tiles_merge", "cmsbwt_shard_emit", "cmsbwt_shard_apply",
    "cmsbwt_shard_rle",
    "cmsbwt_nccl_unique_id", "cmsbwt_comm_init", "cmsbwt_comm_destroy", "cmsbwt_run_sharded",
    "cmsbwt_sharded_after_ms", "cmsbwt_sharded_rle", "cmsbwt_sharded_stats",
    "cmsbwt_fm_build", "cmsbwt_fm_count", "cmsbwt_fm_free", "cmsbwt_rle_encode_static",
    "cmsbwt_parse_collection_mt", "cmsbwt_parse_piece", "cmsbwt_line_start",
    "cmsbwt_host_alloc", "cmsbwt_host_free", "cmsbwt_text_begin", "cmsbwt_text_append",
    "cmsbwt_ms_run_staged", "cmsbwt_result_read", "cmsbwt_copy_fence", "cmsbwt_copy_wait",
]


class Stats(C.Structure):
    _fields_ = [("sa_build_ms", C.c_double), ("index_build_ms", C.c_double), ("ms_ms", C.c_double),
                ("sort_ms", C.c_double), ("emit_ms", C.c_double), ("h2d_ms", C.c_double),
                ("d2h_ms", C.c_double), ("ms_kernel_ms", C.c_double), ("n_heads", C.c_uint64), ("n_docs", C.c_uint64),
                ("n_tails", C.c_uint64), ("n_items", C.c_uint64), ("n_runs", C.c_uint64),
                ("rank_rounds", C.c_uint32), ("kmer_k", C.c_uint32),
                ("kernel_launches", C.c_uint64), ("pack_ms", C.c_double),
                ("fast_path", C.c_uint32), ("light_tiles", C.c_uint32), ("pair_tests", C.c_uint64),
                ("rank_tied", C.c_uint64), ("group_ms", C.c_double * 12), ("n_escapes", C.c_uint64)]

    GROUPS = ("pack", "ms", "gather", "rank", "tiles", "offsets", "fill", "place", "rle")

    def asdict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k != "group_ms"}
        d["group_ms"] = {n: self.group_ms[i] for i, n in enumerate(self.GROUPS)}
        return d


class CmsBwtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("cmsbwt error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    """Loads libcmsbwt_b200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "or `make -C cms_bwt_b200/csrc`" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, u64, i64 = C.c_void_p, C.c_uint64, C.c_int64
        L.cmsbwt_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.cmsbwt_ctx_destroy.argtypes = [vp]
        L.cmsbwt_ctx_destroy.restype = None
        L.cmsbwt_last_error.argtypes = [vp]
        L.cmsbwt_last_error.restype = C.c_char_p
        L.cmsbwt_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.cmsbwt_augment.argtypes = [vp, i64, vp]
        L.cmsbwt_augment.restype = i64
        L.cmsbwt_read_reference.argtypes = [vp, i64, vp]
        L.cmsbwt_read_reference.restype = i64
        L.cmsbwt_parse_collection.argtypes = [vp, i64, u64, vp, C.POINTER(i64)]
        L.cmsbwt_parse_collection.restype = i64
        L.cmsbwt_build_sa.argtypes = [vp, vp, C.c_uint32, vp]
        L.cmsbwt_index_upload.argtypes = [vp, vp, C.c_uint32, vp]
        L.cmsbwt_index_download.argtypes = [vp, vp, vp, vp, vp]
        L.cmsbwt_ms_run.argtypes = [vp, vp, u64]
        L.cmsbwt_ms_run_device.argtypes = [vp, vp, u64]
        L.cmsbwt_ms_run_device_ordered.argtypes = [vp, vp, u64, vp]
        L.cmsbwt_set_ms_dump.argtypes = [vp, C.c_int]
        L.cmsbwt_ms_download.argtypes = [vp, vp, vp, vp, vp]
        L.cmsbwt_heads_sort.argtypes = [vp]
        L.cmsbwt_heads_download.argtypes = [vp, vp, vp, vp, vp, vp]
        L.cmsbwt_emit.argtypes = [vp, C.c_int, C.POINTER(u64)]
        L.cmsbwt_result_device.argtypes = [vp]
        L.cmsbwt_result_device.restype = vp
        L.cmsbwt_result_download.argtypes = [vp, vp, u64]
        L.cmsbwt_run.argtypes = [vp, vp, u64, C.c_int, vp, u64, C.POINTER(u64)]
        L.cmsbwt_set_param.argtypes = [vp, C.c_char_p, i64]
        L.cmsbwt_parse_collection_mt.argtypes = [vp, i64, u64, vp, C.POINTER(i64), C.c_int]
        L.cmsbwt_parse_collection_mt.restype = i64
        L.cmsbwt_parse_piece.argtypes = [vp, i64, i64, vp, C.POINTER(i64), C.POINTER(C.c_int32)]
        L.cmsbwt_parse_piece.restype = i64
        L.cmsbwt_line_start.argtypes = [vp, i64, i64]
        L.cmsbwt_line_start.restype = i64
        L.cmsbwt_host_alloc.argtypes = [C.POINTER(vp), u64]
        L.cmsbwt_host_free.argtypes = [vp]
        L.cmsbwt_host_free.restype = None
        L.cmsbwt_text_begin.argtypes = [vp, u64]
        L.cmsbwt_text_append.argtypes = [vp, vp, u64]
        L.cmsbwt_ms_run_staged.argtypes = [vp]
        L.cmsbwt_result_read.argtypes = [vp, vp, u64, u64]
        L.cmsbwt_copy_fence.argtypes = [vp, C.c_int]
        L.cmsbwt_copy_wait.argtypes = [vp, C.c_int]
        L.cmsbwt_fm_build.argtypes = [vp, C.POINTER(C.c_uint32)]
        L.cmsbwt_fm_count.argtypes = [vp, vp, vp, C.c_uint32, vp]
        L.cmsbwt_fm_free.argtypes = [vp]
        L.cmsbwt_rle_encode_static.argtypes = [vp, vp, u64, vp, u64, vp, C.POINTER(u64), vp]
        L.cmsbwt_nccl_unique_id.argtypes = [vp]
        L.cmsbwt_comm_init.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
        L.cmsbwt_comm_destroy.argtypes = [vp]
        L.cmsbwt_run_sharded.argtypes = [vp, vp, u64, u64, C.c_int, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
        L.cmsbwt_sharded_after_ms.argtypes = [vp, u64, C.c_int, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
        L.cmsbwt_sharded_rle.argtypes = [vp, C.POINTER(vp), C.POINTER(u64), C.POINTER(u64)]
        L.cmsbwt_sharded_stats.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(C.c_double), C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ---- host loaders (reference semantics; pure host code inside the library) -----------------
def read_reference(file_bytes):
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.empty(len(src) + 1, np.uint8)
    m = lib().cmsbwt_read_reference(_p(src), len(src), _p(out))
    if m < 0:
        raise CmsBwtError(m, "read_reference")
    return out[:m].copy()


def augment(ref):
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    out = np.empty(len(ref) + 128, np.uint8)
    m = lib().cmsbwt_augment(_p(ref), len(ref), _p(out))
    if m < 0:
        raise CmsBwtError(m, "reference holds a reserved byte (0,1,2 or >127)")
    return out[:m].copy()


def parse_collection(file_bytes, prefix=UINT64_MAX):
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.empty(len(src) + 2, np.uint8)
    nd = C.c_int64(0)
    N = lib().cmsbwt_parse_collection(_p(src), len(src), C.c_uint64(prefix), _p(out), C.byref(nd))
    if N < 0:
        raise CmsBwtError(N, "parse_collection")
    return out[:N].copy(), nd.value


def parse_collection_mt(file_bytes, prefix=UINT64_MAX, threads=0):
    """cmsbwt_parse_collection on several host threads (same bytes, same document count)."""
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.empty(len(src) + 2, np.uint8)
    nd = C.c_int64(0)
    N = lib().cmsbwt_parse_collection_mt(_p(src), len(src), C.c_uint64(prefix), _p(out), C.byref(nd),
                                         threads)
    if N < 0:
        raise CmsBwtError(N, "parse_collection_mt")
    return out[:N].copy(), nd.value


def parse_pieces(file_bytes, piece):
    """The streamed-ingest form of the parse: pieces of about `piece` bytes cut at line starts,
    each parsed by cmsbwt_parse_piece; yields the text bytes of every piece, the closing separator
    last.  Concatenated they equal parse_collection(file_bytes)[0]."""
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    L = lib()
    size, a = len(src), 0
    op = C.c_int32(0)
    nd = C.c_int64(0)
    while a < size:
        b = L.cmsbwt_line_start(_p(src), size, min(size, a + max(1, piece)))
        out = np.empty(max(1, b - a), np.uint8)
        w = L.cmsbwt_parse_piece(_p(src), a, b, _p(out), C.byref(nd), C.byref(op))
        if w < 0:
            raise CmsBwtError(w, "parse_piece")
        yield out[:w]
        a = b
    if op.value == 1:
        yield np.array([2], np.uint8)


class Context:
    """One GPU context (one per device)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        rc = lib().cmsbwt_ctx_create(device, C.byref(self.h))
        if rc != 0:
            self.h = None
            raise CmsBwtError(rc, "no usable CUDA device %d (this package has no CPU fallback)" % device)
        self.n = 0
        self.N = 0

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.cmsbwt_ctx_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise CmsBwtError(rc, lib().cmsbwt_last_error(self.h).decode())

    def set_param(self, name, value):
        self._ck(lib().cmsbwt_set_param(self.h, name.encode(), int(value)))

    def stats(self):
        s = Stats()
        self._ck(lib().cmsbwt_get_stats(self.h, C.byref(s)))
        return s.asdict()

    def build_sa(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint8)
        sa = np.empty(len(x), np.int32)
        self._ck(lib().cmsbwt_build_sa(self.h, _p(x), len(x), _p(sa)))
        return sa

    def index_upload(self, x, sa=None):
        x = np.ascontiguousarray(x, dtype=np.uint8)
        if sa is not None:
            sa = np.ascontiguousarray(sa, dtype=np.int32)
            assert len(sa) == len(x)
        self._ck(lib().cmsbwt_index_upload(self.h, _p(x), len(x), _p(sa)))
        self.n = len(x)

    def index_download(self):
        n = self.n
        sa = np.empty(n, np.int32); isa = np.empty(n, np.uint32)
        pp = np.empty(n, np.uint32); rb = np.empty(n, np.uint8)
        self._ck(lib().cmsbwt_index_download(self.h, _p(sa), _p(isa), _p(pp), _p(rb)))
        return dict(sa=sa, isa=isa, plcp_prime=pp, ref_bwt=rb)

    def set_ms_dump(self, on=True):
        self._ck(lib().cmsbwt_set_ms_dump(self.h, 1 if on else 0))

    def ms_run(self, text):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        self.N = len(text)
        self._ck(lib().cmsbwt_ms_run(self.h, _p(text), len(text)))

    def ms_run_device(self, dptr, N):
        self.N = int(N)
        self._ck(lib().cmsbwt_ms_run_device(self.h, C.c_void_p(int(dptr)), int(N)))

    def ms_run_device_ordered(self, dptr, N, producer_stream):
        """producer_stream: raw cudaStream_t (int; 0 = legacy default stream) the text was written on"""
        self.N = int(N)
        self._ck(lib().cmsbwt_ms_run_device_ordered(self.h, C.c_void_p(int(dptr)), int(N),
                                                    C.c_void_p(int(producer_stream))))

    def ms_download(self):
        N = self.N
        pos = np.empty(N, np.uint32); ln = np.empty(N, np.uint32)
        sm = np.empty(N, np.uint8); hd = np.empty(N, np.uint8)
        self._ck(lib().cmsbwt_ms_download(self.h, _p(pos), _p(ln), _p(sm), _p(hd)))
        return pos, ln, sm, hd

    def heads_sort(self):
        self._ck(lib().cmsbwt_heads_sort(self.h))

    def heads_download(self, within=False, offset=False):
        H = self.stats()["n_heads"]
        g = np.empty(H, np.uint64); pos = np.empty(H, np.uint32); ln = np.empty(H, np.uint32)
        w = np.empty(H, np.uint64) if within else None
        o = np.empty(H, np.uint64) if offset else None
        self._ck(lib().cmsbwt_heads_download(self.h, _p(g), _p(pos), _p(ln), _p(w), _p(o)))
        return dict(gstart=g, pos=pos, len=ln, within=w, offset=o)

    def emit(self, mode=MODE_PLAIN):
        ob = C.c_uint64(0)
        self._ck(lib().cmsbwt_emit(self.h, mode, C.byref(ob)))
        self.out_bytes = ob.value
        return ob.value

    def result_device(self):
        return lib().cmsbwt_result_device(self.h)

    def result_download(self, out=None):
        if out is None:
            out = np.empty(self.out_bytes, np.uint8)
        self._ck(lib().cmsbwt_result_download(self.h, _p(out), out.nbytes))
        return out[:self.out_bytes]

    def fm_build(self):
        """FM-index over the plain BWT the last emit left on the device -> number of distinct symbols"""
        k = C.c_uint32()
        self._ck(lib().cmsbwt_fm_build(self.h, C.byref(k)))
        return k.value

    def fm_count(self, patterns):
        """occurrences of every pattern (bytes / uint8 arrays) in the collection, by backward search on the GPU"""
        pats = [np.frombuffer(p, np.uint8) if isinstance(p, (bytes, bytearray)) else np.ascontiguousarray(p, np.uint8)
                for p in patterns]
        off = np.zeros(len(pats) + 1, np.uint64)
        off[1:] = np.cumsum([len(p) for p in pats])
        flat = np.concatenate(pats) if pats else np.zeros(0, np.uint8)
        flat = np.ascontiguousarray(np.concatenate([flat, np.zeros(1, np.uint8)]))
        out = np.zeros(len(pats), np.uint64)
        self._ck(lib().cmsbwt_fm_count(self.h, _p(flat), _p(off), len(pats), _p(out)))
        return out

    def rle_encode_static(self, d_bwt, N, d_records, cap_records, d_n_records=None, stream=None, wait=True):
        """9-byte run-length records of the device buffer d_bwt[0, N) into the device buffer d_records, with no
        allocation in the library (static scratch, N <= RLE_STATIC_MAX); device pointers as integers.
        wait: also return the number of records (synchronises the stream); else the call only enqueues."""
        k = C.c_uint64(0)
        self._ck(lib().cmsbwt_rle_encode_static(self.h, d_bwt, N, d_records, cap_records, d_n_records,
                                                C.byref(k) if wait else None, stream))
        return k.value if wait else None

    def run(self, text, mode=MODE_PLAIN, out=None):
        """ms_run + heads_sort + emit + download on host buffers (the drop-in call)."""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        self.N = len(text)
        if out is None:
            out = np.empty(len(text) if mode == MODE_PLAIN else 9 * len(text), np.uint8)
        ob = C.c_uint64(0)
        self._ck(lib().cmsbwt_run(self.h, _p(text), len(text), mode, _p(out), out.nbytes, C.byref(ob)))
        self.out_bytes = ob.value
        return out[:ob.value]


    def run_streamed(self, pieces, capacity, mode=MODE_PLAIN, chunk=1 << 20):
        """The path with streamed ingest / egress (what the cms_bwt command line does): `pieces`
        (uint8 arrays) are appended to the device text through two pinned staging slots, the result
        is read back `chunk` bytes at a time through two more."""
        L = lib()
        slots = []
        try:
            for _ in range(4):
                hp = C.c_void_p()
                self._ck(L.cmsbwt_host_alloc(C.byref(hp), chunk))
                slots.append(hp)
            self._ck(L.cmsbwt_text_begin(self.h, capacity))
            k = 0
            for pc in pieces:
                pc = np.ascontiguousarray(pc, dtype=np.uint8)
                for a in range(0, len(pc), chunk):
                    s = k & 1
                    k += 1
                    self._ck(L.cmsbwt_copy_wait(self.h, s))
                    m = min(chunk, len(pc) - a)
                    C.memmove(slots[s], pc[a:a + m].ctypes.data, m)
                    self._ck(L.cmsbwt_text_append(self.h, slots[s], m))
                    self._ck(L.cmsbwt_copy_fence(self.h, s))
            self._ck(L.cmsbwt_ms_run_staged(self.h))
            self._ck(L.cmsbwt_heads_sort(self.h))
            ob = C.c_uint64(0)
            self._ck(L.cmsbwt_emit(self.h, mode, C.byref(ob)))
            self.out_bytes = ob.value
            out = np.empty(ob.value, np.uint8)
            pend = [None, None]
            for j, a in enumerate(range(0, ob.value, chunk)):
                s = 2 + (j & 1)
                if pend[j & 1] is not None:
                    self._ck(L.cmsbwt_copy_wait(self.h, s))
                    pa, pm = pend[j & 1]
                    C.memmove(out[pa:pa + pm].ctypes.data, slots[s], pm)
                m = min(chunk, ob.value - a)
                self._ck(L.cmsbwt_result_read(self.h, slots[s], a, m))
                self._ck(L.cmsbwt_copy_fence(self.h, s))
                pend[j & 1] = (a, m)
            for i in (0, 1):
                if pend[i] is not None:
                    self._ck(L.cmsbwt_copy_wait(self.h, 2 + i))
                    pa, pm = pend[i]
                    C.memmove(out[pa:pa + pm].ctypes.data, slots[2 + i], pm)
            return out
        finally:
            for hp in slots:
                L.cmsbwt_host_free(hp)


class Args:
    """The reference's Args (utils.h:8-14)."""

    def __init__(self, filename="", outname="", prefixLength=UINT64_MAX, buffer=2, format=0):
        self.filename, self.outname = filename, outname or filename
        self.prefixLength, self.buffer, self.format = prefixLength, buffer, format


def compute_bwt(args, ref_file, coll_file, device=0, sa=None):
    """Mirror of computeBWT (CMS-BWT-functions.cpp:1788-1797): reads the two files, builds the
    index, runs the path on the GPU and writes <outname>.bwt or <outname>.rl_bwt.  Returns stats."""
    with open(ref_file, "rb") as f:
        ref = read_reference(f.read())
    x = augment(ref)
    with open(coll_file, "rb") as f:
        text, ndocs = parse_collection(f.read(), args.prefixLength)
    ctx = Context(device)
    try:
        ctx.index_upload(x, sa)
        out = ctx.run(text, MODE_RLE if args.format else MODE_PLAIN)
        with open(args.outname + (".rl_bwt" if args.format else ".bwt"), "wb") as f:
            f.write(out.tobytes())
        return ctx.stats()
    finally:
        ctx.close()
