"""ctypes binding of oracle/liboracle.so (TEST INFRASTRUCTURE; see oracle/cms_oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
_lib = None


def build(force=False):
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "cms_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"])
    return so


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        u8p, i32p, u32p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_int32),
                                 C.POINTER(C.c_uint32), C.POINTER(C.c_int64))
        L.cms_oracle_augment.restype = C.c_int64
        L.cms_oracle_augment.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.cms_oracle_read_reference.restype = C.c_int64
        L.cms_oracle_read_reference.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.cms_oracle_parse_collection.restype = C.c_int64
        L.cms_oracle_parse_collection.argtypes = [C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p,
                                                  C.c_void_p, C.c_void_p]
        L.cms_oracle_create.restype = C.c_void_p
        L.cms_oracle_create.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.cms_oracle_destroy.argtypes = [C.c_void_p]
        for name, rt in (("sa", i32p), ("isa", u32p), ("lcp", i32p), ("plcp_prime", i32p),
                         ("ref_bwt", u8p)):
            f = getattr(L, "cms_oracle_" + name)
            f.restype = rt
            f.argtypes = [C.c_void_p]
        L.cms_oracle_ms.restype = C.c_int
        L.cms_oracle_ms.argtypes = [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 4
        L.cms_oracle_bwt.restype = C.c_int
        L.cms_oracle_bwt.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.cms_oracle_rle.restype = C.c_int64
        L.cms_oracle_rle.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def read_reference(file_bytes):
    """reference :155-204 -> raw sequence bytes (before trimming/augmentation)."""
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.empty(len(src) + 1, dtype=np.uint8)
    m = lib().cms_oracle_read_reference(_p(src), len(src), _p(out))
    return out[:m].copy()


def augment(ref):
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    out = np.empty(len(ref) + 128, dtype=np.uint8)
    m = lib().cms_oracle_augment(_p(ref), len(ref), _p(out))
    return out[:m].copy()


def parse_collection(file_bytes, prefix=2 ** 64 - 1):
    """-> (text uint8[N] with 0x02 terminators, doc_start int64[D+1])."""
    src = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.empty(len(src) + 2, dtype=np.uint8)
    ds = np.empty(len(src) + 3, dtype=np.int64)
    nd = C.c_int64(0)
    N = lib().cms_oracle_parse_collection(_p(src), len(src), C.c_uint64(prefix), _p(out), _p(ds),
                                          C.byref(nd))
    return out[:N].copy(), ds[:nd.value + 1].copy()


def docs_to_text(docs):
    """list of uint8 arrays -> concatenated text with 0x02 terminators."""
    parts = []
    for d in docs:
        parts.append(np.asarray(d, dtype=np.uint8))
        parts.append(np.array([2], dtype=np.uint8))
    return np.concatenate(parts) if parts else np.zeros(0, np.uint8)


class Oracle:
    def __init__(self, x_aug, sa=None):
        self.x = np.ascontiguousarray(x_aug, dtype=np.uint8)
        self.n = len(self.x)
        sa_p = None
        if sa is not None:
            self._sa_in = np.ascontiguousarray(sa, dtype=np.int32)
            sa_p = _p(self._sa_in)
        self.h = lib().cms_oracle_create(_p(self.x), self.n, sa_p)

    def __del__(self):
        if getattr(self, "h", None):
            lib().cms_oracle_destroy(self.h)
            self.h = None

    def _arr(self, name, count, dtype):
        ptr = getattr(lib(), "cms_oracle_" + name)(self.h)
        return np.ctypeslib.as_array(ptr, shape=(count,)).astype(dtype, copy=True)

    @property
    def sa(self): return self._arr("sa", self.n, np.int32)
    @property
    def isa(self): return self._arr("isa", self.n, np.uint32)
    @property
    def lcp(self): return self._arr("lcp", self.n + 1, np.int32)
    @property
    def plcp_prime(self): return self._arr("plcp_prime", self.n, np.int32)
    @property
    def ref_bwt(self): return self._arr("ref_bwt", self.n, np.uint8)

    def ms(self, text):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        N = len(text)
        pos = np.zeros(N, np.uint32); ln = np.zeros(N, np.uint32)
        sm = np.zeros(N, np.uint8); hd = np.zeros(N, np.uint8)
        lib().cms_oracle_ms(self.h, _p(text), N, _p(pos), _p(ln), _p(sm), _p(hd))
        return pos, ln, sm, hd

    def bwt(self, text):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        N = len(text)
        out = np.zeros(N, np.uint8)
        nh = C.c_uint64(0)
        rc = lib().cms_oracle_bwt(self.h, _p(text), N, _p(out), C.byref(nh))
        if rc != 0:
            raise RuntimeError("oracle bwt size mismatch")
        self.last_heads = nh.value
        return out


def rle(bwt):
    bwt = np.ascontiguousarray(bwt, dtype=np.uint8)
    R = lib().cms_oracle_rle(_p(bwt), len(bwt), None)
    out = np.zeros(9 * R, np.uint8)
    lib().cms_oracle_rle(_p(bwt), len(bwt), _p(out))
    return out


def rle_decode(rec):
    rec = np.asarray(rec, dtype=np.uint8).reshape(-1, 9)
    runs = rec[:, :8].copy().view("<u8").ravel()
    return np.repeat(rec[:, 8], runs.astype(np.int64))


# ---- the real reference (oracle/_ref), when it has been built ------------------------------

def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "cmsbwt"))


def have_gsabwt():
    return os.path.exists(os.path.join(REF_DIR, "gsabwt"))


def run_reference(ref_path, coll_path, workdir, rle_mode=False, prefix=None, name="refrun"):
    """Runs the unmodified reference binary; returns the output bytes (.bwt or .rl_bwt)."""
    lst = os.path.join(workdir, name + ".txt")
    with open(lst, "w") as f:
        f.write(ref_path + "\n" + coll_path + "\n")
    out = os.path.join(workdir, name)
    cmd = [os.path.join(REF_DIR, "cmsbwt"), "-o", out]
    if rle_mode:
        cmd.append("-r")
    if prefix is not None:
        cmd += ["-p", str(prefix)]
    cmd.append(lst)
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
    with open(out + (".rl_bwt" if rle_mode else ".bwt"), "rb") as f:
        return np.frombuffer(f.read(), dtype=np.uint8)


def run_gsabwt(coll_path, workdir, name="gsa"):
    out = os.path.join(workdir, name + ".bwt")
    subprocess.run([os.path.join(REF_DIR, "gsabwt"), coll_path, out], check=True,
                   stderr=subprocess.DEVNULL)
    with open(out, "rb") as f:
        return np.frombuffer(f.read(), dtype=np.uint8)


_sais = None


def libsais_sa(x):
    """Suffix array via the reference's vendored libsais (oracle/_ref/libsais.so) -- test helper."""
    global _sais
    if _sais is None:
        _sais = C.CDLL(os.path.join(REF_DIR, "libsais.so"))
        _sais.libsais.restype = C.c_int32
        _sais.libsais.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    x = np.ascontiguousarray(x, dtype=np.uint8)
    sa = np.empty(len(x), dtype=np.int32)
    rc = _sais.libsais(_p(x), _p(sa), len(x), 0, None)
    assert rc == 0
    return sa
