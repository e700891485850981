"""Host-side mirror of the reference API over the C ABI of libnmf_b200.so (include/nmf.h).

Same names and argument meaning as the reference's README.md:31-54: `matrix`, `read_matrix`,
`write_matrix`, `update_div(W, H, X, CONVERGE_THRESH, max_iter, t, verbose)`.  Everything here is
ctypes plumbing: all arithmetic runs in the library's CUDA kernels, and if the library or a GPU is
missing the calls raise — there is no CPU fallback.
"""
import ctypes
import os

import numpy as np

from . import _build

_f32p = ctypes.POINTER(ctypes.c_float)
_f64p = ctypes.POINTER(ctypes.c_double)

EXACT, REFCOMPAT = 0, 1
ENGINE_TCGEN05, ENGINE_SIMT = 0, 1


class matrix(ctypes.Structure):
    """`typedef struct { float *mat; float *mat_d; int dim[2]; } matrix;` (column-major)."""
    _fields_ = [("mat", _f32p), ("mat_d", _f32p), ("dim", ctypes.c_int * 2)]

    @classmethod
    def from_numpy(cls, a):
        a = np.asfortranarray(a, dtype=np.float32)
        m = cls()
        m.mat = a.ctypes.data_as(_f32p)
        m.mat_d = None
        m.dim[0], m.dim[1] = a.shape
        m._keep = a
        return m

    def to_numpy(self):
        n = self.dim[0] * self.dim[1]
        flat = np.ctypeslib.as_array(self.mat, shape=(n,))
        return flat.reshape((self.dim[0], self.dim[1]), order="F")


class nmf_options(ctypes.Structure):
    _fields_ = [("semantics", ctypes.c_int), ("engine", ctypes.c_int), ("device", ctypes.c_int),
                ("check_every", ctypes.c_int), ("use_graph", ctypes.c_int), ("tf32_passes", ctypes.c_int),
                ("restarts", ctypes.c_int), ("reserved", ctypes.c_int * 9)]


class nmf_result(ctypes.Structure):
    _fields_ = [("iterations", ctypes.c_int), ("converged", ctypes.c_int), ("cost", ctypes.c_double),
                ("device_ms", ctypes.c_double), ("gpu_launches", ctypes.c_longlong)]


class NmfError(RuntimeError):
    pass


_lib = None

# every symbol include/nmf.h declares (tests/test_abi.py checks the export list against the header)
_SIGNATURES = {
    "read_matrix": (None, [ctypes.POINTER(matrix), ctypes.c_char_p]),
    "write_matrix": (None, [matrix, ctypes.c_char_p]),
    "create_matrix": (None, [ctypes.POINTER(matrix), ctypes.c_int, ctypes.c_int, ctypes.c_float]),
    "destroy_matrix": (None, [ctypes.POINTER(matrix)]),
    "update_div": (None, [matrix, matrix, matrix, ctypes.c_float, ctypes.c_int, _f64p, ctypes.c_int]),
    "nmf_default_options": (None, [ctypes.POINTER(nmf_options)]),
    "nmf_set_error_mode": (ctypes.c_int, [ctypes.c_int]),
    "nmf_last_error": (ctypes.c_char_p, []),
    "nmf_version": (ctypes.c_char_p, []),
    "nmf_update_div_ex": (ctypes.c_int, [matrix, matrix, matrix, ctypes.c_float, ctypes.c_int, _f64p, ctypes.c_int,
                                         ctypes.POINTER(nmf_options), ctypes.POINTER(nmf_result)]),
    "nmf_update_div_batch": (ctypes.c_int, [ctypes.POINTER(matrix), ctypes.POINTER(matrix), ctypes.c_int, matrix, ctypes.c_float,
                                            ctypes.c_int, _f64p, ctypes.c_int, ctypes.POINTER(nmf_options),
                                            ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), _f64p]),
    "nmf_session_create": (ctypes.c_void_p, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_longlong,
                                             ctypes.c_longlong, ctypes.POINTER(nmf_options)]),
    "nmf_session_destroy": (None, [ctypes.c_void_p]),
    "nmf_session_upload": (ctypes.c_int, [ctypes.c_void_p, _f32p, _f32p, _f32p]),
    "nmf_session_download": (ctypes.c_int, [ctypes.c_void_p, _f32p, _f32p]),
    "nmf_session_upload_restart": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _f32p, _f32p]),
    "nmf_session_download_restart": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _f32p, _f32p]),
    "nmf_session_restart_results": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), _f64p]),
    "nmf_session_download_cols": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.c_int, _f32p, _f32p]),
    "nmf_session_download_rows": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.c_int, _f32p, _f32p]),
    "nmf_session_half_step": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nmf_session_fill_uniform": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_ulonglong, ctypes.c_int]),
    "nmf_nccl_unique_id": (ctypes.c_int, [ctypes.c_void_p]),
    "nmf_session_join": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "nmf_session_share_comm": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "nmf_session_fill_planted": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_float]),
    "nmf_session_fused_steps": (ctypes.c_int, [ctypes.c_void_p]),
    "nmf_session_corr_bf16": (ctypes.c_int, [ctypes.c_void_p]),
    "nmf_session_run": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_float, ctypes.c_int, _f64p, ctypes.c_int,
                                       ctypes.POINTER(nmf_result)]),
    "nmf_session_cost": (ctypes.c_int, [ctypes.c_void_p, _f64p]),
    "nmf_dbg_quotient": (ctypes.c_int, [_f32p, _f32p, _f32p, _f32p, _f64p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.POINTER(nmf_options)]),
    "nmf_dbg_h_step": (ctypes.c_int, [_f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                      ctypes.POINTER(nmf_options)]),
    "nmf_dbg_w_step": (ctypes.c_int, [_f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                      ctypes.POINTER(nmf_options)]),
    "nmf_dbg_sums": (ctypes.c_int, [_f32p, _f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.POINTER(nmf_options)]),
    "nmf_rowsum_noalloc": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "nmf_dbg_gemm": (ctypes.c_int, [ctypes.c_int, _f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.POINTER(nmf_options)]),
}


def lib_path():
    return _build.LIB


def lib():
    """Loads (building if needed) libnmf_b200.so; raises if it cannot — no fallback."""
    global _lib
    if _lib is None:
        path = os.environ.get("NMF_B200_LIB") or _build.build()      # override: A/B builds of the same ABI (development)
        L = ctypes.CDLL(path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the library misses a declared symbol
            fn.restype, fn.argtypes = res, args
        L.nmf_set_error_mode(1)            # status codes -> Python exceptions (C hosts keep exit())
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise NmfError(lib().nmf_last_error().decode(errors="replace"))


def options(semantics="exact", engine="tcgen05", device=-1, check_every=25, use_graph=True, tf32_passes=3, restarts=1):
    o = nmf_options()
    lib().nmf_default_options(ctypes.byref(o))
    o.semantics = REFCOMPAT if semantics in ("refcompat", REFCOMPAT) else EXACT
    o.engine = ENGINE_SIMT if engine in ("simt", ENGINE_SIMT) else ENGINE_TCGEN05
    o.device, o.check_every, o.use_graph, o.tf32_passes = device, check_every, int(use_graph), tf32_passes
    o.restarts = int(restarts)
    return o


def _p(a):
    return None if a is None else a.ctypes.data_as(_f32p)


def _f(a):
    return np.asfortranarray(a, dtype=np.float32)


# ------------------------------------------------------------------------------------------
# the reference's API
# ------------------------------------------------------------------------------------------
def read_matrix(path):
    """File -> column-major numpy array (values < EPS raised to EPS, as cuda/nmf.cu:188-218)."""
    m = matrix()
    lib().read_matrix(ctypes.byref(m), os.fsencode(path))
    if not m.mat:
        raise NmfError(lib().nmf_last_error().decode(errors="replace"))
    out = m.to_numpy().copy(order="F")
    lib().destroy_matrix(ctypes.byref(m))
    return out


def write_matrix(a, path):
    m = matrix.from_numpy(a)
    lib().write_matrix(m, os.fsencode(path))


def update_div(W, H, X, CONVERGE_THRESH=0.0, max_iter=200, t=None, verbose=0, opt=None):
    """In-place multiplicative KL-NMF updates of W (MxK) and H (KxN) for target X (MxN).

    W and H must be float32 Fortran-ordered numpy arrays (they are updated in place, like the
    reference's matrix structs); t, if given, is a float64 array of >= 10 entries.
    Returns an nmf_result."""
    for a in (W, H):
        if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.flags.f_contiguous and a.flags.writeable):
            raise TypeError("W and H must be writable float32 column-major (order='F') numpy arrays")
    Xf = _f(X)
    mw, mh, mx = matrix.from_numpy(W), matrix.from_numpy(H), matrix.from_numpy(Xf)
    res = nmf_result()
    tp = None
    if t is not None:
        assert t.dtype == np.float64 and t.size >= 10
        tp = t.ctypes.data_as(_f64p)
    _check(lib().nmf_update_div_ex(mw, mh, mx, float(CONVERGE_THRESH), int(max_iter), tp, int(verbose),
                                   ctypes.byref(opt) if opt is not None else None, ctypes.byref(res)))
    return res


def update_div_plain(W, H, X, CONVERGE_THRESH=0.0, max_iter=200, t=None, verbose=0):
    """The reference-facing symbol itself — `void update_div(matrix W, matrix H, matrix X, float, int, double t[10],
    int verbose)`, structs by value, no options, no result record (README.md:38-54).  Semantics / engine come from the
    environment (NMF_SEMANTICS, ...), as for a C caller.  W and H are updated in place."""
    for a in (W, H):
        if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.flags.f_contiguous and a.flags.writeable):
            raise TypeError("W and H must be writable float32 column-major (order='F') numpy arrays")
    Xf = _f(X)
    mw, mh, mx = matrix.from_numpy(W), matrix.from_numpy(H), matrix.from_numpy(Xf)
    tp = t.ctypes.data_as(_f64p) if t is not None else None
    L = lib()
    L.update_div(mw, mh, mx, float(CONVERGE_THRESH), int(max_iter), tp, int(verbose))   # void: in error mode 1 it returns early
    err = L.nmf_last_error()
    if err:
        raise NmfError(err.decode(errors="replace"))


def update_div_batch(Ws, Hs, X, CONVERGE_THRESH=0.0, max_iter=200, t=None, verbose=0, opt=None):
    """R restarts (lists of R float32 Fortran-ordered W (MxK) and H (KxN) arrays, updated in place) on the same X in the
    same kernel launches; returns (best index, iterations[R], converged[R], cost[R])."""
    R = len(Ws)
    assert R == len(Hs) and R >= 1
    for a in list(Ws) + list(Hs):
        if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.flags.f_contiguous and a.flags.writeable):
            raise TypeError("W[r] and H[r] must be writable float32 column-major (order='F') numpy arrays")
    Xf = _f(X)
    mws = (matrix * R)(*[matrix.from_numpy(a) for a in Ws])
    mhs = (matrix * R)(*[matrix.from_numpy(a) for a in Hs])
    its = (ctypes.c_int * R)(); conv = (ctypes.c_int * R)(); cost = (ctypes.c_double * R)()
    tp = t.ctypes.data_as(_f64p) if t is not None else None
    best = lib().nmf_update_div_batch(mws, mhs, R, matrix.from_numpy(Xf), float(CONVERGE_THRESH), int(max_iter), tp,
                                      int(verbose), ctypes.byref(opt) if opt is not None else None, its, conv, cost)
    if best < 0:
        raise NmfError(lib().nmf_last_error().decode(errors="replace"))
    return best, list(its), list(conv), list(cost)


# ------------------------------------------------------------------------------------------
# sessions: device-resident shard of one rank
# ------------------------------------------------------------------------------------------
class Session:
    def __init__(self, M, n_local, K, n_offset=0, n_global=None, opt=None):
        self.M, self.N, self.K = int(M), int(n_local), int(K)
        self.R = max(1, int(opt.restarts)) if opt is not None else 1
        self._opt = opt
        h = lib().nmf_session_create(self.M, self.N, self.K, int(n_offset),
                                     int(n_global if n_global is not None else n_local),
                                     ctypes.byref(opt) if opt is not None else None)
        if not h:
            raise NmfError(lib().nmf_last_error().decode(errors="replace"))
        self._h = ctypes.c_void_p(h)

    def close(self):
        if getattr(self, "_h", None):
            lib().nmf_session_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def upload(self, W=None, H=None, X=None):
        W = None if W is None else _f(W); H = None if H is None else _f(H); X = None if X is None else _f(X)
        _check(lib().nmf_session_upload(self._h, _p(W), _p(H), _p(X)))

    def download(self):
        W = np.empty((self.M, self.K), dtype=np.float32, order="F")
        H = np.empty((self.K, self.N), dtype=np.float32, order="F")
        _check(lib().nmf_session_download(self._h, _p(W), _p(H)))
        return W, H

    def upload_restart(self, r, W=None, H=None):
        W = None if W is None else _f(W); H = None if H is None else _f(H)
        _check(lib().nmf_session_upload_restart(self._h, int(r), _p(W), _p(H)))

    def download_restart(self, r):
        W = np.empty((self.M, self.K), dtype=np.float32, order="F")
        H = np.empty((self.K, self.N), dtype=np.float32, order="F")
        _check(lib().nmf_session_download_restart(self._h, int(r), _p(W), _p(H)))
        return W, H

    def restart_results(self):
        its = (ctypes.c_int * self.R)(); conv = (ctypes.c_int * self.R)(); cost = (ctypes.c_double * self.R)()
        _check(lib().nmf_session_restart_results(self._h, its, conv, cost))
        return list(its), list(conv), list(cost)

    def download_cols(self, cols):
        """columns `cols` (global indices inside this shard) of the resident X (M x len) and H (K x len)"""
        c = np.ascontiguousarray(cols, dtype=np.int64)
        Xc = np.empty((self.M, c.size), dtype=np.float32, order="F")
        Hc = np.empty((self.K, c.size), dtype=np.float32, order="F")
        _check(lib().nmf_session_download_cols(self._h, c.ctypes.data_as(ctypes.POINTER(ctypes.c_longlong)), int(c.size), _p(Xc), _p(Hc)))
        return Xc, Hc

    def download_rows(self, rows):
        """rows `rows` of the resident X (len x N) and W (len x K)"""
        r = np.ascontiguousarray(rows, dtype=np.int32)
        Xr = np.empty((r.size, self.N), dtype=np.float32, order="F")
        Wr = np.empty((r.size, self.K), dtype=np.float32, order="F")
        _check(lib().nmf_session_download_rows(self._h, r.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), int(r.size), _p(Xr), _p(Wr)))
        return Xr, Wr

    def half_step(self, which):
        """which = "h" | "w": one half of an iteration on the resident data"""
        _check(lib().nmf_session_half_step(self._h, 0 if which in (0, "h", "H") else 1))

    def fill_uniform(self, seed, which=7):
        _check(lib().nmf_session_fill_uniform(self._h, int(seed), int(which)))

    def fused_steps(self):
        """bit 0: W step fused, bit 1: H step fused (one kernel per half-iteration, the quotient never in HBM)"""
        return int(lib().nmf_session_fused_steps(self._h))

    def corr_bf16(self):
        """True when the fused kernels run the correction passes of the 3xTF32 products as BF16 MMAs (K = 256 by default)"""
        return bool(lib().nmf_session_corr_bf16(self._h))

    def fill_planted(self, seed, rank, noise=0.05):
        """X <- planted rank-`rank` product with +-noise multiplicative uniform noise (W, H untouched)"""
        _check(lib().nmf_session_fill_planted(self._h, int(seed), int(rank), float(noise)))

    def join(self, rank, nranks, unique_id):
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        _check(lib().nmf_session_join(self._h, int(rank), int(nranks), buf))

    def share_comm(self, other):
        """reuse `other`'s NCCL communicator (same process, same device) instead of a new ncclCommInitRank"""
        _check(lib().nmf_session_share_comm(self._h, other._h))

    def run(self, CONVERGE_THRESH=0.0, max_iter=1, t=None, verbose=0):
        res = nmf_result()
        tp = t.ctypes.data_as(_f64p) if t is not None else None
        _check(lib().nmf_session_run(self._h, float(CONVERGE_THRESH), int(max_iter), tp, int(verbose), ctypes.byref(res)))
        return res

    def cost(self):
        """KL divergence of the resident factors (a list of R values for a session with R > 1 restarts)"""
        c = (ctypes.c_double * self.R)()
        _check(lib().nmf_session_cost(self._h, c))
        return c[0] if self.R == 1 else list(c)


def nccl_unique_id():
    buf = ctypes.create_string_buffer(128)
    _check(lib().nmf_nccl_unique_id(buf))
    return buf.raw


# ------------------------------------------------------------------------------------------
# kernel-level entry points (parity tests)
# ------------------------------------------------------------------------------------------
def _o(opt):
    return ctypes.byref(opt) if opt is not None else None


def dbg_quotient(W, H, X, opt=None, with_cost=True):
    W, H, X = _f(W), _f(H), _f(X)
    M, K = W.shape; N = H.shape[1]
    Y = np.empty((M, N), dtype=np.float32, order="F")
    c = ctypes.c_double(float("nan"))
    _check(lib().nmf_dbg_quotient(_p(W), _p(H), _p(X), _p(Y), ctypes.byref(c) if with_cost else None, M, N, K, _o(opt)))
    return Y, c.value


def dbg_h_step(W, H, X, opt=None):
    W, H, X = _f(W), _f(H).copy(order="F"), _f(X)
    M, K = W.shape; N = H.shape[1]
    _check(lib().nmf_dbg_h_step(_p(W), _p(H), _p(X), M, N, K, _o(opt)))
    return H


def dbg_w_step(W, H, X, opt=None):
    W, H, X = _f(W).copy(order="F"), _f(H), _f(X)
    M, K = W.shape; N = H.shape[1]
    _check(lib().nmf_dbg_w_step(_p(W), _p(H), _p(X), M, N, K, _o(opt)))
    return W


def dbg_sums(W, H, opt=None):
    W, H = _f(W), _f(H)
    M, K = W.shape; N = H.shape[1]
    sw = np.empty(K, dtype=np.float32); sh = np.empty(K, dtype=np.float32)
    _check(lib().nmf_dbg_sums(_p(W), _p(H), _p(sw), _p(sh), M, N, K, _o(opt)))
    return sw, sh


def rowsum_noalloc(H_dev_ptr, K, N, sumH_dev_ptr, stream=None):
    """nmf_rowsum_noalloc on raw DEVICE addresses (ints): row sums of a K x N column-major matrix, nothing allocated"""
    _check(lib().nmf_rowsum_noalloc(ctypes.c_void_p(H_dev_ptr), int(K), int(N), ctypes.c_void_p(sumH_dev_ptr),
                                    ctypes.c_void_p(stream) if stream else None))


def dbg_gemm(which, A, B, M, N, K, passes=3, opt=None):
    A, B = _f(A), _f(B)
    shape = {0: (M, N), 1: (K, N), 2: (M, K)}[which]
    C = np.empty(shape, dtype=np.float32, order="F")
    _check(lib().nmf_dbg_gemm(int(which), _p(A), _p(B), _p(C), M, N, K, int(passes), _o(opt)))
    return C
