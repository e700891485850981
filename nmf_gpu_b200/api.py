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
                ("reserved", ctypes.c_int * 10)]


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
    "nmf_session_create": (ctypes.c_void_p, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_longlong,
                                             ctypes.c_longlong, ctypes.POINTER(nmf_options)]),
    "nmf_session_destroy": (None, [ctypes.c_void_p]),
    "nmf_session_upload": (ctypes.c_int, [ctypes.c_void_p, _f32p, _f32p, _f32p]),
    "nmf_session_download": (ctypes.c_int, [ctypes.c_void_p, _f32p, _f32p]),
    "nmf_session_fill_uniform": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_ulonglong, ctypes.c_int]),
    "nmf_nccl_unique_id": (ctypes.c_int, [ctypes.c_void_p]),
    "nmf_session_join": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "nmf_session_share_comm": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "nmf_session_fill_planted": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_float]),
    "nmf_session_fused_steps": (ctypes.c_int, [ctypes.c_void_p]),
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


def options(semantics="exact", engine="tcgen05", device=-1, check_every=25, use_graph=True, tf32_passes=3):
    o = nmf_options()
    lib().nmf_default_options(ctypes.byref(o))
    o.semantics = REFCOMPAT if semantics in ("refcompat", REFCOMPAT) else EXACT
    o.engine = ENGINE_SIMT if engine in ("simt", ENGINE_SIMT) else ENGINE_TCGEN05
    o.device, o.check_every, o.use_graph, o.tf32_passes = device, check_every, int(use_graph), tf32_passes
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


# ------------------------------------------------------------------------------------------
# sessions: device-resident shard of one rank
# ------------------------------------------------------------------------------------------
class Session:
    def __init__(self, M, n_local, K, n_offset=0, n_global=None, opt=None):
        self.M, self.N, self.K = int(M), int(n_local), int(K)
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

    def fill_uniform(self, seed, which=7):
        _check(lib().nmf_session_fill_uniform(self._h, int(seed), int(which)))

    def fused_steps(self):
        """bit 0: W step fused, bit 1: H step fused (one kernel per half-iteration, the quotient never in HBM)"""
        return int(lib().nmf_session_fused_steps(self._h))

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
        c = ctypes.c_double()
        _check(lib().nmf_session_cost(self._h, ctypes.byref(c)))
        return c.value


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


def dbg_gemm(which, A, B, M, N, K, passes=3, opt=None):
    A, B = _f(A), _f(B)
    shape = {0: (M, N), 1: (K, N), 2: (M, K)}[which]
    C = np.empty(shape, dtype=np.float32, order="F")
    _check(lib().nmf_dbg_gemm(int(which), _p(A), _p(B), _p(C), M, N, K, int(passes), _o(opt)))
    return C
