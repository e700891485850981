"""ctypes loader for the CPU oracle (oracle/nmf_oracle.cpp).

TEST INFRASTRUCTURE.  Import only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under nmf_gpu_b200/ may import this.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libnmf_oracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)


def build(force=False):
    src = os.path.join(_HERE, "nmf_oracle.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    return _LIB


def lib():
    global _lib
    if _lib is None:
        build()
        try:
            _lib = ctypes.CDLL(_LIB)
        except OSError:
            build(force=True)
            _lib = ctypes.CDLL(_LIB)
        L = _lib
        L.nmf_oracle_threads.restype = ctypes.c_int
        L.nmf_oracle_set_threads.restype = None
        L.nmf_oracle_set_threads.argtypes = [ctypes.c_int]
        L.nmf_oracle_eps.restype = ctypes.c_float
        L.nmf_oracle_update_div.restype = ctypes.c_int
        L.nmf_oracle_update_div.argtypes = [_f32p, _f32p, _f32p] + [ctypes.c_int] * 3 + [
            ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.POINTER(ctypes.c_double)]
        L.nmf_oracle_kl.restype = ctypes.c_double
        L.nmf_oracle_kl.argtypes = [_f32p, _f32p, _f32p] + [ctypes.c_int] * 3
        L.nmf_oracle_quotient.restype = None
        L.nmf_oracle_quotient.argtypes = [_f32p, _f32p, _f32p, _f32p] + [ctypes.c_int] * 4
        L.nmf_oracle_h_step.restype = None
        L.nmf_oracle_h_step.argtypes = [_f32p, _f32p, _f32p] + [ctypes.c_int] * 5
        L.nmf_oracle_w_step.restype = None
        L.nmf_oracle_w_step.argtypes = [_f32p, _f32p, _f32p] + [ctypes.c_int] * 5
        L.nmf_oracle_time_iterations.restype = ctypes.c_double
        L.nmf_oracle_time_iterations.argtypes = [_f32p, _f32p, _f32p] + [ctypes.c_int] * 4
    return _lib


def _colmajor(a):
    """float32 Fortran-ordered copy (the oracle takes column-major arrays)."""
    return np.asfortranarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(_f32p)


MODES = {"exact": 0, "refcompat": 1}


def update_div(W, H, X, thresh=0.0, max_iter=200, mode="exact", acc64=False, check_every=25, verbose=0):
    """Returns (W, H, iterations, final fp64 KL cost); inputs are not modified."""
    W = _colmajor(W).copy(order="F"); H = _colmajor(H).copy(order="F"); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    assert H.shape[0] == K and X.shape == (M, N)
    cost = ctypes.c_double(0.0)
    it = lib().nmf_oracle_update_div(_p(W), _p(H), _p(X), M, N, K, float(thresh), int(max_iter),
                                     MODES[mode], int(acc64), int(check_every), int(verbose),
                                     ctypes.byref(cost))
    return W, H, it, cost.value


def kl(W, H, X):
    W = _colmajor(W); H = _colmajor(H); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    return lib().nmf_oracle_kl(_p(W), _p(H), _p(X), M, N, K)


def quotient(W, H, X, acc64=False):
    W = _colmajor(W); H = _colmajor(H); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    Y = np.empty((M, N), dtype=np.float32, order="F")
    lib().nmf_oracle_quotient(_p(W), _p(H), _p(X), _p(Y), M, N, K, int(acc64))
    return Y


def h_step(W, H, X, mode="exact", acc64=False):
    W = _colmajor(W); H = _colmajor(H).copy(order="F"); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    lib().nmf_oracle_h_step(_p(W), _p(H), _p(X), M, N, K, MODES[mode], int(acc64))
    return H


def w_step(W, H, X, mode="exact", acc64=False):
    W = _colmajor(W).copy(order="F"); H = _colmajor(H); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    lib().nmf_oracle_w_step(_p(W), _p(H), _p(X), M, N, K, MODES[mode], int(acc64))
    return W


def time_iterations(W, H, X, iters):
    """Seconds for `iters` exact/fp32 iterations on all host threads; returns (seconds, threads)."""
    W = _colmajor(W).copy(order="F"); H = _colmajor(H).copy(order="F"); X = _colmajor(X)
    M, K = W.shape; N = H.shape[1]
    s = lib().nmf_oracle_time_iterations(_p(W), _p(H), _p(X), M, N, K, int(iters))
    return s, lib().nmf_oracle_threads()


def threads():
    return lib().nmf_oracle_threads()


def set_threads(n):
    """OpenMP threads for every later call (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    lib().nmf_oracle_set_threads(int(n))
