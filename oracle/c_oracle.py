"""ctypes binding of oracle/logmodel_oracle.c -- TEST INFRASTRUCTURE ONLY (see that file)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "liblogmodel_oracle.so")
_lib = None

_dp = ctypes.POINTER(ctypes.c_double)


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            build()
        _lib = ctypes.CDLL(_PATH)
        _lib.orc_logmodel_solve.restype = ctypes.c_int
        _lib.orc_logmodel_solve.argtypes = [_dp, ctypes.c_int64, ctypes.c_double, ctypes.c_double, _dp,
                                            ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                            ctypes.c_double, ctypes.c_int, _dp, _dp]
        _lib.orc_pdist_sq.argtypes = [_dp, ctypes.c_int64, ctypes.c_int64, _dp]
        _lib.orc_sum_op.argtypes = [_dp, ctypes.c_int64, _dp]
        _lib.orc_sum_op_adj.argtypes = [_dp, ctypes.c_int64, _dp]
        _lib.orc_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(_dp)


def threads():
    return lib().orc_threads()


def set_threads(n: int) -> int:
    """Override the OpenMP thread count (torchrun exports OMP_NUM_THREADS=1 to every rank; the parity leg
    of bench.py runs the oracle on rank 0 only and may use the idle host cores).  Returns the new count."""
    lib()
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(max(1, n)))
    except OSError:
        pass
    return threads()


def pdist_sq(x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    m, n = x.shape
    out = np.empty(n * (n - 1) // 2)
    lib().orc_pdist_sq(_p(x), m, n, _p(out))
    return out


def solve(z, alpha=1.0, beta=1.0, edge_init=None, maxit=1000, tol=1e-5, step_size=0.5,
          edge_tol=1e-3, threshold=True):
    """Returns (w_packed, n_iter, rel_primal, rel_dual)."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    n = int(round((1 + np.sqrt(1 + 8 * z.size)) / 2))
    w = np.zeros_like(z) if edge_init is None else np.array(edge_init, dtype=np.float64, order="C")
    rp, rd = ctypes.c_double(), ctypes.c_double()
    it = lib().orc_logmodel_solve(_p(z), n, alpha, beta, _p(w), maxit, tol, step_size, edge_tol,
                                  int(threshold), ctypes.byref(rp), ctypes.byref(rd))
    return w, it, rp.value, rd.value
