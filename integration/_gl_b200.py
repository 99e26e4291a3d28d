"""Reference-side binding of libgl_b200.so -- the file a maintainer of LTS4/graph-learning would drop
in as ``graph_learn/_gl_b200.py`` (INTEGRATION.md route B).  It depends on NumPy, SciPy and ctypes
only -- no torch, nothing from graph_learning_b200 -- and replaces
``gsp_learn_graph_log_degrees`` (reference graph_learn/smooth_learning.py:47-174) with one call of
the C-ABI entry point ``gl_logmodel_solve_host`` (include/gl_b200.h).

``tests/test_gpu_boundary.py`` drives exactly this file against the goldens of the unmodified
reference, so the stub printed in INTEGRATION.md is tested code.
"""
import ctypes
import os

import numpy as np
from scipy.spatial.distance import squareform

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = ctypes.CDLL(os.environ.get(
    "GL_B200_LIB", os.path.join(_HERE, os.pardir, "graph_learning_b200", "libgl_b200.so")))
_dp, _ip = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
_lib.gl_logmodel_solve_host.restype = ctypes.c_int
_lib.gl_logmodel_solve_host.argtypes = [
    ctypes.c_void_p, _dp, ctypes.c_int, ctypes.c_int, ctypes.c_int,    # Z (K,E) host, z_scale, n, K, dtype
    ctypes.c_double, ctypes.c_double, ctypes.c_void_p,                # alpha, beta, edge_init (K,E) host | NULL
    ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double,  # maxit, tol, step_size, edge_tol
    ctypes.c_void_p, _ip, _dp, _dp]                                   # W (K,E) host out, n_iter, residuals
_lib.gl_last_error.restype = ctypes.c_char_p

GL_F32, GL_F64 = 0, 1
last_info = {}  # n_iter / rel_primal / rel_dual of the last call (the reference returns neither)


def gsp_learn_graph_log_degrees(distances, alpha, beta, edge_init=None, maxit=1000, tol=1e-5,
                                step_size=0.5, edge_tol=1e-3, dtype=np.float64):
    distances = np.asarray(distances)
    if distances.ndim > 2:
        raise ValueError("Distances must be vector or matrix")           # smooth_learning.py:130-131
    z = np.ascontiguousarray(np.atleast_2d(distances), dtype=dtype)       # (K, E)
    k, e = z.shape
    n = int(np.round((1 + np.sqrt(1 + 8 * e)) / 2))                       # :133
    # the reference's edge_init is in its INTERNAL layout, (E,) or (E, K) (:140-141, glmm.py:58-59);
    # the C-ABI takes (K, E) like the distances
    w0 = None if edge_init is None else np.ascontiguousarray(np.atleast_2d(np.asarray(edge_init).T), dtype=dtype)
    w = np.empty_like(z)
    it, rp, rd = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
    rc = _lib.gl_logmodel_solve_host(z.ctypes.data, None, n, k, GL_F64 if z.dtype == np.float64 else GL_F32,
                                     alpha, beta, None if w0 is None else w0.ctypes.data, maxit, tol,
                                     step_size, edge_tol, w.ctypes.data, ctypes.byref(it), ctypes.byref(rp),
                                     ctypes.byref(rd))
    if rc == 1:
        raise ValueError(_lib.gl_last_error().decode())
    if rc != 0:
        raise RuntimeError(_lib.gl_last_error().decode())
    last_info.update(n_iter=it.value, rel_primal=rp.value, rel_dual=rd.value)
    return squareform(w[0]) if distances.ndim == 1 else w                # :171-174
