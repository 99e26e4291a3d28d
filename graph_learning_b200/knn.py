"""kNN-restricted sparse-support LogModel (BASELINE.json config 5; Kalofolias & Perraudin 2019).

The reference has no code for this variant: it is the same
forward-backward-forward iteration as ``gsp_learn_graph_log_degrees`` with the weight vector
supported on a given edge list.  Parity is pinned against ``oracle/logmodel_oracle.py::solve_knn``
and, with the complete edge list, against the dense solver.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _native as nat
from .smooth_learning import SolveInfo, _gl_dtype, _ptr, _stream, _to_device, _torch_dtype

__all__ = ["gsp_learn_graph_log_degrees_knn"]


def gsp_learn_graph_log_degrees_knn(ei, ej, distances, n_nodes: int, alpha: float = 1.0, beta: float = 1.0,
                                    edge_init=None, maxit: int = 1000, tol: float = 1e-5, step_size: float = 0.5,
                                    edge_tol: float = 1e-3, *, dtype=None, return_info: bool = False):
    """Edge weights on the support {(ei[k], ej[k])}, ei < ej, from the (theta-scaled) squared
    distances of those edges.  Returns an array shaped like ``distances`` (NumPy in -> NumPy out).
    Step size: gamma = step_size / (2 beta + sqrt(2 d_max)), d_max the largest support degree."""
    is_np = not isinstance(distances, torch.Tensor)
    if dtype is None and getattr(distances, "dtype", None) in (np.float32, torch.float32):
        dtype = torch.float32
    dt = _torch_dtype(dtype)
    z = _to_device(distances, dt)
    if z.ndim != 1:
        raise ValueError("Distances must be a vector over the edge list")
    dev = z.device
    i32 = torch.as_tensor(np.asarray(ei) if is_np else ei, device=dev).to(torch.int32).contiguous()
    j32 = torch.as_tensor(np.asarray(ej) if is_np else ej, device=dev).to(torch.int32).contiguous()
    if i32.numel() != z.numel() or j32.numel() != z.numel():
        raise ValueError("edge list and distances must have the same length")
    w = torch.zeros_like(z) if edge_init is None else _to_device(edge_init, dt).clone()
    it = ctypes.c_int()
    rp, rd = ctypes.c_double(), ctypes.c_double()
    nat.check(nat.lib().gl_knn_logmodel_solve(_ptr(i32), _ptr(j32), _ptr(z), z.numel(), int(n_nodes), _gl_dtype(dt),
                                              float(alpha), float(beta), int(maxit), float(tol), float(step_size),
                                              float(edge_tol), _ptr(w), ctypes.byref(it), ctypes.byref(rp),
                                              ctypes.byref(rd), _stream()))
    info = SolveInfo(it.value, rp.value, rd.value, bool(rp.value < tol and rd.value < tol))
    out = w.cpu().numpy() if is_np else w
    return (out, info) if return_info else out
