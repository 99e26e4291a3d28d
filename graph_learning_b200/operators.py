"""Dense-Laplacian forms of the edge <-> degree operator pair -- B200-native drop-in for the part
of the reference module ``graph_learn/operators.py`` that lies on the LogModel path (:39-94).

``laplacian_squareform_vec`` is also the output step of GLMM / k-Graphs (clustering/glmm.py:83,
clustering/kgraphs.py:108).  The rest of the reference module (dictionary-model math, :97-260) is
out of scope.  NumPy in -> NumPy out, CUDA tensor in -> CUDA tensor out.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _native as nat
from .smooth_learning import _gl_dtype, _ptr, _stream, _to_device

__all__ = [
    "square_to_vec",
    "laplacian_squareform",
    "laplacian_squareform_vec",
    "laplacian_squareform_adj",
    "laplacian_squareform_adj_vec",
]


def _dt(x):
    return torch.float32 if getattr(x, "dtype", None) in (np.float32, torch.float32) else torch.float64


def _n_nodes(n_edges: int) -> int:
    return int(np.sqrt(2 * n_edges + 0.25) + 0.5)  # operators.py:60


def square_to_vec(x):
    """(K, n, n) -> (K, E): strict upper triangle, row-major (operators.py:39-41).  Pure indexing."""
    if isinstance(x, torch.Tensor):
        iu = torch.triu_indices(x.shape[-1], x.shape[-1], 1, device=x.device)
        return x[:, iu[0], iu[1]]
    triu = np.triu_indices(x.shape[-1], k=1)
    return x[:, triu[0], triu[1]]


def laplacian_squareform_vec(x):
    """(K, E) edge weights -> (K, n, n) Laplacians D - W; the diagonal is S w (operators.py:58-70)."""
    is_np = not isinstance(x, torch.Tensor)
    dt = _dt(x)
    xd = _to_device(x, dt)
    k, e = xd.shape
    n = _n_nodes(e)
    out = torch.empty((k, n, n), dtype=dt, device=xd.device)
    nat.check(nat.lib().gl_laplacian_from_packed(_ptr(xd), _ptr(out), n, k, _gl_dtype(dt), _stream()))
    return out.cpu().numpy() if is_np else out


def laplacian_squareform(x):
    """(E,) edge weights -> (n, n) Laplacian (operators.py:44-55)."""
    return laplacian_squareform_vec(x[None, :])[0]


def laplacian_squareform_adj_vec(laplacians):
    """Adjoint of :func:`laplacian_squareform_vec`: L -> L_ii + L_jj - L_ij - L_ji over i<j
    (operators.py:85-94)."""
    is_np = not isinstance(laplacians, torch.Tensor)
    dt = _dt(laplacians)
    ld = _to_device(laplacians, dt)
    k, n, _ = ld.shape
    out = torch.empty((k, n * (n - 1) // 2), dtype=dt, device=ld.device)
    nat.check(nat.lib().gl_laplacian_adjoint(_ptr(ld), _ptr(out), n, k, _gl_dtype(dt), _stream()))
    return out.cpu().numpy() if is_np else out


def laplacian_squareform_adj(laplacian):
    """Adjoint of :func:`laplacian_squareform` (operators.py:73-82)."""
    return laplacian_squareform_adj_vec(laplacian[None])[0]
