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
from sklearn.base import BaseEstimator

from .smooth_learning import SolveInfo, _gl_dtype, _ptr, _stream, _theta_dense, _to_device, _torch_dtype

__all__ = ["gsp_learn_graph_log_degrees_knn", "knn_search", "knn_graph", "theta_from_knn", "locality_keys",
           "KNNLogModel"]


def _part1by2(v: torch.Tensor) -> torch.Tensor:
    """Spread the low 10 bits of v (int64) so that two zero bits follow each of them."""
    v = v & 0x3FF
    v = (v | (v << 16)) & 0x030000FF
    v = (v | (v << 8)) & 0x0300F00F
    v = (v | (v << 4)) & 0x030C30C3
    v = (v | (v << 2)) & 0x09249249
    return v


def locality_keys(x) -> torch.Tensor:
    """Morton (Z-order) codes of the nodes (COLUMNS of x, shape (m, n)) as non-negative int32, for the
    ``order_key`` of :func:`gsp_learn_graph_log_degrees_knn`: nodes with close codes are close in
    space, so rows that share a slice of the solver's layout have overlapping neighbourhoods.  The
    three features of largest variance are quantised to 10 bits each (all of them when m <= 3); for
    high-dimensional signals this is only a coarse arrangement.  Plain torch ops on x's device; the
    codes steer an internal row order and never change a result."""
    xt = x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x))
    if xt.ndim != 2:
        raise ValueError("signals must be a (n_samples, n_nodes) matrix")
    xt = xt.to(torch.float64)
    m, n = xt.shape
    if m > 3:
        top = torch.argsort(xt.var(dim=1, unbiased=False), descending=True, stable=True)[:3]
        xt = xt[torch.sort(top).values]
    lo = xt.min(dim=1, keepdim=True).values
    span = (xt.max(dim=1, keepdim=True).values - lo).clamp_min(torch.finfo(torch.float64).tiny)
    q = torch.clamp(((xt - lo) / span * 1024.0).to(torch.int64), 0, 1023)
    key = torch.zeros(n, dtype=torch.int64, device=xt.device)
    for a in range(q.shape[0]):
        key |= _part1by2(q[a]) << a
    return key.to(torch.int32)


def knn_search(x, k: int):
    """k nearest neighbours between the COLUMNS of x (m, n) by exact brute force on the device
    (``gl_knn_build``).  Returns CUDA tensors ``idx`` (n, k) int32 and ``d2`` (n, k) float32 of
    squared distances, ascending per row, self excluded."""
    xd = _to_device(x, torch.float32)
    if xd.ndim != 2:
        raise ValueError("signals must be a (n_samples, n_nodes) matrix")
    m, n = xd.shape
    idx = torch.empty((n, k), dtype=torch.int32, device=xd.device)
    d2 = torch.empty((n, k), dtype=torch.float32, device=xd.device)
    nat.check(nat.lib().gl_knn_build(_ptr(xd), m, n, xd.stride(0), int(k), _ptr(idx), _ptr(d2), _stream()))
    return idx, d2


def knn_graph(x, k: int):
    """Symmetrised kNN support: the undirected edges {i, j} with j among the k nearest of i or i
    among those of j, as ``(ei, ej, d2)`` with ei < ej, sorted by (ei, ej), plus the raw search
    result.  The edge assembly (canonical order, de-duplication) is torch glue on the device."""
    idx, d2 = knn_search(x, k)
    n = idx.shape[0]
    rows = torch.arange(n, device=idx.device, dtype=torch.int64)[:, None].expand(-1, k)
    cols = idx.to(torch.int64)
    ok = cols >= 0
    lo = torch.minimum(rows, cols)[ok]
    hi = torch.maximum(rows, cols)[ok]
    key, first = _unique_first(lo * n + hi)
    ei = torch.div(key, n, rounding_mode="floor").to(torch.int32)
    ej = (key % n).to(torch.int32)
    return ei, ej, d2[ok][first], (idx, d2)


def _unique_first(key: torch.Tensor):
    """Sorted unique keys and, for each, the position of one occurrence (d(i,j) == d(j,i) bit for
    bit, so any occurrence carries the same distance)."""
    order = torch.argsort(key)
    ks = key[order]
    keep = torch.ones_like(ks, dtype=torch.bool)
    keep[1:] = ks[1:] != ks[:-1]
    return ks[keep], order[keep]


def theta_from_knn(d2_sorted: torch.Tensor, avg_degree: int) -> float:
    """The reference's theta heuristic (smooth_learning.py:201-219) from kNN lists: the statistic only
    reads the avg_degree + 1 smallest entries of every row of the square distance matrix (the zero
    diagonal included), which the lists hold exactly when k >= avg_degree."""
    n, k = d2_sorted.shape
    if k < avg_degree:
        raise ValueError(f"need k >= avg_degree neighbours per node (k={k}, avg_degree={avg_degree})")
    mat = torch.cat([torch.zeros((n, 1), dtype=d2_sorted.dtype, device=d2_sorted.device), d2_sorted], dim=1)
    return _theta_dense(mat.contiguous(), int(avg_degree))


def gsp_learn_graph_log_degrees_knn(ei, ej, distances, n_nodes: int, alpha: float = 1.0, beta: float = 1.0,
                                    edge_init=None, maxit: int = 1000, tol: float = 1e-5, step_size: float = 0.5,
                                    edge_tol: float = 1e-3, *, dtype=None, return_info: bool = False,
                                    order_key=None, sort_window: int = 0):
    """Edge weights on the support {(ei[k], ej[k])}, ei < ej, from the (theta-scaled) squared
    distances of those edges.  Returns an array shaped like ``distances`` (NumPy in -> NumPy out).
    Step size: gamma = step_size / (2 beta + sqrt(2 d_max)), d_max the largest support degree.

    ``order_key`` ((n_nodes,) non-negative int32, e.g. :func:`locality_keys`) and ``sort_window``
    choose the solver's INTERNAL row order (``gl_knn_create_ordered`` in include/gl_b200.h); without a
    key the rows are sorted by decreasing degree over the caller's numbering.  With Morton keys the far
    ends of a CTA's rows fall into a few KB of the node array (0.085 against 0.127 ms per iteration at
    n = 1e6 on B200).  The order changes memory traffic and locality, never the result beyond the
    rounding of the degree sums' summation order."""
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
    if order_key is None and int(sort_window) == 0:
        nat.check(nat.lib().gl_knn_logmodel_solve(_ptr(i32), _ptr(j32), _ptr(z), z.numel(), int(n_nodes),
                                                  _gl_dtype(dt), float(alpha), float(beta), int(maxit), float(tol),
                                                  float(step_size), float(edge_tol), _ptr(w), ctypes.byref(it),
                                                  ctypes.byref(rp), ctypes.byref(rd), _stream()))
    else:
        key = None
        if order_key is not None:
            key = torch.as_tensor(order_key, device=dev).to(torch.int32).contiguous()
            if key.ndim != 1 or key.numel() != int(n_nodes):
                raise ValueError("order_key must hold one key per node")
            if key.numel() and int(key.min()) < 0:
                raise ValueError("order_key must be non-negative")
        nat.check(nat.lib().gl_knn_logmodel_solve_ordered(
            _ptr(i32), _ptr(j32), _ptr(z), z.numel(), int(n_nodes), _gl_dtype(dt), float(alpha), float(beta),
            int(maxit), float(tol), float(step_size), float(edge_tol), _ptr(w), ctypes.byref(it), ctypes.byref(rp),
            ctypes.byref(rd), _ptr(key) if key is not None else None, int(sort_window), 0, _stream()))
    info = SolveInfo(it.value, rp.value, rd.value, bool(rp.value < tol and rd.value < tol))
    out = w.cpu().numpy() if is_np else w
    return (out, info) if return_info else out


class KNNLogModel(BaseEstimator):
    """LogModel on a kNN-restricted support (BASELINE config 5; Kalofolias & Perraudin 2019): kNN
    search, theta for the requested degree from the kNN lists, and the sparse-support solve, all on
    the device.  No reference estimator exists for it; the parameters mirror ``LogModel``.

    Attributes: ``edges_`` (2, nnz) int32 with ei < ej, ``weights_`` (nnz,), ``theta_``,
    ``n_iter_``; :meth:`to_sparse` gives the symmetric scipy CSR matrix.

    ``locality_order`` (default True: Morton arrangement of the rows from the three features of largest
    variance, then a stable global degree sort) / ``sort_window`` choose the solver's internal row
    order (see :func:`gsp_learn_graph_log_degrees_knn`); they never change what is returned."""

    def __init__(self, avg_degree: int = 10, n_neighbors: int = 20, *, maxit: int = 1000, tol: float = 1e-5,
                 step_size: float = 0.5, edge_tol: float = 1e-3, locality_order: bool = True,
                 sort_window: int = 0) -> None:
        self.locality_order = locality_order   # Morton arrangement of the rows (locality_keys)
        self.sort_window = sort_window         # degree sort inside windows of this many rows (0: global)
        self.avg_degree = avg_degree
        self.n_neighbors = n_neighbors
        self.maxit = maxit
        self.tol = tol
        self.step_size = step_size
        self.edge_tol = edge_tol

    def fit(self, x):
        is_np = not isinstance(x, torch.Tensor)
        ei, ej, d2, (_, d2_lists) = knn_graph(x, self.n_neighbors)
        self.n_nodes_ = int(d2_lists.shape[0])
        self.theta_ = theta_from_knn(d2_lists, self.avg_degree)
        w, info = gsp_learn_graph_log_degrees_knn(ei, ej, d2 * self.theta_, self.n_nodes_, maxit=self.maxit,
                                                  tol=self.tol, step_size=self.step_size, edge_tol=self.edge_tol,
                                                  return_info=True,
                                                  order_key=locality_keys(x).to(d2.device) if self.locality_order
                                                  else None, sort_window=self.sort_window)
        edges = torch.stack([ei, ej])
        self.edges_ = edges.cpu().numpy() if is_np else edges
        self.weights_ = w.cpu().numpy() if is_np else w
        self.n_iter_, self.rel_primal_, self.rel_dual_, self.converged_ = (info.n_iter, info.rel_primal,
                                                                           info.rel_dual, info.converged)
        return self

    def to_sparse(self):
        from scipy import sparse
        e = np.asarray(self.edges_.cpu() if isinstance(self.edges_, torch.Tensor) else self.edges_)
        w = np.asarray(self.weights_.cpu() if isinstance(self.weights_, torch.Tensor) else self.weights_)
        a = sparse.coo_matrix((w, (e[0], e[1])), shape=(self.n_nodes_, self.n_nodes_))
        return (a + a.T).tocsr()
