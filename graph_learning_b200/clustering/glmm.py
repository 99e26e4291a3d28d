"""Graph Laplacian Mixture Model -- drop-in for the reference ``graph_learn/clustering/glmm.py``.

The M-step (reference ``_estimate_gauss_laplacian_parameters``, :18-85) is the batched caller of
the LogModel hot path: K responsibility-weighted, mean-centred distance
vectors, K thetas, ONE batched (K, E) solve with a shared stopping decision, then the Laplacians.
Here the distances come from ``gl_pdist_sq`` with ``row_scale = resp[:, k]`` and
``col_shift = means[k]`` (never materialising the reference's (m, K, n, n) temporary, :67-68), the
solve from the fused CUDA iteration, and the Laplacians from ``gl_laplacian_from_packed``.

The EM driver is scikit-learn's private ``BaseMixture`` exactly as in the reference (:88); the
hooks accept the ``xp=`` keyword scikit-learn >= 1.8 passes.  The E-step (:238-279; SURVEY.md section
8(f) N1) keeps the reference's ``eigh`` (LAPACK for small n, cuSOLVER through torch otherwise) and
replaces the (m, n) x (n, n-1) projection + weighted square sum by the sparse quadratic-form kernel
(``assign.py``): sum_{j>=1} (lam_j + r)(v_j . y)^2 = y^T (L + r I) y - (lam_0 + r)(v_0 . y)^2.  The signal
matrix is uploaded once per ``fit`` and reused by every E- and M-step.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
from sklearn.mixture._base import BaseMixture

from ..operators import laplacian_squareform_vec, square_to_vec
from ..smooth_learning import (_theta_packed, _to_device, _torch_dtype, get_theta,
                               gsp_learn_graph_log_degrees, pdist_sq)

EIGH_DEVICE_MIN_NODES = 256   # below this LAPACK on the host is faster than the cuSOLVER launch chain


def weighted_sq_dists(x, resp, means, dtype=None) -> torch.Tensor:
    """(K, E) packed squared distances of y_k = resp[:, k, None] * (x - means[k]) between node
    columns: the quantity of reference :67-68, one ``gl_pdist_sq`` launch per component."""
    dt = _torch_dtype(dtype)
    xd = _to_device(x, dt)
    rd = _to_device(resp, dt).t().contiguous()      # (K, m)
    md = _to_device(means, dt).contiguous()         # (K, n)
    return torch.stack([pdist_sq(xd, row_scale=rd[k], col_shift=md[k], dtype=dt) for k in range(rd.shape[0])])


def _estimate_gauss_laplacian_parameters(
    x,
    resp,
    delta: float,
    *,
    theta=None,
    avg_degree=None,
    blocks=None,
    laplacians=None,
    dtype="float64",
    return_device: bool = False,
):
    """Weights, means and Laplacians of every component given the associations.

    Same contract as the reference (:18-85): exactly one of ``theta`` / ``avg_degree``;
    ``x`` (n_samples, n_nodes), ``resp`` (n_samples, n_components).  Returns NumPy arrays.
    """
    if (theta is None) == (avg_degree is None):
        raise ValueError("Exactly one of theta and avg_degree should be provided")
    resp = np.asarray(resp, dtype=np.float64)
    x_dev = x if isinstance(x, torch.Tensor) else None      # already resident (GLMM keeps it for the whole fit)
    if x_dev is None:
        x = np.asarray(x, dtype=np.float64)
    _n_samples, n_nodes = x.shape
    n_components = resp.shape[1]

    edge_init = None
    if laplacians is not None:
        # the solver expects the edge axis first (:58-59)
        edge_init = -square_to_vec(np.asarray(laplacians)).T

    weights = resp.sum(axis=0)                       # (K,)
    if x_dev is None:
        means = resp.T @ x / weights[:, None]        # (K, n)                      :62
    else:
        rt = _to_device(resp, x_dev.dtype).t()
        means = ((rt @ x_dev).double().cpu().numpy()) / weights[:, None]
    weights = weights / n_nodes                      #                             :63

    sq_dist = weighted_sq_dists(x, resp, means, dtype)   # (K, E) on the device    :67-68

    if avg_degree is not None:
        if blocks is None:
            theta_inv = torch.tensor([[_theta_packed(sq_dist[k], n_nodes, avg_degree)] for k in range(n_components)],
                                     dtype=sq_dist.dtype, device=sq_dist.device)
        else:
            iu = torch.triu_indices(n_nodes, n_nodes, 1, device=sq_dist.device)
            rows = []
            for k in range(n_components):
                sq = torch.zeros((n_nodes, n_nodes), dtype=sq_dist.dtype, device=sq_dist.device)
                sq[iu[0], iu[1]] = sq_dist[k]
                sq = sq + sq.t()
                rows.append(get_theta(sq, avg_degree, blocks=blocks))
            theta_inv = torch.stack(rows)            # (K, E) per-edge thetas        :73
    else:
        theta_inv = _to_device(np.asarray(1.0 / np.asarray(theta, dtype=np.float64)), sq_dist.dtype)

    if theta_inv.numel() in (1, n_components):
        # scalar / per-component theta rides along as z_scale (multiplied into Z while it is tiled)
        sc = theta_inv.double().reshape(-1).cpu().numpy()
        edge_weights = gsp_learn_graph_log_degrees(sq_dist, alpha=1, beta=1, edge_init=edge_init, dtype=dtype,
                                                   z_scale=np.broadcast_to(sc, (n_components,)))      # :79
    else:   # per-edge thetas of the block form (smooth_learning.py:220-235): an elementwise product
        edge_weights = gsp_learn_graph_log_degrees(sq_dist * theta_inv, alpha=1, beta=1, edge_init=edge_init,
                                                   dtype=dtype)                      # :79
    lapl = laplacian_squareform_vec(delta * edge_weights)                           # :83
    lapl_host = lapl.cpu().numpy().astype(np.float64)
    if return_device:
        return weights, means, lapl_host, lapl
    return weights, means, lapl_host


def _random_laplacians(n_nodes: int, n_components: int, random_state):
    """Laplacians with uniform weights in [0, 1], one ``default_rng(random_state)`` per component as the
    reference does (glmm.py:202-207 -> sampling/graphs.py:10-14)."""
    e = n_nodes * (n_nodes - 1) // 2
    return laplacian_squareform_vec(np.stack(
        [np.random.default_rng(random_state).uniform(size=e) for _ in range(n_components)]))


class GLMM(BaseMixture):
    """Graph Laplacian Mixture Model (Maretic & Frossard 2020); same constructor, attributes
    (``weights_``, ``means_``, ``laplacians_``, ``n_nodes_``) and ``fit`` / ``fit_predict`` /
    ``predict`` behaviour as the reference class (:88-301).  ``dtype`` is an extra keyword."""

    def __init__(
        self,
        n_components=1,
        avg_degree=2,
        *,
        tol=1e-3,
        reg_covar=1e-6,
        max_iter=100,
        n_init=1,
        init_params="kmeans",
        regul: float = 0.15,
        theta=None,
        delta: float = 2,
        laplacian_init: Optional[float | str] = None,
        blocks=None,
        random_state=None,
        warm_start=False,
        verbose=0,
        verbose_interval=10,
        dtype="float64",
    ):
        super().__init__(
            n_components=n_components, tol=tol, reg_covar=reg_covar, max_iter=max_iter, n_init=n_init,
            init_params=init_params, random_state=random_state, warm_start=warm_start, verbose=verbose,
            verbose_interval=verbose_interval,
        )
        self.regul = regul
        self.theta = theta
        self.avg_degree = avg_degree
        self.delta = delta
        self.blocks = blocks
        self.laplacian_init = laplacian_init
        self.dtype = dtype
        self._propagate_laplacians = False

    # -- BaseMixture hooks (xp= is passed by scikit-learn >= 1.8) ---------------------------------
    def _check_parameters(self, X, xp=None):
        pass

    def _mstep(self, x, resp):
        weights, means, lapl, lapl_dev = _estimate_gauss_laplacian_parameters(
            x, resp, self.delta, theta=self.theta, avg_degree=self.avg_degree, blocks=self.blocks,
            laplacians=self.laplacians_ if self._propagate_laplacians else None, dtype=self.dtype,
            return_device=True,
        )
        self._lap_cache = (lapl, lapl_dev)      # device twin of the array about to become laplacians_
        return weights, means, lapl

    def _initialize(self, x, resp, xp=None):
        _n_samples, self.n_nodes_ = x.shape
        x = self._resident(x)
        init = self.laplacian_init
        if init is None:
            self.laplacians_ = None
        elif isinstance(init, np.ndarray):
            if init.shape != (self.n_components, self.n_nodes_, self.n_nodes_):
                raise ValueError("Laplacians must have shape (n_components, n_nodes, n_nodes)")
            self._propagate_laplacians = True
            self.laplacians_ = init
            raise NotImplementedError("Laplacian order should be related to resp to make sense")
        elif isinstance(init, float):
            n = self.n_nodes_
            self.laplacians_ = np.tile(init * np.eye(n) - init / n * np.ones((n, n)), (self.n_components, 1, 1))
        elif init == "random":
            # the reference's sampler semantics (glmm.py:202-207 -> sampling/graphs.py:10-14): one
            # default_rng(random_state) PER component -- an integer seed therefore gives every component the
            # same Laplacian, a Generator is advanced from one component to the next
            self.laplacians_ = _random_laplacians(self.n_nodes_, self.n_components, self.random_state)
        else:
            raise ValueError("Invalid Laplacian init")
        self.weights_, self.means_, self.laplacians_ = self._mstep(x, resp)

    def _m_step(self, x, log_resp, xp=None):
        self.weights_, self.means_, self.laplacians_ = self._mstep(self._resident(x), np.exp(log_resp))
        self.weights_ /= self.weights_.sum()

    def _resident(self, x) -> torch.Tensor:
        """Device copy of the signal matrix, uploaded once per ``fit`` (the cache holds a reference to
        the host array, so the identity test cannot be fooled by a recycled address; do not mutate
        ``x`` in place between calls)."""
        if isinstance(x, torch.Tensor):
            return x
        cache = getattr(self, "_x_cache", None)
        if cache is not None and cache[0] is x:
            return cache[1]
        xd = _to_device(x, _torch_dtype(self.dtype))
        if getattr(self, "_fitting", False):
            self._x_cache = (x, xd)
        return xd

    def fit_predict(self, X, y=None):
        self._fitting = True
        try:
            return super().fit_predict(X, y)
        finally:
            self._fitting = False
            self._x_cache = None
            self._lap_cache = None

    def _estimate_log_prob(self, x, xp=None):
        """Assignment log-probabilities under N(mean_k, pinv(L_k + regul)) restricted to the
        non-constant eigenvectors (reference :238-279)."""
        from .assign import QuadForm
        xd = self._resident(x)
        lap = np.asarray(self.laplacians_, dtype=np.float64)
        cache = getattr(self, "_lap_cache", None)
        lap_dev = cache[1] if cache is not None and cache[0] is self.laplacians_ else _to_device(lap, xd.dtype)
        if self.n_nodes_ < EIGH_DEVICE_MIN_NODES:
            eigval, eigvec = np.linalg.eigh(lap)                                     # :253
            first = np.ascontiguousarray(eigvec[:, :, 0])
        else:
            ev, evec = torch.linalg.eigh(lap_dev.double())
            eigval, first = ev.cpu().numpy(), evec[:, :, 0].contiguous()
        eigval = eigval + self.regul                                                 # :254
        with QuadForm(lap_dev, diag_add=self.regul, dtype=xd.dtype) as qf:
            quad, dot = qf.apply(xd, shift=self.means_, vectors=first)
        lam0 = torch.as_tensor(eigval[:, 0], device=quad.device, dtype=torch.float64)
        data = quad.double() - lam0[None, :] * dot.double() ** 2                      # :267-274
        # :272 is log(prod(.)); the sum of logs is the same number wherever that product is finite and
        # stays finite where it overflows (n >~ 400), which config 4 (n = 1,000) needs
        with np.errstate(divide="ignore", invalid="ignore"):
            logdet = np.sum(np.log(eigval[:, 1:]), axis=1)
        const = -(self.n_nodes_ - 1) * np.log(2 * np.pi)
        return (const + logdet[None, :] - data.cpu().numpy()) / 2

    def _estimate_log_weights(self, xp=None):
        return np.log(self.weights_)

    def _get_parameters(self):
        return self.weights_, self.means_, self.laplacians_

    def _set_parameters(self, params, xp=None):
        self.weights_, self.means_, self.laplacians_ = params
        self.n_nodes_ = self.means_.shape[-1]

    def _compute_lower_bound(self, _, log_prob_norm):
        return log_prob_norm
