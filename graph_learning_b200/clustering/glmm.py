"""Graph Laplacian Mixture Model -- drop-in for the reference ``graph_learn/clustering/glmm.py``.

The M-step (reference ``_estimate_gauss_laplacian_parameters``, :18-85) is the batched caller of
the LogModel hot path: K responsibility-weighted, mean-centred distance
vectors, K thetas, ONE batched (K, E) solve with a shared stopping decision, then the Laplacians.
Here the distances come from ``gl_pdist_sq`` with ``row_scale = resp[:, k]`` and
``col_shift = means[k]`` (never materialising the reference's (m, K, n, n) temporary, :67-68), the
solve from the fused CUDA iteration, and the Laplacians from ``gl_laplacian_from_packed``.

The EM driver is scikit-learn's private ``BaseMixture`` exactly as in the reference (:88); the
hooks accept the ``xp=`` keyword scikit-learn >= 1.8 passes.  The E-step (:238-279) is host NumPy
here as in the reference -- it is a "next" row, not the hot path.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
from sklearn.mixture._base import BaseMixture

from ..operators import laplacian_squareform_vec, square_to_vec
from ..smooth_learning import (_theta_packed, _to_device, _torch_dtype, get_theta,
                               gsp_learn_graph_log_degrees, pdist_sq)


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
):
    """Weights, means and Laplacians of every component given the associations.

    Same contract as the reference (:18-85): exactly one of ``theta`` / ``avg_degree``;
    ``x`` (n_samples, n_nodes), ``resp`` (n_samples, n_components).  Returns NumPy arrays.
    """
    if (theta is None) == (avg_degree is None):
        raise ValueError("Exactly one of theta and avg_degree should be provided")
    x = np.asarray(x, dtype=np.float64)
    resp = np.asarray(resp, dtype=np.float64)
    _n_samples, n_nodes = x.shape
    n_components = resp.shape[1]

    edge_init = None
    if laplacians is not None:
        # the solver expects the edge axis first (:58-59)
        edge_init = -square_to_vec(np.asarray(laplacians)).T

    weights = resp.sum(axis=0)                       # (K,)
    means = resp.T @ x / weights[:, None]            # (K, n)                      :62
    weights = weights / n_nodes                      #                             :63

    sq_dist = weighted_sq_dists(x, resp, means, dtype)   # (K, E) on the device    :67-68

    if avg_degree is not None:
        if blocks is None:
            theta_inv = torch.tensor([[_theta_packed(sq_dist[k], n_nodes, avg_degree)] for k in range(n_components)],
                                     dtype=sq_dist.dtype, device=sq_dist.device)
        else:
            from .. import _native as nat
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

    edge_weights = gsp_learn_graph_log_degrees(sq_dist * theta_inv, alpha=1, beta=1, edge_init=edge_init,
                                               dtype=dtype)                          # :79
    lapl = laplacian_squareform_vec(delta * edge_weights)                           # :83
    return weights, means, lapl.cpu().numpy().astype(np.float64)


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
        return _estimate_gauss_laplacian_parameters(
            x, resp, self.delta, theta=self.theta, avg_degree=self.avg_degree, blocks=self.blocks,
            laplacians=self.laplacians_ if self._propagate_laplacians else None, dtype=self.dtype,
        )

    def _initialize(self, x, resp, xp=None):
        _n_samples, self.n_nodes_ = x.shape
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
            rng = np.random.default_rng(self.random_state if isinstance(self.random_state, (int, np.integer)) else None)
            e = self.n_nodes_ * (self.n_nodes_ - 1) // 2
            self.laplacians_ = laplacian_squareform_vec(rng.uniform(size=(self.n_components, e)))
        else:
            raise ValueError("Invalid Laplacian init")
        self.weights_, self.means_, self.laplacians_ = self._mstep(x, resp)

    def _m_step(self, x, log_resp, xp=None):
        self.weights_, self.means_, self.laplacians_ = self._mstep(x, np.exp(log_resp))
        self.weights_ /= self.weights_.sum()

    def _estimate_log_prob(self, x, xp=None):
        """Assignment log-probabilities under N(mean_k, pinv(L_k + regul)) restricted to the
        non-constant eigenvectors (reference :238-279)."""
        x = np.asarray(x)
        eigval, eigvec = np.linalg.eigh(self.laplacians_)
        eigval = eigval + self.regul
        log_prob = np.empty((x.shape[0], self.n_components))
        for k in range(self.n_components):
            proj = (x - self.means_[k]) @ eigvec[k, :, 1:]
            log_prob[:, k] = (
                -(self.n_nodes_ - 1) * np.log(2 * np.pi)
                + np.log(np.prod(eigval[k, 1:]))
                - np.sum(proj * proj * eigval[k, 1:], axis=-1)
            ) / 2
        return log_prob

    def _estimate_log_weights(self, xp=None):
        return np.log(self.weights_)

    def _get_parameters(self):
        return self.weights_, self.means_, self.laplacians_

    def _set_parameters(self, params, xp=None):
        self.weights_, self.means_, self.laplacians_ = params
        self.n_nodes_ = self.means_.shape[-1]

    def _compute_lower_bound(self, _, log_prob_norm):
        return log_prob_norm
