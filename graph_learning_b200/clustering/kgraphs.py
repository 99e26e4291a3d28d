"""K-Graphs clustering -- drop-in for the reference ``graph_learn/clustering/kgraphs.py``.

The graph-update step (reference ``KGraphs._single_fit``, :93-108) is a batched caller of the
LogModel hot path: per-cluster squared distances (:96), per-cluster
theta (:103), ONE batched (K, E) solve (:102-106) and the Laplacians (:108), all on the device.
The assignment step (:87-88, :112-113: ``einsum('ni,kij,nj->nk')`` + ``argmin``; SURVEY.md section
8(f) N2) runs as the sparse quadratic-form kernel of ``assign.py`` / ``csrc/quadform.cu``; the signals,
labels and Laplacians stay on the device for the whole outer loop.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
from scipy.special import softmax
from sklearn.base import BaseEstimator, ClusterMixin
from sklearn.utils import check_random_state

from ..operators import laplacian_squareform_vec
from ..smooth_learning import _theta_packed, _to_device, _torch_dtype, gsp_learn_graph_log_degrees, pdist_sq
from .assign import QuadForm
from .glmm import _estimate_gauss_laplacian_parameters
from .utils import init_labels, one_hot


class KGraphs(BaseEstimator, ClusterMixin):
    """K-Graphs (Araghi et al. 2019).  Same constructor / attributes (``labels_``, ``laplacians_``,
    ``converged_``, ``score_``) / methods as the reference class (:20-149); ``dtype`` is extra."""

    def __init__(
        self,
        n_clusters=1,
        avg_degree: int = 2,
        *,
        max_iter=100,
        n_init=1,
        init_params="kmeans",
        delta: float = 1,
        random_state=None,
        dtype="float64",
    ) -> None:
        self.n_clusters = n_clusters
        self.max_iter = max_iter
        self.n_init = n_init
        self.init_params = init_params
        self.delta = delta
        self.avg_degree = avg_degree
        self.random_state = random_state
        self.dtype = dtype

    def _init_parameters(self, x):
        _n_samples, n_nodes = x.shape
        self.random_state = check_random_state(self.random_state)
        self.labels_ = init_labels(x, self.n_clusters, init_params=self.init_params,
                                   random_state=self.random_state)
        self.laplacians_ = np.empty((self.n_clusters, n_nodes, n_nodes))
        self.converged_ = False
        self.score_ = np.inf

    def _smoothness(self, x) -> np.ndarray:
        """x_n^T L_k x_n for every sample and cluster, shape (n_samples, n_clusters) (:87-88)."""
        with QuadForm(self.laplacians_, dtype=torch.float64) as qf:
            return qf.apply(x).cpu().numpy()

    def _graph_step_dev(self, x_dev: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Laplacians (K, n, n) of all clusters from the current labels, on the device (:96-108)."""
        n_nodes = x_dev.shape[1]
        sq_dist = torch.stack([pdist_sq(x_dev[labels == k], dtype=x_dev.dtype) for k in range(self.n_clusters)])
        # theta per cluster (:98-101) rides along as z_scale: multiplied into Z while it is tiled, no extra pass
        theta = [_theta_packed(sq_dist[k], n_nodes, self.avg_degree) for k in range(self.n_clusters)]
        edge_weights = gsp_learn_graph_log_degrees(sq_dist, alpha=1, beta=1, dtype=self.dtype, z_scale=theta)
        return laplacian_squareform_vec(self.delta * edge_weights)

    def _graph_step(self, x_dev: torch.Tensor) -> np.ndarray:
        labels = torch.as_tensor(self.labels_, device=x_dev.device)
        return self._graph_step_dev(x_dev, labels).cpu().numpy().astype(np.float64)

    def _single_fit(self, x, _y=None) -> None:
        self._init_parameters(x)
        x_dev = _to_device(x, _torch_dtype(self.dtype))
        labels_dev = torch.as_tensor(np.asarray(self.labels_), device=x_dev.device).to(torch.int32)
        lap_dev = None
        for _i in range(self.max_iter):
            lap_dev = self._graph_step_dev(x_dev, labels_dev)
            with QuadForm(lap_dev, dtype=x_dev.dtype) as qf:
                smooth, new_labels = qf.apply(x_dev, labels=True)          # :112-113, argmin fused
            self.score_ = float(smooth.gather(1, new_labels.long()[:, None]).sum())
            if torch.equal(new_labels, labels_dev):
                self.converged_ = True
                break
            labels_dev = new_labels
        self.labels_ = labels_dev.cpu().numpy().astype(np.int64)
        if lap_dev is not None:
            self.laplacians_ = lap_dev.cpu().numpy().astype(np.float64)

    def fit_predict(self, x, _y=None):
        best_score, best_params = np.inf, {}
        for _n in range(self.n_init):
            self._single_fit(x)
            if self.score_ < best_score:
                best_score = self.score_
                best_params = self.get_params()
        self.score_ = best_score
        self.set_params(**best_params)
        return self.labels_

    def fit(self, x, _y=None):
        self.fit_predict(x)
        return self

    def predict(self, x):
        return np.argmin(self._smoothness(x), axis=1)

    def predict_proba(self, x):
        return softmax(-self._smoothness(x), axis=1)


class KGraphsV2(KGraphs):
    """K-Graphs with cluster means: the graph step is the GLMM M-step with one-hot
    responsibilities (reference :152-233)."""

    def __init__(
        self,
        n_clusters=1,
        avg_degree: int = 2,
        *,
        max_iter=100,
        n_init=1,
        init_params="kmeans",
        delta: float = 1,
        theta: Optional[float] = None,
        random_state=None,
        dtype="float64",
    ) -> None:
        super().__init__(n_clusters, avg_degree, max_iter=max_iter, n_init=n_init, init_params=init_params,
                         delta=delta, random_state=random_state, dtype=dtype)
        self.theta = theta

    def _mstep(self, x, labels):
        # NOTE the reference passes BOTH theta and avg_degree (:221-227); with its default
        # theta=None that is the avg_degree parametrisation, otherwise it raises -- kept.
        return _estimate_gauss_laplacian_parameters(x, one_hot(labels, self.n_clusters), self.delta,
                                                    theta=self.theta, avg_degree=self.avg_degree,
                                                    dtype=self.dtype)

    def _init_parameters(self, x):
        self.random_state = check_random_state(self.random_state)
        self.labels_ = init_labels(x, self.n_clusters, init_params=self.init_params,
                                   random_state=self.random_state)
        _, self.means_, self.laplacians_ = self._mstep(x, self.labels_)
        self.converged_ = False
        self.score_ = np.inf

    def _centered_smoothness(self, x) -> np.ndarray:
        with QuadForm(self.laplacians_, dtype=torch.float64) as qf:
            return qf.apply(x, shift=self.means_).cpu().numpy()

    def _single_fit(self, x, _y=None) -> None:
        self._init_parameters(x)
        x_dev = _to_device(x, torch.float64)
        for _i in range(self.max_iter):
            with QuadForm(self.laplacians_, dtype=torch.float64) as qf:
                smooth, lab = qf.apply(x_dev, labels=True)                  # uncentred, as the reference (:214)
            self.score_ = float(smooth.gather(1, lab.long()[:, None]).sum())
            labels = lab.cpu().numpy().astype(np.int64)
            _, self.means_, self.laplacians_ = self._mstep(x, labels)
            if np.allclose(labels, self.labels_):
                self.converged_ = True
                return
            self.labels_ = labels
