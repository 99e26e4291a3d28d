"""Windowed LogModel -- drop-in for the reference ``graph_learn/temporal/smooth.py``.

``WindowLogModel`` splits the samples into non-overlapping windows, builds one squared-distance
vector and one theta per window (reference ``_initialize``, :41-61) and hands the (n_windows, E)
batch to the solver's 2-D path through the inherited ``LogModel.fit``.  Here each window's
distances are one ``gl_pdist_sq`` launch over a row slice of x (zero-padding the last window
changes nothing: zero rows add zero to every distance, :44-45).
"""
from __future__ import annotations

import numpy as np
import torch

from ..smooth_learning import LogModel, _theta_packed, _to_device, _torch_dtype, pdist_sq


class WindowLogModel(LogModel):
    """Learn a sequence of graphs from signals by applying LogModel on windows.

    ``theta_`` is an (n_windows, 1) array (or 1 when ``avg_degree`` is None) and ``weights_`` is
    (n_windows, E) packed, as in the reference (:12-61)."""

    def __init__(
        self,
        window_size: int = 10,
        avg_degree: int = None,
        *,
        edge_init=None,
        maxit: int = 1000,
        tol: float = 1e-5,
        step_size: float = 0.5,
        edge_tol: float = 1e-3,
        random_state=None,
        dtype="float64",
    ) -> None:
        super().__init__(avg_degree=avg_degree, edge_init=edge_init, maxit=maxit, tol=tol,
                         step_size=step_size, edge_tol=edge_tol, dtype=dtype)
        self.window_size = window_size
        self.random_state = random_state  # unused, kept for compatibility with the reference

    def _initialize(self, x):
        dt = _torch_dtype(self.dtype)
        xd = _to_device(x, dt)
        n_samples, n_nodes = xd.shape
        n_windows = -(-n_samples // self.window_size)
        sq_pdists = torch.stack([
            pdist_sq(xd[w * self.window_size:(w + 1) * self.window_size], dtype=dt) for w in range(n_windows)
        ])
        if self.avg_degree is None:
            self.theta_ = 1
        else:
            self.theta_ = np.array([[_theta_packed(sq_pdists[w], n_nodes, self.avg_degree)]
                                    for w in range(n_windows)])
        return sq_pdists
