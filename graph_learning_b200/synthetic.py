"""Synthetic inputs for the LogModel benchmarks and parity tests.

BASELINE.json asks for "synthetic smooth signals drawn from a random-geometric-graph
Laplacian".  Two generators:

* :func:`rgg_signals_exact` (NumPy, small n): LGMRF samples x ~ N(0, pinv(L)), the same
  distribution the reference's ``sampling/signals.py:7-28`` draws from, via an ``eigh`` of L.
* :func:`rgg_signals_filtered` (torch, any device, large n): white noise low-pass filtered by
  repeated heat-diffusion steps ``X <- X - eps * X L``; avoids ``eigh``/``pinv`` at n = 20,000
  (SURVEY.md section 8(d)).  Not the same distribution as the exact sampler, only smooth on
  the same graph; used where no reference comparison of the *signals* is involved.

Both are host-side utilities (data generation, not the hot path).
"""

from __future__ import annotations

import math

import numpy as np

__all__ = ["rgg_laplacian", "rgg_signals_exact", "rgg_signals_filtered", "rgg_points"]


def rgg_points(n: int, seed: int = 0, dim: int = 2) -> np.ndarray:
    return np.random.default_rng(seed).uniform(size=(n, dim))


def rgg_laplacian(n: int, seed: int = 0, c: float = 1.5) -> np.ndarray:
    """Dense Laplacian of a random geometric graph: n uniform points in the unit square,
    pairs closer than r = c*sqrt(log n / n) connected with Gaussian-kernel weights."""
    pts = rgg_points(n, seed)
    r = c * math.sqrt(math.log(n) / n)
    d2 = ((pts[:, None, :] - pts[None, :, :]) ** 2).sum(-1)
    w = np.exp(-d2 / (2.0 * (r / 2.0) ** 2)) * (d2 < r * r)
    np.fill_diagonal(w, 0.0)
    return np.diag(w.sum(1)) - w


def rgg_signals_exact(n: int, m: int, seed: int = 0, c: float = 1.5) -> np.ndarray:
    """(m, n) float64 LGMRF samples on the RGG Laplacian (nodes are COLUMNS, as the reference's
    estimators expect, smooth_learning.py:277-278)."""
    lap = rgg_laplacian(n, seed, c)
    lam, u = np.linalg.eigh(lap)
    keep = lam > 1e-9 * lam[-1]
    basis = u[:, keep] / np.sqrt(lam[keep])
    g = np.random.default_rng(seed + 1).standard_normal((m, int(keep.sum())))
    return g @ basis.T


def rgg_signals_filtered(n: int, m: int, seed: int = 0, c: float = 1.5, steps: int = 20,
                         device="cpu", dtype=None, chunk: int = 4096):
    """(m, n) smooth signals for large n, generated with torch on ``device``.

    White noise G (m, n) is diffused ``steps`` times with X <- X - eps * X L, eps = 0.9/d_max,
    i.e. X = G (I - eps L)^steps: a low-pass graph filter.  The graph is built in row chunks so
    the n x n distance matrix is never held at once.
    """
    import torch

    dtype = dtype or torch.float32
    gen = torch.Generator(device="cpu").manual_seed(seed)
    pts = torch.rand((n, 2), generator=gen, dtype=torch.float64).to(device)
    r = c * math.sqrt(math.log(n) / n)
    rows, cols, vals = [], [], []
    for s in range(0, n, chunk):
        d2 = torch.cdist(pts[s : s + chunk], pts).square_()
        mask = d2 < r * r
        idx = mask.nonzero(as_tuple=False)
        keep = idx[:, 0] + s != idx[:, 1]
        idx = idx[keep]
        rows.append(idx[:, 0] + s)
        cols.append(idx[:, 1])
        vals.append(torch.exp(-d2[idx[:, 0], idx[:, 1]] / (2.0 * (r / 2.0) ** 2)))
    rows, cols, vals = torch.cat(rows), torch.cat(cols), torch.cat(vals).to(dtype)
    deg = torch.zeros(n, dtype=dtype, device=device).index_add_(0, rows, vals)
    eps = 0.9 / float(deg.max())
    # A = I - eps*L = I - eps*D + eps*W  (symmetric), applied to X^T (n, m)
    adj = torch.sparse_coo_tensor(torch.stack([rows, cols]), eps * vals, (n, n)).coalesce()
    gen_d = torch.Generator(device=device).manual_seed(seed + 1)
    xt = torch.randn((n, m), generator=gen_d, dtype=dtype, device=device)
    diag = (1.0 - eps * deg)[:, None]
    for _ in range(steps):
        xt = diag * xt + torch.sparse.mm(adj, xt)
    return xt.t().contiguous()
