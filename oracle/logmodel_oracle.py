"""CPU oracle for the LogModel hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module.  Nothing under
``graph_learning_b200/`` imports it; the product path has no CPU fallback.

This is a NumPy/SciPy *restatement* of the reference algorithm (LTS4/graph-learning),
written from the maths, every function citing the reference lines it follows
(paths relative to /root/reference).  Parity is PINNED: ``tests/golden/make_golden.py``
imports the unmodified reference in the build container and stores its outputs
for seeded inputs in ``tests/golden/*.npz``; ``tests/test_oracle.py`` checks every
function here against those vectors (bit-for-bit where the operation order is the same,
<=1e-12 relative otherwise).

Two forms of the solver are given:

* ``solve_two_pass``  -- the reference's own formulation: four sparse mat-vecs per
  iteration with an explicit CSR ``S`` (graph_learn/smooth_learning.py:146-167).  This is
  what ``bench.py --impl reference`` times.
* ``solve_one_pass``  -- the algebraically identical single sweep over the edge vector
  that the CUDA kernels implement (degrees of ``p`` and of the new ``w`` accumulated while
  the edges stream by).  Used to validate the reformulation against ``solve_two_pass``.
"""

from __future__ import annotations

from dataclasses import dataclass
from itertools import combinations_with_replacement

import numpy as np
from scipy import sparse

__all__ = [
    "n_nodes_from_edges",
    "edge_index",
    "sum_operator",
    "pdist_sq",
    "theta_for_degree",
    "solve_two_pass",
    "solve_one_pass",
    "solve_knn",
    "knn_support",
    "to_square",
    "to_packed",
    "laplacian_from_packed",
    "laplacian_adjoint",
    "smoothness",
    "glmm_log_prob",
    "SolveInfo",
]


# --------------------------------------------------------------------------------------
# packed <-> square helpers
# --------------------------------------------------------------------------------------
def n_nodes_from_edges(n_edges: int) -> int:
    """n such that n(n-1)/2 == n_edges (smooth_learning.py:133 rounds the same root)."""
    return int(np.round((1.0 + np.sqrt(1.0 + 8.0 * n_edges)) / 2.0))


def edge_index(n: int) -> tuple[np.ndarray, np.ndarray]:
    """(i, j) of every packed edge, row-major upper triangle (operators.py:39-41)."""
    i, j = np.triu_indices(n, k=1)
    return i.astype(np.int64), j.astype(np.int64)


def to_square(w: np.ndarray, n: int | None = None) -> np.ndarray:
    """Packed (..., E) -> symmetric (..., n, n) with zero diagonal (scipy squareform order)."""
    w = np.asarray(w)
    n = n or n_nodes_from_edges(w.shape[-1])
    i, j = edge_index(n)
    out = np.zeros(w.shape[:-1] + (n, n), dtype=w.dtype)
    out[..., i, j] = w
    out[..., j, i] = w
    return out


def to_packed(sq: np.ndarray) -> np.ndarray:
    """Square (..., n, n) -> packed (..., E) (operators.py:39-41)."""
    sq = np.asarray(sq)
    i, j = edge_index(sq.shape[-1])
    return sq[..., i, j]


def sum_operator(n: int) -> tuple[sparse.csr_array, sparse.csr_array]:
    """S (n x E) and S^T (E x n): S @ w = weighted degrees, S^T @ d = d_i + d_j.

    Same matrices as smooth_learning.py:13-44, built from the edge index instead of
    Python lists (the reference's construction is O(n^2) interpreter work).
    """
    i, j = edge_index(n)
    n_edges = i.size
    cols = np.arange(n_edges, dtype=np.int64)
    s = sparse.coo_array(
        (np.ones(2 * n_edges), (np.concatenate([i, j]), np.concatenate([cols, cols]))),
        shape=(n, n_edges),
    ).tocsr()
    return s, s.T.tocsr()


# --------------------------------------------------------------------------------------
# front end: distances and theta
# --------------------------------------------------------------------------------------
def pdist_sq(x: np.ndarray) -> np.ndarray:
    """Packed squared Euclidean distances between the COLUMNS of x (m, n).

    smooth_learning.py:278 computes ``pdist(x.T) ** 2``; this is the same quantity by
    direct differences, accumulated in float64 one row-block at a time.
    """
    x = np.asarray(x, dtype=np.float64)
    n = x.shape[1]
    i, j = edge_index(n)
    out = np.zeros(i.size)
    step = max(1, (1 << 24) // max(1, i.size))
    for s in range(0, x.shape[0], step):
        blk = x[s : s + step]
        diff = blk[:, i] - blk[:, j]
        out += np.einsum("te,te->e", diff, diff)
    return out


def theta_for_degree(z_sq: np.ndarray, avg_degree, blocks=None):
    """Kalofolias-Perraudin theta for a requested mean degree.

    smooth_learning.py:201-219: rows of the SQUARE matrix are sorted with the zero diagonal
    included; b = sum of the first k, theta_min uses sorted[k], theta_max sorted[k-1],
    result sqrt(mean(theta_min) * mean(theta_max)).  Block form :220-235.
    """
    z_sq = np.asarray(z_sq, dtype=np.float64)
    if blocks is None:
        k = avg_degree
        srt = np.sort(z_sq, axis=1)
        b = srt[:, :k].sum(axis=1)
        with np.errstate(divide="ignore", invalid="ignore"):
            lo = np.mean((k * srt[:, k] ** 2 - b * srt[:, k]) ** -0.5)
            hi = np.mean((k * srt[:, k - 1] ** 2 - b * srt[:, k - 1]) ** -0.5)
        return float(np.sqrt(lo * hi))
    blocks = np.asarray(blocks)
    full = np.zeros((blocks.size, blocks.size))
    for a, b_ in combinations_with_replacement(np.unique(blocks).tolist(), 2):
        ra = np.where(blocks == a)[0][:, None]
        rb = np.where(blocks == b_)[0][None, :]
        k = avg_degree[(a, b_)] if isinstance(avg_degree, dict) else avg_degree
        t = theta_for_degree(z_sq[ra, rb], k)
        full[ra, rb] = t
        full[rb.T, ra.T] = t
    return to_packed(full)


# --------------------------------------------------------------------------------------
# solver
# --------------------------------------------------------------------------------------
@dataclass
class SolveInfo:
    n_iter: int
    rel_primal: float
    rel_dual: float
    converged: bool


def _norm2(a: np.ndarray) -> float:
    """2-norm of a vector, spectral norm of a matrix -- what np.linalg.norm(., ord=2) returns
    for the (E,) and (E,K) operands at smooth_learning.py:158-159."""
    return float(np.linalg.norm(a, ord=2))


def _gamma(n: int, beta: float, step_size: float) -> float:
    # smooth_learning.py:135-137; sqrt(2(n-1)) is the exact 2-norm of S.
    return step_size / (2.0 * beta + np.sqrt(2.0 * (n - 1)))


def solve_two_pass(
    distances,
    alpha=1.0,
    beta=1.0,
    edge_init=None,
    maxit=1000,
    tol=1e-5,
    step_size=0.5,
    edge_tol=1e-3,
    *,
    threshold=True,
    history=None,
    operator=None,
):
    """FBF primal-dual iteration exactly as the reference schedules it
    (smooth_learning.py:125-174): two applications of S and two of S^T per iteration.

    distances: (E,) or (K, E).  Returns (w, info) with w shaped like ``distances``
    (NOT expanded to (n, n); callers use ``to_square``).  ``history`` (a list) receives
    (w, d) copies after every iteration when given.  ``operator`` = a prebuilt (S, S^T) pair
    (bench.py times the iterations without the operator construction, which the reference redoes
    on every call with O(n^2) interpreter work, smooth_learning.py:133).
    """
    z = np.asarray(distances, dtype=np.float64)
    if z.ndim > 2:
        raise ValueError("Distances must be vector or matrix")
    batched = z.ndim == 2
    zt = z.T if batched else z  # (E, K) internally, smooth_learning.py:128-129
    n = n_nodes_from_edges(zt.shape[0])
    s_op, st_op = operator if operator is not None else sum_operator(n)
    g = _gamma(n, beta, step_size)
    c = 1.0 - 2.0 * g * beta

    w = np.zeros_like(zt) if edge_init is None else np.asarray(edge_init, dtype=np.float64)
    d = s_op @ w
    rp = rd = np.inf
    it = 0
    converged = False
    for it in range(1, maxit + 1):
        y = c * w - g * (st_op @ d)  # :147
        yb = d + g * (s_op @ w)  # :148
        p = np.maximum(y - 2.0 * g * zt, 0.0)  # :150
        pb = (yb - np.sqrt(yb * yb + 4.0 * alpha * g)) / 2.0  # :151
        q = c * p - g * (st_op @ pb)  # :153
        qb = pb + g * (s_op @ p)  # :154
        if it > 1:  # :156-161 -- denominators are the pre-update iterates
            with np.errstate(divide="ignore", invalid="ignore"):
                rp = _norm2(q - y) / _norm2(w)
                rd = _norm2(qb - yb) / _norm2(d)
        w = w - y + q  # :163
        d = d - yb + qb  # :164
        if history is not None:
            history.append((w.copy(), d.copy()))
        if rp < tol and rd < tol:  # :166
            converged = True
            break
    else:
        it = maxit
    if threshold:
        w = np.where(w < edge_tol, 0.0, w)  # :169
    out = w.T if batched else w
    return out, SolveInfo(it, float(rp), float(rd), converged)


def solve_one_pass(
    distances,
    alpha=1.0,
    beta=1.0,
    edge_init=None,
    maxit=1000,
    tol=1e-5,
    step_size=0.5,
    edge_tol=1e-3,
    *,
    threshold=True,
):
    """Same iteration with ONE sweep over the edges per iteration (the CUDA schedule).

    The dual prox input yb = d + g*S w is known before the sweep because S w was
    accumulated by the previous sweep (and S w0 = d0 by smooth_learning.py:144).  Each
    sweep reads w and z, writes w, and accumulates S p and S w_new; the node tail finishes
    qb and d.  Matrix-free: S and S^T are applied with bincount / fancy indexing.
    """
    z = np.asarray(distances, dtype=np.float64)
    if z.ndim > 2:
        raise ValueError("Distances must be vector or matrix")
    batched = z.ndim == 2
    zk = z if batched else z[None, :]  # (K, E)
    n_graphs, n_edges = zk.shape
    n = n_nodes_from_edges(n_edges)
    ei, ej = edge_index(n)
    g = _gamma(n, beta, step_size)
    c = 1.0 - 2.0 * g * beta

    def deg(v):  # S v for every graph
        out = np.zeros((n_graphs, n))
        for k in range(n_graphs):
            out[k] = np.bincount(ei, v[k], n) + np.bincount(ej, v[k], n)
        return out

    if edge_init is None:
        w = np.zeros_like(zk)
    else:
        w0 = np.asarray(edge_init, dtype=np.float64)
        w = (w0.T if batched else w0[None, :]).copy()
    d = deg(w)
    sw = d.copy()
    rp = rd = np.inf
    it = 0
    converged = False
    for it in range(1, maxit + 1):
        yb = d + g * sw
        pb = (yb - np.sqrt(yb * yb + 4.0 * alpha * g)) / 2.0
        a, b = g * d, g * pb
        y = c * w - (a[:, ei] + a[:, ej])
        p = np.maximum(y - 2.0 * g * zk, 0.0)
        q = c * p - (b[:, ei] + b[:, ej])
        r = q - y
        w_new = w + r
        sp, sw = deg(p), deg(w_new)
        qb = pb + g * sp
        rb = qb - yb
        if it > 1:
            with np.errstate(divide="ignore", invalid="ignore"):
                rp = _norm2(r.T) / _norm2(w.T)
                rd = _norm2(rb.T) / _norm2(d.T)
        w, d = w_new, d + rb
        if rp < tol and rd < tol:
            converged = True
            break
    else:
        it = maxit
    if threshold:
        w = np.where(w < edge_tol, 0.0, w)
    out = w if batched else w[0]
    return out, SolveInfo(it, float(rp), float(rd), converged)


# --------------------------------------------------------------------------------------
# kNN-restricted support (config 5) -- NO reference code exists for this; PARITY UNPINNED
# by the reference.  Restatement = the same iteration with S restricted to the columns of
# the support (S[:, mask]) and step bound ||S_mask||_2 <= sqrt(2 * d_max).
# --------------------------------------------------------------------------------------
def knn_support(z_sq: np.ndarray, k: int) -> tuple[np.ndarray, np.ndarray]:
    """Symmetrised k-nearest-neighbour edge list (i<j, sorted by (i, j)) of a square
    distance matrix.  Ties are broken by column index (stable argsort)."""
    z_sq = np.asarray(z_sq, dtype=np.float64).copy()
    n = z_sq.shape[0]
    np.fill_diagonal(z_sq, np.inf)
    nbr = np.argsort(z_sq, axis=1, kind="stable")[:, :k]
    rows = np.repeat(np.arange(n), k)
    lo = np.minimum(rows, nbr.ravel())
    hi = np.maximum(rows, nbr.ravel())
    key = np.unique(lo.astype(np.int64) * n + hi)
    return (key // n).astype(np.int64), (key % n).astype(np.int64)


def solve_knn(
    ei,
    ej,
    z,
    n,
    alpha=1.0,
    beta=1.0,
    edge_init=None,
    maxit=1000,
    tol=1e-5,
    step_size=0.5,
    edge_tol=1e-3,
    *,
    threshold=True,
):
    """LogModel restricted to the edge list (ei, ej); z holds the (theta-scaled) squared
    distances of those edges.  With the complete edge list this IS solve_one_pass except for
    the step bound: gamma = step / (2 beta + sqrt(2 d_max)), d_max = largest support degree
    (= n-1 on the complete graph, recovering smooth_learning.py:135-137)."""
    ei = np.asarray(ei, dtype=np.int64)
    ej = np.asarray(ej, dtype=np.int64)
    z = np.asarray(z, dtype=np.float64)
    dmax = int(np.max(np.bincount(ei, minlength=n) + np.bincount(ej, minlength=n)))
    g = step_size / (2.0 * beta + np.sqrt(2.0 * dmax))
    c = 1.0 - 2.0 * g * beta

    def deg(v):
        return np.bincount(ei, v, n) + np.bincount(ej, v, n)

    w = np.zeros_like(z) if edge_init is None else np.asarray(edge_init, dtype=np.float64).copy()
    d = deg(w)
    sw = d.copy()
    rp = rd = np.inf
    it = 0
    converged = False
    for it in range(1, maxit + 1):
        yb = d + g * sw
        pb = (yb - np.sqrt(yb * yb + 4.0 * alpha * g)) / 2.0
        y = c * w - g * (d[ei] + d[ej])
        p = np.maximum(y - 2.0 * g * z, 0.0)
        q = c * p - g * (pb[ei] + pb[ej])
        r = q - y
        w_new = w + r
        sp, sw = deg(p), deg(w_new)
        rb = pb + g * sp - yb
        if it > 1:
            with np.errstate(divide="ignore", invalid="ignore"):
                rp = _norm2(r) / _norm2(w)
                rd = _norm2(rb) / _norm2(d)
        w, d = w_new, d + rb
        if rp < tol and rd < tol:
            converged = True
            break
    else:
        it = maxit
    if threshold:
        w = np.where(w < edge_tol, 0.0, w)
    return w, SolveInfo(it, float(rp), float(rd), converged)


def knn_row_order(ei, ej, n: int, order_key=None, sort_window: int = 0) -> np.ndarray:
    """Internal row order of the kNN solver's sliced layout (include/gl_b200.h, gl_knn_create_ordered):
    nodes arranged by ascending ``order_key`` (stable; None = their numbering), then sorted by
    decreasing support degree inside windows of ``sort_window`` positions (0: one window; < 0: no
    degree sort).  Returns ``order`` with order[p] = node stored in row p.  No reference code (A9)."""
    deg = np.bincount(np.asarray(ei, dtype=np.int64), minlength=n) + np.bincount(np.asarray(ej, dtype=np.int64), minlength=n)
    arr = np.arange(n) if order_key is None else np.argsort(np.asarray(order_key), kind="stable")
    if sort_window < 0:
        return arr
    win = np.arange(n) // sort_window if sort_window > 0 else np.zeros(n, dtype=np.int64)
    # stable: ties keep the arrangement
    return arr[np.lexsort((np.arange(n), -deg[arr], win))]


def knn_padded_slots(ei, ej, n: int, order: np.ndarray, slice_rows: int = 32) -> int:
    """Entries of a sliced-ELL layout with ``slice_rows`` rows per slice in the given row order,
    padding included: every slice is as wide as its longest row."""
    deg = np.bincount(np.asarray(ei, dtype=np.int64), minlength=n) + np.bincount(np.asarray(ej, dtype=np.int64), minlength=n)
    d = deg[order]
    pad = (-len(d)) % slice_rows
    d = np.concatenate([d, np.zeros(pad, dtype=d.dtype)]).reshape(-1, slice_rows)
    return int(slice_rows * d.max(axis=1).sum())


def morton_keys(x: np.ndarray) -> np.ndarray:
    """Z-order codes of the columns of x (m, n): the (up to) three features of largest variance
    quantised to 10 bits each, bits interleaved with feature 0 lowest -- a bit-by-bit restatement of
    graph_learning_b200.knn.locality_keys (host-side glue, no reference code)."""
    x = np.asarray(x, dtype=np.float64)
    if x.shape[0] > 3:
        top = np.argsort(-x.var(axis=1), kind="stable")[:3]
        x = x[np.sort(top)]
    lo = x.min(axis=1, keepdims=True)
    span = np.maximum(x.max(axis=1, keepdims=True) - lo, np.finfo(np.float64).tiny)
    q = np.clip(((x - lo) / span * 1024.0).astype(np.int64), 0, 1023)
    key = np.zeros(x.shape[1], dtype=np.int64)
    for a in range(q.shape[0]):
        for b in range(10):
            key |= ((q[a] >> b) & 1) << (3 * b + a)
    return key.astype(np.int32)


# --------------------------------------------------------------------------------------
# dense-Laplacian forms of the same operator pair (operators.py:44-94)
# --------------------------------------------------------------------------------------
def laplacian_from_packed(w: np.ndarray) -> np.ndarray:
    """(K, E) or (E,) edge weights -> Laplacians D - W; diagonal = S w (operators.py:44-70)."""
    sq = to_square(w)
    lap = -sq
    idx = np.arange(sq.shape[-1])
    lap[..., idx, idx] = sq.sum(axis=-1)
    return lap


def laplacian_adjoint(lap: np.ndarray) -> np.ndarray:
    """Adjoint of ``laplacian_from_packed``: L -> L_ii + L_jj - L_ij - L_ji over i<j
    (operators.py:73-94)."""
    lap = np.asarray(lap)
    dg = np.diagonal(lap, axis1=-2, axis2=-1)
    full = dg[..., :, None] + dg[..., None, :] - lap - np.swapaxes(lap, -1, -2)
    return to_packed(full)


# --------------------------------------------------------------------------------------
# assignment step of the batched callers
# --------------------------------------------------------------------------------------
def smoothness(x: np.ndarray, laplacians: np.ndarray, means: np.ndarray | None = None) -> np.ndarray:
    """(m, n) signals x (K, n, n) matrices -> (m, K) quadratic forms y^T L_k y with y = x (k-Graphs,
    kgraphs.py:87-88) or y = x - means[k] (k-Graphs V2, kgraphs.py:203-205).  Dense contraction."""
    x = np.asarray(x, dtype=np.float64)
    out = np.empty((x.shape[0], laplacians.shape[0]))
    for k, lap in enumerate(np.asarray(laplacians, dtype=np.float64)):
        y = x if means is None else x - means[k]
        out[:, k] = np.sum((y @ lap) * y, axis=1)
    return out


def glmm_log_prob(x: np.ndarray, laplacians: np.ndarray, means: np.ndarray, regul: float) -> np.ndarray:
    """GLMM E-step log-probabilities (glmm.py:238-279): eigendecompose each Laplacian, add ``regul``
    to the spectrum, drop the first eigenpair, Gaussian log-density in the remaining n - 1 directions.
    ``log(prod(.))`` is kept as written there (it overflows for large n exactly like the reference)."""
    x = np.asarray(x, dtype=np.float64)
    n = x.shape[1]
    eigval, eigvec = np.linalg.eigh(np.asarray(laplacians, dtype=np.float64))
    eigval = eigval + regul
    out = np.empty((x.shape[0], laplacians.shape[0]))
    for k in range(laplacians.shape[0]):
        proj = (x - means[k]) @ eigvec[k][:, 1:]
        with np.errstate(over="ignore", divide="ignore"):
            logdet = np.log(np.prod(eigval[k, 1:]))
        out[:, k] = (-(n - 1) * np.log(2 * np.pi) + logdet - np.sum(proj * proj * eigval[k, 1:], axis=1)) / 2
    return out
