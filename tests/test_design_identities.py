"""CPU checks of the algebraic identities the CUDA kernels rely on where they do NOT follow the
reference's operation order (DESIGN.md K4 batch, K6, K8).  NumPy only; the kernels themselves are
tested against the oracle and the goldens in the -m gpu files."""
import numpy as np
import pytest

from oracle import logmodel_oracle as orc


@pytest.mark.parametrize("seed", range(8))
def test_lazy_stop_rule_bounds_are_lower_bounds(seed):
    """K4 batch: rel = sigma_max(R) / sigma_max(W) >= sqrt(max_k ||r_k||^2) / sqrt(sum_k ||w_k||^2), so
    'bound >= tol' rules the stop out exactly (the spectral norms are what smooth_learning.py:158-159
    takes of the 2-D operands)."""
    rng = np.random.default_rng(seed)
    e, k = int(rng.integers(5, 400)), int(rng.integers(2, 9))
    r = rng.standard_normal((e, k)) * rng.uniform(1e-6, 1.0, size=(1, k))
    w = rng.standard_normal((e, k))
    if seed % 2:                      # strongly correlated graphs: the bound is loosest there
        w = w[:, :1] + 1e-3 * w
        r = r[:, :1] + 1e-3 * r
    exact = np.linalg.norm(r, 2) / np.linalg.norm(w, 2)
    bound = np.sqrt((r * r).sum(0).max()) / np.sqrt((w * w).sum())
    assert bound <= exact * (1 + 1e-12)
    assert exact <= bound * np.sqrt(k) * np.sqrt(k) * (1 + 1e-12)     # never looser than a factor K


@pytest.mark.parametrize("seed", range(4))
def test_previous_iterate_is_recovered_from_the_update(seed):
    """K4 batch: the deciding iterations form Gram(W' - R) for the previous iterate (w' = w - y + q,
    r = q - y, smooth_learning.py:163); in float64 the difference reproduces ||W||_2 to rounding."""
    rng = np.random.default_rng(seed)
    w = np.abs(rng.standard_normal((300, 4)))
    y, q = rng.standard_normal((300, 4)), rng.standard_normal((300, 4))
    w_new, r = (w - y) + q, q - y
    assert abs(np.linalg.norm(w_new - r, 2) - np.linalg.norm(w, 2)) < 1e-13 * np.linalg.norm(w, 2)


@pytest.mark.parametrize("seed", range(4))
def test_entry_form_of_the_quadratic_form_holds_for_any_matrix(seed):
    """K8: sum_i A_ii y_i^2 + sum_{i<j} (A_ij + A_ji) y_i y_j == y^T A y for ANY square A, evaluated
    row by row as y_i * (A_ii y_i + sum_{j>i} (A_ij + A_ji) y_j)."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(2, 40))
    a = rng.standard_normal((n, n)) * (rng.uniform(size=(n, n)) < 0.3)
    y = rng.standard_normal(n)
    sym = np.triu(a + a.T, 1)
    rows = sum(y[i] * (a[i, i] * y[i] + sym[i] @ y) for i in range(n))
    assert abs(rows - y @ a @ y) < 1e-12 * max(1.0, abs(y @ a @ y))


@pytest.mark.parametrize("seed", range(4))
def test_glmm_projection_identity(seed):
    """K8 / N1: sum_{j>=1} (lam_j + r)(v_j . y)^2 = y^T (L + r I) y - (lam_0 + r)(v_0 . y)^2
    (glmm.py:253-274 drops the first eigenpair)."""
    rng = np.random.default_rng(seed)
    n, regul = 30, 0.15
    w = rng.uniform(0.1, 2.0, size=n * (n - 1) // 2) * (rng.uniform(size=n * (n - 1) // 2) < 0.3)
    lap = orc.laplacian_from_packed(w)
    lam, vec = np.linalg.eigh(lap)
    y = rng.standard_normal(n)
    proj = y @ vec[:, 1:]
    lhs = np.sum(proj * proj * (lam[1:] + regul))
    rhs = y @ (lap + regul * np.eye(n)) @ y - (lam[0] + regul) * (vec[:, 0] @ y) ** 2
    assert abs(lhs - rhs) < 1e-11 * abs(lhs)
    # the log-determinant: the reference's log(prod(.)) equals the sum of logs where the product is finite
    assert abs(np.log(np.prod(lam[1:] + regul)) - np.sum(np.log(lam[1:] + regul))) < 1e-12 * n


def test_knn_relabelling_leaves_edge_results_unchanged():
    """K6: renumbering the nodes (here: by decreasing degree) permutes only the node vectors; the
    per-edge weights of the sparse-support solve are the same."""
    rng = np.random.default_rng(3)
    n, k = 60, 4
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    z = zsq[ei, ej] * n ** (2 / 3)
    w0, i0 = orc.solve_knn(ei, ej, z, n, maxit=200)
    deg = np.bincount(np.concatenate([ei, ej]), minlength=n)
    order = np.argsort(-deg, kind="stable")
    lab = np.empty(n, dtype=np.int64)
    lab[order] = np.arange(n)
    w1, i1 = orc.solve_knn(lab[ei], lab[ej], z, n, maxit=200)
    assert i0.n_iter == i1.n_iter
    assert np.linalg.norm(w0 - w1) < 1e-12 * np.linalg.norm(w0)
