"""GPU parity tests of the kNN-support variant (K6).  No reference code exists for it (parity
unpinned by the reference): pinned against the CPU restatement
oracle.solve_knn, and against the DENSE solver (which is pinned to the reference) when the edge
list is the complete graph."""
import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def test_complete_support_equals_dense(golden):
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    g = golden("solver_n100")
    z = g["z"] * g["theta"]
    ei, ej = orc.edge_index(100)
    w, info = knn(ei, ej, z, 100, maxit=50, edge_tol=-np.inf, return_info=True)
    assert info.n_iter == 50
    assert rel(orc.to_square(w), g["w_raw_it50"]) < 1e-11
    w = knn(ei, ej, z, 100)
    assert rel(orc.to_square(w), g["w_final"]) < 1e-9


@pytest.mark.parametrize("n,k", [(200, 5), (1000, 10)])
def test_knn_support_vs_oracle(n, k):
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    rng = np.random.default_rng(n)
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    z = zsq[ei, ej] * (n ** (2 / 3))
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=300)
    w, info = knn(ei, ej, z, n, maxit=300, return_info=True)
    assert info.n_iter == io.n_iter
    assert rel(w, wo) < 1e-9
    assert np.array_equal(w > 0, wo > 0)
    w32 = knn(ei, ej, z.astype(np.float32), n, maxit=300)
    assert w32.dtype == np.float32 and rel(w32, wo) < 1e-4


def test_knn_warm_start_and_empty():
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    rng = np.random.default_rng(5)
    n = 50
    pts = rng.uniform(size=(n, 2))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, 4)
    z = zsq[ei, ej] * 20
    init = rng.uniform(0, 0.2, size=z.shape)
    wo, _ = orc.solve_knn(ei, ej, z, n, edge_init=init, maxit=20, threshold=False)
    w = knn(ei, ej, z, n, edge_init=init, maxit=20, edge_tol=-np.inf)
    assert rel(w, wo) < 1e-12
    assert knn(np.zeros(0, int), np.zeros(0, int), np.zeros(0), n, maxit=3).shape == (0,)
