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


@pytest.mark.parametrize("maxit", [1, 2, 15, 16, 17, 33, 100])
def test_one_launch_iteration_any_iteration_count(maxit):
    """The one-shot solve runs the whole iteration in one launch with double-buffered (a, b) pairs whose
    roles are baked into a 16-iteration graph: odd counts, counts around the graph length and stops in
    the middle of a replay must leave the same iterate as the oracle (raw weights, no threshold)."""
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    n, k = 300, 6
    rng = np.random.default_rng(7)
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    z = zsq[ei, ej] * (n ** (2 / 3))
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=maxit, tol=0.0, threshold=False)
    w, info = knn(ei, ej, z, n, maxit=maxit, tol=0.0, edge_tol=-np.inf, return_info=True)
    assert info.n_iter == io.n_iter == maxit
    assert rel(w, wo) < 1e-12
    if maxit > 1:
        assert abs(info.rel_primal - io.rel_primal) < 1e-9 * io.rel_primal
        assert abs(info.rel_dual - io.rel_dual) < 1e-9 * io.rel_dual


def test_one_launch_mode_needs_a_single_shard():
    import ctypes
    import torch
    from graph_learning_b200 import _native as nat
    ei = torch.tensor([0, 1], dtype=torch.int32, device="cuda")
    ej = torch.tensor([1, 2], dtype=torch.int32, device="cuda")
    z = torch.ones(2, dtype=torch.float32, device="cuda")
    h = ctypes.c_void_p()
    nat.check(nat.lib().gl_knn_create(ctypes.byref(h), ei.data_ptr(), ej.data_ptr(), z.data_ptr(), None, 2, 300,
                                      nat.GL_F32, 0, 128, 384, None))
    try:
        assert nat.lib().gl_knn_set_fused(h, 2) == nat.GL_E_STATE      # rows [0, 128) of 384
        assert nat.lib().gl_knn_set_fused(h, 3) == nat.GL_E_ARG
        assert nat.lib().gl_knn_set_fused(h, 1) == nat.GL_OK
    finally:
        nat.lib().gl_knn_destroy(h)


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


# ---- kNN front end (N3): search, symmetrised support, theta from the lists, estimator ---------------
def _points(n, dim, seed):
    return np.random.default_rng(seed).uniform(size=(dim, n)).astype(np.float32)  # (m, n): nodes = columns


@pytest.mark.parametrize("n,dim,k", [(50, 2, 4), (300, 3, 10), (1000, 8, 20), (129, 40, 32), (20, 3, 32)])
def test_knn_search_vs_bruteforce(n, dim, k):
    from graph_learning_b200.knn import knn_search
    x = _points(n, dim, n + k)
    idx, d2 = knn_search(x, k)
    idx, d2 = idx.cpu().numpy(), d2.cpu().numpy()
    zsq = orc.to_square(orc.pdist_sq(x.astype(np.float64)))
    np.fill_diagonal(zsq, np.inf)
    kk = min(k, n - 1)
    order = np.argsort(zsq, axis=1, kind="stable")[:, :kk]
    ref_d = np.take_along_axis(zsq, order, axis=1)
    assert np.all(np.diff(d2[:, :kk], axis=1) >= 0)                      # ascending
    assert np.allclose(d2[:, :kk], ref_d, rtol=2e-6, atol=1e-12)
    assert (idx[:, :kk] != np.arange(n)[:, None]).all()                  # self excluded
    # the index sets agree wherever the k-th and (k+1)-th distances are separated
    if kk < n - 1:
        nxt = np.sort(zsq, axis=1)[:, kk]
        clear = (nxt - ref_d[:, -1]) > 1e-5 * nxt
        same = np.array([set(a) == set(b) for a, b in zip(idx[:, :kk], order)])
        assert same[clear].all() and clear.mean() > 0.9
    else:
        assert (np.sort(idx[:, :kk], axis=1) == np.sort(order, axis=1)).all()
        assert (idx[:, kk:] == -1).all() and np.isinf(d2[:, kk:]).all()


def test_knn_graph_matches_oracle_support():
    from graph_learning_b200.knn import knn_graph
    n, k = 400, 6
    x = _points(n, 3, 7)
    ei, ej, d2, _ = knn_graph(x, k)
    zsq = orc.to_square(orc.pdist_sq(x.astype(np.float64)))
    oi, oj = orc.knn_support(zsq, k)
    assert np.array_equal(ei.cpu().numpy(), oi) and np.array_equal(ej.cpu().numpy(), oj)
    assert np.allclose(d2.cpu().numpy(), zsq[oi, oj], rtol=2e-6)


def test_theta_from_knn_equals_dense_theta():
    from graph_learning_b200.knn import knn_search, theta_from_knn
    n, k = 300, 12
    x = _points(n, 3, 3)
    _, d2 = knn_search(x, k)
    zsq = orc.to_square(orc.pdist_sq(x.astype(np.float64)))
    for deg in (3, 8, 12):
        ref = orc.theta_for_degree(zsq, deg)
        assert abs(theta_from_knn(d2, deg) - ref) < 5e-6 * ref
    with pytest.raises(ValueError):
        theta_from_knn(d2, 13)


def test_knn_logmodel_estimator_end_to_end():
    """KNNLogModel.fit = search + theta + sparse solve; against the oracle chain on the same data."""
    from graph_learning_b200.knn import KNNLogModel
    n, k, deg = 500, 10, 5
    x = _points(n, 3, 11)
    mdl = KNNLogModel(avg_degree=deg, n_neighbors=k, maxit=300).fit(x)
    zsq = orc.to_square(orc.pdist_sq(x.astype(np.float64)))
    oi, oj = orc.knn_support(zsq, k)
    theta = orc.theta_for_degree(zsq, deg)
    wo, io = orc.solve_knn(oi, oj, zsq[oi, oj] * theta, n, maxit=300)
    assert np.array_equal(mdl.edges_[0], oi) and np.array_equal(mdl.edges_[1], oj)
    assert abs(mdl.theta_ - theta) < 5e-6 * theta
    assert mdl.n_iter_ == io.n_iter
    assert rel(mdl.weights_, wo) < 1e-4
    a = mdl.to_sparse()
    assert a.shape == (n, n) and abs(a - a.T).max() == 0
    mean_deg = (a > 0).sum() / n
    assert 0.5 * deg < mean_deg < 2.0 * deg   # the heuristic targets avg_degree neighbours per node


# ---- row-sharded stepwise solver (gl_knn_*), SURVEY 8(e) ------------------------------------------------
def _knn_case(n, k):
    rng = np.random.default_rng(n)
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    return ei, ej, zsq[ei, ej] * (n ** (2 / 3))


@pytest.mark.parametrize("dtype,tol", [("float64", 1e-10), ("float32", 1e-4)])
def test_stepwise_knn_solver_single_shard(dtype, tol):
    from graph_learning_b200.sharded import ShardedKNNSolver
    n, k, maxit = 700, 8, 250
    ei, ej, z = _knn_case(n, k)
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=maxit)
    dt = torch.float64 if dtype == "float64" else torch.float32
    sv = ShardedKNNSolver(ei, ej, z, n, dtype=dt)
    sv.begin(maxit=maxit)
    sv.run(poll_every=16)
    st = sv.status()
    w = sv.gather().cpu().numpy()
    sv.close()
    assert st.n_iter == io.n_iter
    assert rel(w, wo) < tol
    if dtype == "float64":
        assert np.array_equal(w > 0, wo > 0)


@pytest.mark.parametrize("world", [2, 3])
def test_stepwise_knn_solver_emulated_shards(world):
    """Several row shards in one process on one GPU: the exchange (all-gather of the (a, b) slices,
    sum of the four scalars) is done by hand between the shards' buffers; results must equal the
    oracle and every shard must take the same stop decision."""
    from graph_learning_b200.sharded import ShardedKNNSolver
    n, k, maxit = 900, 6, 1000
    ei, ej, z = _knn_case(n, k)
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=maxit)
    shards = [ShardedKNNSolver(ei, ej, z, n, dtype=torch.float64, rank=r, world=world) for r in range(world)]
    per = shards[0].per
    assert sum(s.lib.gl_knn_local_slots(s._h) for s in shards) == 2 * len(ei)

    def exchange():
        for r, s in enumerate(shards):
            for t in shards:
                if t is not s:
                    t.ab()[r * per:(r + 1) * per].copy_(s.ab()[r * per:(r + 1) * per])
        tot = sum(s.scalars().clone() for s in shards)
        for s in shards:
            s.scalars().copy_(tot)

    for s in shards:
        s.begin(maxit=maxit)
    exchange()
    for s in shards:
        s.decide(True)
    for it in range(maxit):
        for s in shards:
            s.edge_step()
            s.node_step()
        exchange()
        for s in shards:
            s.decide(False)
        if it % 25 == 24 and shards[0].done():
            break
    sts = [s.status() for s in shards]
    w = sum(s.gather() for s in shards).cpu().numpy()
    for s in shards:
        s.close()
    assert all(st.n_iter == sts[0].n_iter and st.converged == sts[0].converged for st in sts)
    assert sts[0].n_iter == io.n_iter
    assert rel(w, wo) < 1e-10
    assert np.array_equal(w > 0, wo > 0)
