"""Row ordering of the kNN solver's sliced layout (gl_knn_create_ordered / locality_keys).

CPU part: the Morton keys against a bit-by-bit restatement, and the properties of the windowed
degree sort the library implements (restated in oracle.knn_row_order).

GPU part: WRITTEN AFTER THE ROUND'S GPU MINUTES WERE SPENT -- the ordered entry points have been
compiled but never run on a device.  The default path (global degree sort) is untouched and keeps its
own tests in test_gpu_knn.py; the tests below are skipped unless GL_RUN_UNVALIDATED=1 so that an
unmeasured option cannot turn the GPU suite red.  First thing next round: run them, then drop the gate.
"""
import os

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

torch = pytest.importorskip("torch")

unvalidated = pytest.mark.skipif(os.environ.get("GL_RUN_UNVALIDATED") != "1",
                                 reason="ordered kNN layout not yet run on a GPU (set GL_RUN_UNVALIDATED=1)")


def _case(n, k, dim=3, seed=0):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(size=(n, dim))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    return pts, ei, ej, zsq[ei, ej] * (n ** (2 / dim))


# ---- CPU ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("m", [1, 2, 3, 8])
def test_locality_keys_match_bitwise_restatement(m):
    from graph_learning_b200.knn import locality_keys
    rng = np.random.default_rng(m)
    x = rng.standard_normal((m, 500)) * rng.uniform(0.1, 5.0, size=(m, 1))
    got = locality_keys(x)
    assert got.dtype == torch.int32 and got.shape == (500,)
    want = orc.morton_keys(x)
    assert np.array_equal(got.numpy(), want)
    assert int(got.min()) >= 0 and int(got.max()) < 2 ** 30
    # tensors in, same keys
    assert torch.equal(locality_keys(torch.as_tensor(x)), got)


def test_locality_keys_constant_feature_and_errors():
    from graph_learning_b200.knn import locality_keys
    x = np.stack([np.linspace(0, 1, 64), np.full(64, 3.0)])
    keys = locality_keys(x).numpy()
    assert np.array_equal(keys, orc.morton_keys(x))
    assert np.all(np.diff(keys) >= 0)            # one varying coordinate: keys follow it
    with pytest.raises(ValueError):
        locality_keys(np.zeros(5))


def test_locality_keys_are_local():
    """Points adjacent in key order are close in space: the mean distance between consecutive points of
    the arrangement is far below that of a random order."""
    from graph_learning_b200.knn import locality_keys
    rng = np.random.default_rng(0)
    pts = rng.uniform(size=(4000, 3))
    order = np.argsort(locality_keys(pts.T).numpy(), kind="stable")
    step = np.linalg.norm(np.diff(pts[order], axis=0), axis=1).mean()
    base = np.linalg.norm(np.diff(pts, axis=0), axis=1).mean()
    assert step < 0.25 * base


@pytest.mark.parametrize("window", [-1, 0, 32, 64, 1000])
def test_row_order_properties(window):
    n = 700
    pts, ei, ej, _ = _case(n, 8)
    keys = orc.morton_keys(pts.T)
    deg = np.bincount(ei, minlength=n) + np.bincount(ej, minlength=n)
    for key in (None, keys):
        order = orc.knn_row_order(ei, ej, n, key, window)
        assert np.array_equal(np.sort(order), np.arange(n))
        arr = np.arange(n) if key is None else np.argsort(key, kind="stable")
        if window < 0:
            assert np.array_equal(order, arr)
            continue
        w = window if window > 0 else n
        for b in range(0, n, w):
            blk = order[b:b + w]
            assert np.array_equal(np.sort(blk), np.sort(arr[b:b + w]))       # windows keep their nodes
            assert np.all(np.diff(deg[blk]) <= 0)                            # decreasing degree inside
    glob = orc.knn_padded_slots(ei, ej, n, orc.knn_row_order(ei, ej, n, None, 0))
    here = orc.knn_padded_slots(ei, ej, n, orc.knn_row_order(ei, ej, n, keys, window))
    assert 2 * len(ei) <= glob <= here        # the global sort has the least padding of all orders


# ---- GPU (gated, see the module docstring) ----------------------------------------------------------
def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.gpu
@unvalidated
@pytest.mark.parametrize("with_key", [False, True])
@pytest.mark.parametrize("window", [-1, 0, 32, 256, 4096])
def test_ordered_solve_equals_oracle(window, with_key):
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    from graph_learning_b200.knn import locality_keys
    n = 1000
    pts, ei, ej, z = _case(n, 10, seed=3)
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=300)
    key = locality_keys(pts.T) if with_key else None
    w, info = knn(ei, ej, z, n, maxit=300, return_info=True, order_key=key, sort_window=window)
    assert info.n_iter == io.n_iter
    assert _rel(w, wo) < 1e-9
    assert np.array_equal(w > 0, wo > 0)
    w32 = knn(ei, ej, z.astype(np.float32), n, maxit=300, order_key=key, sort_window=window)
    assert w32.dtype == np.float32 and _rel(w32, wo) < 1e-4


@pytest.mark.gpu
@unvalidated
@pytest.mark.parametrize("with_key", [False, True])
@pytest.mark.parametrize("window", [-1, 0, 32, 100, 4096])
def test_padded_slots_match_row_order_restatement(window, with_key):
    """The padding of the device layout is a fingerprint of the row order actually built."""
    import ctypes
    from graph_learning_b200 import _native as nat
    from graph_learning_b200.knn import locality_keys
    n = 900
    pts, ei, ej, z = _case(n, 7, seed=5)
    key = locality_keys(pts.T).cuda() if with_key else None
    e_i = torch.as_tensor(ei, dtype=torch.int32, device="cuda")
    e_j = torch.as_tensor(ej, dtype=torch.int32, device="cuda")
    zd = torch.as_tensor(z, dtype=torch.float64, device="cuda")
    npad = -(-n // 128) * 128
    h = ctypes.c_void_p()
    nat.check(nat.lib().gl_knn_create_ordered(ctypes.byref(h), e_i.data_ptr(), e_j.data_ptr(), zd.data_ptr(), None,
                                              len(ei), n, nat.GL_F64, 0, npad, npad,
                                              key.data_ptr() if key is not None else None, window, 0, None))
    try:
        torch.cuda.synchronize()
        order = orc.knn_row_order(ei, ej, n, key.cpu().numpy() if key is not None else None, window)
        assert nat.lib().gl_knn_local_slots(h) == 2 * len(ei)
        assert nat.lib().gl_knn_padded_slots(h) == orc.knn_padded_slots(ei, ej, n, order)
    finally:
        nat.lib().gl_knn_destroy(h)


@pytest.mark.gpu
@unvalidated
def test_estimator_with_locality_order_matches_default():
    from graph_learning_b200.knn import KNNLogModel
    rng = np.random.default_rng(11)
    x = rng.uniform(size=(3, 2000))
    a = KNNLogModel(avg_degree=6, n_neighbors=12, maxit=200).fit(x)
    b = KNNLogModel(avg_degree=6, n_neighbors=12, maxit=200, locality_order=True, sort_window=512).fit(x)
    assert np.array_equal(a.edges_, b.edges_)
    assert a.n_iter_ == b.n_iter_
    assert _rel(b.weights_, a.weights_) < 1e-5        # float32 solve; only the summation order differs
    assert b.get_params()["locality_order"] is True and b.get_params()["sort_window"] == 512


# ---- CPU: executable specification of the planned blocked layout -----------------------
@pytest.mark.parametrize("block,window", [(256, 256), (64, 64), (256, 0), (1024, 1024)])
def test_blocked_layout_invariants(block, window):
    n = 1500
    pts, ei, ej, _ = _case(n, 8, seed=2)
    order = orc.knn_row_order(ei, ej, n, orc.morton_keys(pts.T), window)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block)
    lab = lay["lab"]
    assert np.array_equal(lab[order], np.arange(n))
    # every edge exactly once in an intra list or exactly twice among the slots
    count = np.bincount(lay["intra_eid"], minlength=len(ei)) * 2 + np.bincount(lay["slot_eid"], minlength=len(ei))
    assert np.all(count == 2)
    assert np.all(np.bincount(lay["slot_eid"][lay["slot_first"]], minlength=len(ei)) + np.bincount(
        lay["intra_eid"], minlength=len(ei)) == 1)
    # intra edges: both ends in the list's block, local ids reproduce the rows
    blk = np.repeat(np.arange(lay["nblocks"]), np.diff(lay["intra_ptr"]))
    ra, rb = blk * block + lay["intra_a"], blk * block + lay["intra_b"]
    e = lay["intra_eid"]
    assert np.array_equal(np.sort(np.stack([ra, rb]), axis=0), np.sort(np.stack([lab[ei[e]], lab[ej[e]]]), axis=0))
    assert lay["intra_a"].dtype == (np.uint8 if block <= 256 else np.uint16)
    # slots: source row from slot_ptr, far end is the other end of the edge, never in the same block
    src = np.repeat(np.arange(n), np.diff(lay["slot_ptr"]))
    e = lay["slot_eid"]
    assert np.array_equal(src + lay["slot_dst"], lab[ei[e]] + lab[ej[e]])
    assert np.all(src // block != lay["slot_dst"] // block)
    assert 14.0 <= lay["bytes_per_edge"] <= 32.0
    if window == block == 256:
        assert lay["intra_fraction"] > 0.35          # small n; 0.67 at n = 2e5


@pytest.mark.parametrize("block,window,with_key", [(256, 256, True), (32, 0, True), (128, -1, False), (4096, 0, False)])
def test_blocked_schedule_equals_knn_oracle(block, window, with_key):
    n = 600
    pts, ei, ej, z = _case(n, 7, seed=9)
    order = orc.knn_row_order(ei, ej, n, orc.morton_keys(pts.T) if with_key else None, window)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block)
    for kw in ({"maxit": 1, "tol": 0.0}, {"maxit": 40, "tol": 0.0}, {"maxit": 400}):
        wo, io = orc.solve_knn(ei, ej, z, n, threshold=False, **kw)
        w, info = orc.solve_knn_blocked(ei, ej, z, n, lay, threshold=False, **kw)
        assert info.n_iter == io.n_iter and info.converged == io.converged
        assert _rel(w, wo) < 1e-12
        if io.n_iter > 1:
            assert abs(info.rel_primal - io.rel_primal) < 1e-9 * io.rel_primal
            assert abs(info.rel_dual - io.rel_dual) < 1e-9 * io.rel_dual
    # warm start and thresholding go through the same path
    init = np.random.default_rng(1).uniform(0, 0.2, size=len(ei))
    wo, _ = orc.solve_knn(ei, ej, z, n, edge_init=init, maxit=30)
    w, _ = orc.solve_knn_blocked(ei, ej, z, n, lay, edge_init=init, maxit=30)
    assert _rel(w, wo) < 1e-12 and np.array_equal(w > 0, wo > 0)


@pytest.mark.gpu
@unvalidated
@pytest.mark.parametrize("n,k,window,with_key", [(1000, 10, 256, True), (1000, 10, 0, False), (200, 5, -1, True),
                                                 (3000, 12, 256, True), (257, 6, 256, True)])
def test_blocked_layout_solve_equals_oracle(n, k, window, with_key):
    """layout="blocked": intra-block edges stored once (knn_blocked_kernel).  Same iterates as the oracle;
    the device's split into intra edges and slots matches the executable specification."""
    import ctypes
    from graph_learning_b200 import _native as nat
    from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn as knn
    from graph_learning_b200.knn import locality_keys
    pts, ei, ej, z = _case(n, k, seed=n)
    key = locality_keys(pts.T) if with_key else None
    for kw in ({"maxit": 1, "tol": 0.0}, {"maxit": 17, "tol": 0.0}, {"maxit": 300}):
        wo, io = orc.solve_knn(ei, ej, z, n, threshold=False, **kw)
        w, info = knn(ei, ej, z, n, edge_tol=-np.inf, return_info=True, order_key=key, sort_window=window,
                      layout="blocked", **kw)
        assert info.n_iter == io.n_iter
        assert _rel(w, wo) < 1e-9
    w32 = knn(ei, ej, z.astype(np.float32), n, maxit=300, order_key=key, sort_window=window, layout="blocked")
    wo, _ = orc.solve_knn(ei, ej, z, n, maxit=300)
    assert w32.dtype == np.float32 and _rel(w32, wo) < 1e-4
    # warm start goes through both the intra lists and the slots
    init = np.random.default_rng(1).uniform(0, 0.2, size=len(ei))
    wo, _ = orc.solve_knn(ei, ej, z, n, edge_init=init, maxit=25)
    w = knn(ei, ej, z, n, edge_init=init, maxit=25, order_key=key, sort_window=window, layout="blocked")
    assert _rel(w, wo) < 1e-9
    # the split itself
    order = orc.knn_row_order(ei, ej, n, key.numpy() if key is not None else None, window)
    lay = orc.knn_blocked_layout(ei, ej, n, order, 256)
    e_i = torch.as_tensor(ei, dtype=torch.int32, device="cuda")
    e_j = torch.as_tensor(ej, dtype=torch.int32, device="cuda")
    zd = torch.as_tensor(z, dtype=torch.float64, device="cuda")
    kd = key.cuda() if key is not None else None
    npad = -(-n // 128) * 128
    h = ctypes.c_void_p()
    nat.check(nat.lib().gl_knn_create_ordered(ctypes.byref(h), e_i.data_ptr(), e_j.data_ptr(), zd.data_ptr(), None,
                                              len(ei), n, nat.GL_F64, 0, npad, npad,
                                              kd.data_ptr() if kd is not None else None, window, 1, None))
    try:
        torch.cuda.synchronize()
        assert nat.lib().gl_knn_intra_edges(h) == len(lay["intra_eid"])
        assert nat.lib().gl_knn_local_slots(h) == len(lay["slot_eid"])
        assert nat.lib().gl_knn_set_fused(h, 1) == nat.GL_E_STATE          # one-launch form only
    finally:
        nat.lib().gl_knn_destroy(h)


@pytest.mark.gpu
@unvalidated
def test_blocked_layout_needs_a_single_shard():
    import ctypes
    from graph_learning_b200 import _native as nat
    ei = torch.tensor([0, 1], dtype=torch.int32, device="cuda")
    ej = torch.tensor([1, 2], dtype=torch.int32, device="cuda")
    z = torch.ones(2, dtype=torch.float32, device="cuda")
    h = ctypes.c_void_p()
    rc = nat.lib().gl_knn_create_ordered(ctypes.byref(h), ei.data_ptr(), ej.data_ptr(), z.data_ptr(), None, 2, 300,
                                         nat.GL_F32, 0, 128, 384, None, 0, 1, None)
    assert rc == nat.GL_E_STATE and not h.value
    assert nat.lib().gl_knn_create_ordered(ctypes.byref(h), ei.data_ptr(), ej.data_ptr(), z.data_ptr(), None, 2, 300,
                                           nat.GL_F32, 0, 384, 384, None, 0, 7, None) == nat.GL_E_ARG
