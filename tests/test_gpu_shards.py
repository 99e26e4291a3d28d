"""Single-GPU tests of the pieces the distributed path is made of, with several tile shards emulated
on one device: per-shard theta candidates merged into the full statistic (gl_theta_partial), rows of
any length through the chunked row select, shard loads from the shard's slice of the packed vector
(gl_shard_rows / gl_solver_load_rows), and the resident LogModel.fit route (no packed Z, no eager
scaling pass).  The 2-GPU run of the same code is tests/test_gpu_multi.py.
"""
import ctypes

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def sl():
    import graph_learning_b200.smooth_learning as m
    return m


@pytest.fixture(scope="module")
def nat():
    from graph_learning_b200 import _native
    return _native


def _shards(nat, n, world):
    out = []
    for r in range(world):
        b, e = ctypes.c_int(), ctypes.c_int()
        nat.check(nat.lib().gl_shard_tiles(n, r, world, ctypes.byref(b), ctypes.byref(e)))
        out.append((b.value, e.value))
    return out


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n,world,k", [(300, 3, 6), (129, 2, 2), (700, 8, 10), (50, 4, 3)])
def test_theta_partial_merge_equals_full(sl, nat, n, world, k, dtype):
    lib = nat.lib()
    rng = np.random.default_rng(n + k)
    pts = rng.uniform(size=(n, 3))
    z_np = orc.to_packed(((pts[:, None] - pts[None]) ** 2).sum(-1)).astype(dtype)
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    gdt = nat.GL_F64 if dtype == np.float64 else nat.GL_F32
    z = torch.as_tensor(z_np, device="cuda")
    tt = nat.TILE * nat.TILE
    nt = lib.gl_num_tiles(n)
    zt = torch.empty(nt * tt, dtype=tdt, device="cuda")
    nat.check(lib.gl_pack_to_tiles(z.data_ptr(), zt.data_ptr(), n, 1, gdt, 0, nt, None))
    th_full = ctypes.c_double()
    nat.check(lib.gl_theta(zt.data_ptr(), nat.LAYOUT_TILED, n, k, gdt, ctypes.byref(th_full), None))
    want = orc.theta_for_degree(orc.to_square(z_np.astype(np.float64)), k)
    assert abs(th_full.value - want) <= 1e-12 * want
    cands = []
    for b, e in _shards(nat, n, world):
        c = torch.full((n, k + 1), -1.0, dtype=tdt, device="cuda")
        shard = zt[b * tt:e * tt].clone()          # the shard alone, as a rank holds it
        nat.check(lib.gl_theta_partial(shard.data_ptr(), n, b, e, k, gdt, c.data_ptr(), None))
        cands.append(c)
    merged = torch.cat(cands, dim=1).contiguous()
    th = ctypes.c_double()
    nat.check(lib.gl_theta_dense(merged.data_ptr(), n, merged.shape[1], merged.shape[1], k, gdt, ctypes.byref(th), None))
    assert th.value == th_full.value          # same values selected, same order of operations
    # the helper used by LogModel.fit on one GPU
    assert sl._theta_tiles(zt, n, k) == th_full.value


def test_theta_rows_longer_than_shared_memory(nat):
    """Rows stream through shared memory in chunks: 120,000 float32 columns (a one-chunk row holds
    ~51,000) and 60,000 float64 columns, against NumPy's sort (smooth_learning.py:202-219)."""
    lib = nat.lib()
    for dtype, cols in ((np.float32, 120_000), (np.float64, 60_000)):
        rng = np.random.default_rng(cols)
        rows, k = 5, 7
        mat = rng.uniform(0.5, 2.0, size=(rows, cols)).astype(dtype)
        mat[:, 0] = 0.0                                    # the zero "diagonal" of the reference's rows
        mat[2, cols - 3:] = [0.01, 0.02, 0.03]             # minima in the last chunk
        mat[3, 51_000:51_004] = 0.04                       # ties across a chunk boundary
        t = torch.as_tensor(mat, device="cuda")
        th = ctypes.c_double()
        gdt = nat.GL_F64 if dtype == np.float64 else nat.GL_F32
        nat.check(lib.gl_theta_dense(t.data_ptr(), rows, cols, cols, k, gdt, ctypes.byref(th), None))
        srt = np.sort(mat.astype(np.float64), axis=1)
        b = srt[:, :k].sum(1)
        lo = 1 / np.sqrt(k * srt[:, k] ** 2 - b * srt[:, k])
        hi = 1 / np.sqrt(k * srt[:, k - 1] ** 2 - b * srt[:, k - 1])
        want = np.sqrt(lo.mean() * hi.mean())
        assert abs(th.value - want) <= 1e-12 * want


@pytest.mark.parametrize("n,world", [(300, 2), (513, 3), (129, 2), (1000, 8)])
def test_shard_rows_and_load_rows(sl, nat, n, world):
    """A rank uploads only [edge_begin, edge_end) of the packed vector; the resident tiles equal those of
    a load from the full vector, and the slices of all ranks cover what the ranks need."""
    rng = np.random.default_rng(n)
    E = n * (n - 1) // 2
    z_np = rng.uniform(0.1, 1.0, size=E)
    w_np = rng.uniform(0.0, 1.0, size=E)
    z, w = torch.as_tensor(z_np, device="cuda"), torch.as_tensor(w_np, device="cuda")
    for b, e in _shards(nat, n, world):
        full = sl.Solver(n, 1, torch.float64, tile_range=(b, e))
        full.load(z[None, :].contiguous(), [2.5], w[None, :].contiguous())
        part = sl.Solver(n, 1, torch.float64, tile_range=(b, e))
        r0, r1, e0, e1 = part.shard_rows()
        assert 0 <= r0 <= r1 <= n and 0 <= e0 <= e1 <= E
        if e > b:
            assert e0 == r0 * (n - 1) - r0 * (r0 - 1) // 2      # packed offset of row r0 (operators.py:39-41)
        part.load_rows(z[e0:e1].clone(), 2.5, w[e0:e1].clone())
        assert torch.equal(full.z_view(0), part.z_view(0))
        assert torch.equal(full.w_view(0), part.w_view(0))
        full.close()
        part.close()


def test_batch_cannot_be_tile_sharded(sl, nat):
    """K > 1 shares one spectral stop rule over all edges (smooth_learning.py:158-159): a partial tile
    range is refused instead of silently deciding from shard-local norms."""
    with pytest.raises(nat.GLError, match="tile-sharded"):
        sl.Solver(300, 2, torch.float64, tile_range=(0, 2))
    sl.Solver(300, 2, torch.float64).close()


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_logmodel_fit_resident_route(sl, dtype):
    """LogModel.fit builds Z straight into the solver's tiles, scales in place and solves; same theta and
    weights as packed Z -> get_theta -> gsp_learn_graph_log_degrees(z * theta) (:277-300)."""
    from graph_learning_b200.synthetic import rgg_signals_exact
    x = rgg_signals_exact(150, 400, seed=3)
    mdl = sl.LogModel(avg_degree=5, maxit=300, dtype=dtype).fit(x)
    z = orc.pdist_sq(x)
    theta = orc.theta_for_degree(orc.to_square(z), 5)
    w_ref, info = orc.solve_two_pass(z * theta, maxit=300)
    tol = 1e-9 if dtype == "float64" else 1e-4
    assert abs(mdl.theta_ - theta) <= (1e-12 if dtype == "float64" else 1e-5) * theta
    assert rel(mdl.weights_, orc.to_square(w_ref)) < tol
    assert mdl.weights_.shape == (150, 150) and isinstance(mdl.weights_, np.ndarray)
    if dtype == "float64":
        assert mdl.n_iter_ == info.n_iter
    # warm start through the estimator (edge_init is packed (E,), :261)
    w0 = np.abs(np.random.default_rng(0).standard_normal(z.size)) * 0.01
    mdl2 = sl.LogModel(avg_degree=5, maxit=20, dtype=dtype, edge_init=w0).fit(x)
    w_ref2, _ = orc.solve_two_pass(z * theta, maxit=20, edge_init=w0.copy())
    assert rel(mdl2.weights_, orc.to_square(w_ref2)) < tol
    # out= (extra): preallocated host result
    out = np.empty((150, 150), dtype=np.float64 if dtype == "float64" else np.float32)
    mdl3 = sl.LogModel(avg_degree=5, maxit=300, dtype=dtype).fit(x, out=out)
    assert mdl3.weights_ is out and rel(out, mdl.weights_) < 1e-6   # two runs: atomic summation order differs
    with pytest.raises(ValueError, match="edge_init"):
        sl.LogModel(avg_degree=5, edge_init=np.zeros(7)).fit(x)


def test_solver_function_z_scale(golden, sl):
    """z_scale multiplies theta into Z on load (no eager z * theta pass by the caller)."""
    g = golden("batched_n30")
    w = sl.gsp_learn_graph_log_degrees(g["z"], 1, 1, z_scale=g["theta"].ravel())
    assert rel(w, g["w"]) < 1e-9
    g1 = golden("solver_n100")
    w1 = sl.gsp_learn_graph_log_degrees(g1["z"], 1, 1, z_scale=float(g1["theta"]))
    assert rel(w1, g1["w_final"]) < 1e-9


def test_logmodel_without_avg_degree_and_workspace_cache(sl):
    """avg_degree=None: theta_ = 1, distances unscaled (:279-280).  Alternating problem sizes replace the
    cached single-GPU workspace; release_solver_cache frees it."""
    from graph_learning_b200.synthetic import rgg_signals_exact
    for n, m in ((60, 80), (140, 90), (60, 80)):
        x = rgg_signals_exact(n, m, seed=n)
        mdl = sl.LogModel(maxit=40).fit(x)
        assert mdl.theta_ == 1
        w_ref, info = orc.solve_two_pass(orc.pdist_sq(x), maxit=40)
        assert mdl.n_iter_ == info.n_iter
        assert rel(mdl.weights_, orc.to_square(w_ref)) < 1e-9
        z = orc.pdist_sq(x)
        w = sl.gsp_learn_graph_log_degrees(z, 1, 1, maxit=40)          # the function form shares the cache
        assert rel(w, orc.to_square(w_ref)) < 1e-9
        assert len(sl._solver_cache) == 1
    sl.release_solver_cache()
    assert len(sl._solver_cache) == 0
    # and the next call simply builds a new workspace
    w = sl.gsp_learn_graph_log_degrees(z, 1, 1, maxit=40)
    assert rel(w, orc.to_square(w_ref)) < 1e-9
