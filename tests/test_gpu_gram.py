"""GPU parity tests of the tensor-core distance kernel (K1, gl_pdist_sq_gram): the tcgen05 Gram
GEMM on error-compensated bf16 splits against the float64 oracle (`pdist(x.T)**2`,
smooth_learning.py:278), the GLMM weighted/centred form (glmm.py:61-68), both output layouts, tile
ranges (what a GPU shard builds), every tile variant, and LogModel.fit end to end within the
north_star float32 tolerance (1e-4 relative on the weights, identical support outside a band).
"""
import ctypes

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

TOL3 = 1e-6   # 3-term splits (~24 bits): relative Frobenius error of Z against float64
TOL2 = 4e-6   # 2-term splits (~16 bits)


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


@pytest.fixture(scope="module")
def smooth300():
    from graph_learning_b200.synthetic import rgg_signals_exact
    return rgg_signals_exact(300, 1000, seed=0)


@pytest.mark.parametrize("m,n", [(1, 2), (5, 3), (64, 130), (257, 129), (300, 515), (1000, 100)])
@pytest.mark.parametrize("method,tol", [("gram", TOL3), ("gram2", TOL2), ("gram16", TOL3)])
def test_gram_shapes_vs_oracle(sl, m, n, method, tol):
    """Ragged sizes: n and m off the 128 / 64 grid, a single sample, the smallest graph."""
    rng = np.random.default_rng(m * 1000 + n)
    x = rng.standard_normal((m, n)).astype(np.float32)
    z = sl.pdist_sq(x, method=method).cpu().numpy()
    assert z.dtype == np.float32 and z.shape == (n * (n - 1) // 2,)
    assert (z >= 0).all()
    assert rel(z, orc.pdist_sq(x.astype(np.float64))) < tol


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("method", ["gram", "gram2"])
def test_gram_variants_on_smooth_signals(sl, nat, smooth300, method, variant):
    """Every kernel variant (split / single accumulator, unit width 128/256, k-block 32/64 =
    SWIZZLE_64B/128B, and the CTA-pair cta_group::2 kernel) on signals where the Gram form cancels
    (smooth on the graph)."""
    x = smooth300.astype(np.float32)
    ref = orc.pdist_sq(x.astype(np.float64))
    if method == "gram2" and variant == 6:
        with pytest.raises(nat.GLError, match="terms = 3"):
            sl.pdist_sq(x, method=method, variant=variant)
        return
    if method == "gram" and variant == 4:
        with pytest.raises(nat.GLError, match="does not fit"):
            sl.pdist_sq(x, method=method, variant=variant)
        return
    z = sl.pdist_sq(x, method=method, variant=variant).cpu().numpy()
    assert rel(z, ref) < (TOL3 if method == "gram" else TOL2)
    # chunked accumulation is what keeps the tensor-core accumulator honest: one 64-sample chunk
    # per fold is at least as accurate as the default
    z64 = sl.pdist_sq(x, method=method, variant=variant, chunk=64).cpu().numpy()
    assert rel(z64, ref) < (TOL3 if method == "gram" else TOL2)


def test_gram_weighted_centred(sl):
    """GLMM M-step form (glmm.py:61-68): y = resp[:, None] * (x - mean)."""
    rng = np.random.default_rng(3)
    x = rng.standard_normal((500, 140)).astype(np.float32)
    resp = rng.uniform(0.1, 1.0, size=500).astype(np.float32)
    mean = rng.standard_normal(140).astype(np.float32)
    y = resp.astype(np.float64)[:, None] * (x.astype(np.float64) - mean.astype(np.float64))
    z = sl.pdist_sq(x, row_scale=resp, col_shift=mean, method="gram").cpu().numpy()
    assert rel(z, orc.pdist_sq(y)) < TOL3


def test_gram_strided_input(sl):
    """Leading dimension larger than n (a column slice of a wider matrix)."""
    rng = np.random.default_rng(5)
    wide = torch.from_numpy(rng.standard_normal((300, 400)).astype(np.float32)).cuda()
    view = wide[:, 17:217]
    assert view.stride(0) == 400
    from graph_learning_b200 import _native as nat
    n, m = 200, 300
    out = torch.empty((n * (n - 1) // 2,), dtype=torch.float32, device="cuda")
    ws = torch.empty((nat.lib().gl_gram_workspace_bytes(m, n, 3),), dtype=torch.uint8, device="cuda")
    nat.check(nat.lib().gl_pdist_sq_gram(view.data_ptr(), m, n, view.stride(0), None, None, out.data_ptr(),
                                         nat.LAYOUT_PACKED, 0, 0, 3, 0, 0, ws.data_ptr(), ws.numel(), None))
    assert rel(out.cpu().numpy(), orc.pdist_sq(view.cpu().numpy().astype(np.float64))) < TOL3


@pytest.mark.parametrize("variant", [0, 2, 6])
def test_gram_tiled_ranges_match_direct(sl, nat, variant):
    """Tiled output for a shard's tile range: same slots as the exact kernel, zeros in the invalid
    slots (i >= j, j >= n), nothing written outside the range."""
    rng = np.random.default_rng(11)
    n, m = 700, 320
    x = torch.from_numpy(rng.standard_normal((m, n)).astype(np.float32)).cuda()
    nt = -(-n // 128)
    total = nt * (nt + 1) // 2
    full = sl.pdist_sq_tiles(x, 0, total, method="direct")
    for b, e in [(0, total), (3, 11), (7, 8), (total - 2, total), (5, 5)]:
        guard = 3
        buf = torch.full(((e - b + 2 * guard) * 128 * 128,), -7.0, dtype=torch.float32, device="cuda")
        out = buf[guard * 128 * 128:(guard + e - b) * 128 * 128]
        sl.pdist_sq_tiles(x, b, e, out=out, method="gram", variant=variant)
        torch.cuda.synchronize()
        ref = full[b * 128 * 128:e * 128 * 128]
        assert (buf[:guard * 128 * 128] == -7.0).all() and (buf[(guard + e - b) * 128 * 128:] == -7.0).all()
        if e > b:
            assert rel(out.cpu().numpy(), ref.cpu().numpy()) < TOL3
            assert ((out == 0) == (ref == 0)).float().mean() > 0.9999  # invalid slots are exactly 0
            assert (out[ref == 0] == 0).all()


def test_gram_duplicate_columns_and_translation(sl):
    """Properties: identical columns are at distance ~0 (never negative), and a common per-sample
    offset (which the Gram form sees as a large shared component) does not move the distances
    beyond the float32 budget."""
    rng = np.random.default_rng(2)
    x = rng.standard_normal((600, 150)).astype(np.float32)
    x[:, 77] = x[:, 3]
    z = sl.pdist_sq(x, method="gram").cpu().numpy()
    zsq = orc.to_square(z)
    scale = (x.astype(np.float64) ** 2).sum(0).mean()
    assert 0 <= zsq[3, 77] < 2e-7 * scale
    shifted = x + rng.standard_normal((600, 1)).astype(np.float32)
    z2 = sl.pdist_sq(shifted, method="gram").cpu().numpy()
    assert rel(z2, orc.pdist_sq(shifted.astype(np.float64))) < TOL3


def test_precise_variant_beats_the_plain_accumulator(sl, smooth300):
    """The default (exact leading slice + split accumulators + float64 folds) is at least as close to
    float64 as the exact float32 kernel on smooth signals; the single-accumulator variant shows the
    truncation bias the default removes."""
    x = smooth300.astype(np.float32)
    ref = orc.pdist_sq(x.astype(np.float64))
    scale = (x.astype(np.float64) ** 2).sum(0).mean()
    err = lambda z: np.abs(z.cpu().numpy().astype(np.float64) - ref).max() / scale  # noqa: E731
    e_precise = err(sl.pdist_sq(x, method="gram"))
    e_direct = err(sl.pdist_sq(x, method="direct"))
    e_plain = err(sl.pdist_sq(x, method="gram", variant=2))
    assert e_precise < 5e-7
    assert e_precise <= 2 * e_direct
    assert e_plain < 2e-5


def test_gram_config2_size_against_direct(sl):
    """Config 2 (n=2,000, m=10,000): tensor-core form against the exact float32 kernel and a
    float64 spot check of rows."""
    from graph_learning_b200.synthetic import rgg_signals_filtered
    x = rgg_signals_filtered(2000, 10000, seed=0, device="cuda", dtype=torch.float32)
    zg = sl.pdist_sq(x, method="gram")
    zd = sl.pdist_sq(x, method="direct")
    assert float((zg.double() - zd.double()).norm() / zd.double().norm()) < 1e-6
    z64 = sl.pdist_sq(x.double(), method="direct")
    assert float((zg.double() - z64).norm() / z64.norm()) < 1e-6
    assert float((zg.double() - z64).norm()) <= 3 * float((zd.double() - z64).norm()) + 1e-12


def test_auto_picks_gram_for_float32(sl, smooth300):
    assert sl._resolve_method("auto", torch.float32, 1000) == "gram16"
    assert sl._resolve_method("auto", torch.float32, 10) == "direct"
    assert sl._resolve_method("auto", torch.float64, 1000) == "direct"
    with pytest.raises(ValueError):
        sl.pdist_sq(smooth300, dtype="float64", method="gram")
    with pytest.raises(ValueError):
        sl.pdist_sq(smooth300, method="bogus")


def test_gram_workspace_errors(nat):
    lib = nat.lib()
    x = torch.zeros((64, 130), dtype=torch.float32, device="cuda")
    out = torch.empty((130 * 129 // 2,), dtype=torch.float32, device="cuda")
    ws = torch.empty((1024,), dtype=torch.uint8, device="cuda")
    rc = lib.gl_pdist_sq_gram(x.data_ptr(), 64, 130, 130, None, None, out.data_ptr(), nat.LAYOUT_PACKED, 0, 0, 3, 0, 0,
                              ws.data_ptr(), ws.numel(), None)
    assert rc == nat.GL_E_ARG and b"workspace too small" in lib.gl_last_error()
    rc = lib.gl_pdist_sq_gram(x.data_ptr(), 64, 1, 130, None, None, out.data_ptr(), nat.LAYOUT_PACKED, 0, 0, 3, 0, 0,
                              ws.data_ptr(), ws.numel(), None)
    assert rc == nat.GL_E_SHAPE
    rc = lib.gl_pdist_sq_gram(x.data_ptr(), 64, 130, 130, None, None, out.data_ptr(), nat.LAYOUT_PACKED, 0, 0, 4, 0, 0,
                              ws.data_ptr(), ws.numel(), None)
    assert rc == nat.GL_E_ARG


def test_logmodel_fit_float32_through_the_tensor_core_path(sl, smooth300):
    """End to end (north_star tolerance): LogModel(dtype=float32).fit with Z from the Gram GEMM
    against the float64 two-pass oracle (pinned to the reference's golden vectors): weights within
    1e-4 relative, same iteration count, identical support outside a 1e-5 band around edge_tol."""
    x = smooth300
    z = orc.pdist_sq(x)
    theta = orc.theta_for_degree(orc.to_square(z), 6)
    w_ref, info_ref = orc.solve_two_pass(z * theta, maxit=400)
    w_ref_sq = orc.to_square(w_ref)
    assert sl._resolve_method("auto", torch.float32, x.shape[0]) == "gram16"
    mdl = sl.LogModel(avg_degree=6, maxit=400, dtype="float32").fit(x)
    assert abs(mdl.theta_ - theta) < 1e-5 * theta
    assert mdl.n_iter_ == info_ref.n_iter
    assert rel(mdl.weights_, w_ref_sq) < 1e-4
    raw, _ = orc.solve_two_pass(z * theta, maxit=400, edge_tol=-np.inf)
    away = np.abs(orc.to_square(raw) - 1e-3) > 1e-5
    assert ((mdl.weights_ > 0) == (w_ref_sq > 0))[away].all()


# ---- fp16 two-slice form (variant 7, method "gram16"): normalised planes, four products ------------------
def test_gram16_precision_on_smooth_signals(sl, smooth300):
    """Two fp16 slices (8-bit exact leading slice + 11-bit remainder, per-node power-of-two normalisation)
    against float64 on signals where the Gram form cancels: same error budget as the 3-slice bf16 form."""
    x = smooth300.astype(np.float32)
    ref = orc.pdist_sq(x.astype(np.float64))
    scale = (x.astype(np.float64) ** 2).sum(0).mean()
    err = lambda z: np.abs(z.cpu().numpy().astype(np.float64) - ref).max() / scale  # noqa: E731
    z16 = sl.pdist_sq(x, method="gram16")
    assert rel(z16.cpu().numpy(), ref) < TOL3
    assert err(z16) < 1e-6
    assert err(z16) <= 4 * err(sl.pdist_sq(x, method="gram")) + 1e-7
    assert rel(sl.pdist_sq(x, method="gram16", chunk=64).cpu().numpy(), ref) < TOL3
    # the same scheme on CTA pairs (cta_group::2, each CTA loading half of the shared operand): identical
    # products in the same order, hence bit-identical distances
    zp = sl.pdist_sq(x, method="gram16", variant=8)
    assert torch.equal(zp, z16)


@pytest.mark.parametrize("m,n", [(64, 130), (257, 129), (300, 515), (1000, 300)])
def test_gram16_pair_variant_shapes_and_tiles(sl, m, n):
    rng = np.random.default_rng(m + n)
    x = rng.standard_normal((m, n)).astype(np.float32)
    ref = orc.pdist_sq(x.astype(np.float64))
    assert rel(sl.pdist_sq(x, method="gram16", variant=8).cpu().numpy(), ref) < TOL3
    nt = -(-n // 128)
    total = nt * (nt + 1) // 2
    xt = torch.from_numpy(x).cuda()
    full = sl.pdist_sq_tiles(xt, 0, total, method="gram16")
    for b, e in [(0, total), (1, total), (total - 1, total)]:
        if e > b:
            part = sl.pdist_sq_tiles(xt, b, e, method="gram16", variant=8)
            assert torch.equal(part, full[b * 128 * 128:e * 128 * 128])


def test_gram16_scales_offsets_and_weighted_form(sl):
    """Nodes of very different magnitude (the normalisation is per node), a node of zeros, a huge and a tiny
    overall scale (outside fp16's range in true units), and the GLMM weighted / centred form."""
    rng = np.random.default_rng(8)
    x = rng.standard_normal((400, 160)).astype(np.float32)
    x *= np.exp(rng.uniform(-6, 6, size=160)).astype(np.float32)[None, :]
    x[:, 5] = 0.0
    for factor in (1.0, 3.0e4, 1.0e-6):
        xs = (x * np.float32(factor)).astype(np.float32)
        ref = orc.pdist_sq(xs.astype(np.float64))
        z = sl.pdist_sq(xs, method="gram16").cpu().numpy()
        assert np.isfinite(z).all() and (z >= 0).all()
        assert rel(z, ref) < TOL3
    resp = rng.uniform(0.1, 1.0, size=400).astype(np.float32)
    mean = x.mean(0).astype(np.float32)
    y = resp.astype(np.float64)[:, None] * (x.astype(np.float64) - mean.astype(np.float64))
    z = sl.pdist_sq(x, row_scale=resp, col_shift=mean, method="gram16").cpu().numpy()
    assert rel(z, orc.pdist_sq(y)) < TOL3


def test_gram16_tiled_range_and_weights(sl, smooth300):
    rng = np.random.default_rng(12)
    n, m = 700, 320
    x = torch.from_numpy(rng.standard_normal((m, n)).astype(np.float32)).cuda()
    nt = -(-n // 128)
    total = nt * (nt + 1) // 2
    full = sl.pdist_sq_tiles(x, 0, total, method="direct")
    out = sl.pdist_sq_tiles(x, 3, 11, method="gram16")
    ref = full[3 * 128 * 128:11 * 128 * 128]
    assert rel(out.cpu().numpy(), ref.cpu().numpy()) < TOL3
    assert (out[ref == 0] == 0).all()
    # learned weights from gram16 distances within the float32 tolerance of the float64 solve
    xs = smooth300
    z64 = orc.pdist_sq(xs)
    theta = orc.theta_for_degree(orc.to_square(z64), 6)
    w_ref, info_ref = orc.solve_two_pass(z64 * theta, maxit=400)
    z16 = sl.pdist_sq(xs.astype(np.float32), method="gram16")
    th16 = sl._theta_packed(z16, 300, 6)
    assert abs(th16 - theta) < 1e-5 * theta
    w16, info = sl.gsp_learn_graph_log_degrees(z16, 1, 1, maxit=400, z_scale=th16, return_info=True)
    assert info.n_iter == info_ref.n_iter
    assert rel(w16.cpu().numpy(), orc.to_square(w_ref)) < 1e-4
