"""GPU parity tests of the assignment-step kernel (csrc/quadform.cu)
through the C-ABI: against golden outputs of the unmodified reference (tests/golden/assign_n25.npz:
kgraphs.py:87-88, :203-205, glmm.py:238-279) and against the oracle on seeded inputs."""
import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def asg():
    from graph_learning_b200.clustering import assign
    return assign


@pytest.fixture(scope="module")
def nat():
    from graph_learning_b200 import _native
    return _native


def sparse_laplacians(rng, K, n, deg=6):
    """Random sparse graph Laplacians (about ``deg`` edges per node)."""
    e = n * (n - 1) // 2
    w = rng.uniform(0.2, 2.0, size=(K, e)) * (rng.uniform(size=(K, e)) < deg / (n - 1))
    return orc.laplacian_from_packed(w)


def test_golden_smoothness_and_labels(asg, golden):
    g = golden("assign_n25")
    with asg.QuadForm(g["kgraphs_laplacians"]) as qf:
        q, lab = qf.apply(g["x"], labels=True)
    assert rel(q.cpu().numpy(), g["smoothness"]) < 1e-13
    assert np.array_equal(lab.cpu().numpy(), g["labels"])
    assert lab.dtype == torch.int32


def test_golden_centered_and_dense(asg, golden):
    g = golden("assign_n25")
    got = asg.smoothness(g["x"], g["kgraphs2_laplacians"], g["kgraphs2_means"])
    assert rel(got.cpu().numpy(), g["centered_smoothness"]) < 1e-13
    # dense, non-symmetric, non-Laplacian matrices: the entry form is exact for any square matrix
    with asg.QuadForm(g["dense"]) as qf:
        assert qf.nnz == 2 * (25 + 25 * 24 // 2)
        assert rel(qf.apply(g["x"]).cpu().numpy(), g["smoothness_dense"]) < 1e-13


def test_golden_glmm_log_prob(golden):
    from graph_learning_b200.clustering import GLMM
    g = golden("assign_n25")
    mdl = GLMM(n_components=3, avg_degree=3)
    mdl.laplacians_, mdl.means_, mdl.n_nodes_ = g["glmm_laplacians"], g["glmm_means"], 25
    assert rel(mdl._estimate_log_prob(g["x"]), g["log_prob"]) < 1e-11


def test_glmm_log_prob_device_eigh_matches_oracle():
    """n above the host-LAPACK threshold: cuSOLVER eigh + quadratic-form identity vs the oracle."""
    from graph_learning_b200.clustering import GLMM
    from graph_learning_b200.clustering.glmm import EIGH_DEVICE_MIN_NODES
    rng = np.random.default_rng(3)
    K, n, m = 2, EIGH_DEVICE_MIN_NODES + 44, 77
    laps = sparse_laplacians(rng, K, n, deg=8)        # connected with overwhelming probability
    means = rng.standard_normal((K, n))
    x = rng.standard_normal((m, n))
    mdl = GLMM(n_components=K, regul=0.15)
    mdl.laplacians_, mdl.means_, mdl.n_nodes_ = laps, means, n
    ref = orc.glmm_log_prob(x, laps, means, 0.15)
    got = mdl._estimate_log_prob(x)
    data_ref = ref - ref.max()          # logdet may overflow identically in both; compare finite parts
    if np.all(np.isfinite(ref)):
        assert rel(got, ref) < 1e-10
    else:
        assert np.array_equal(np.isfinite(got), np.isfinite(data_ref))


@pytest.mark.parametrize("dtype,tol", [("float64", 1e-12), ("float32", 2e-5)])
@pytest.mark.parametrize("K,n,m", [(1, 2, 1), (3, 33, 100), (8, 300, 1000), (2, 1000, 257), (2, 1800, 40),
                                   (2, 3100, 70), (1, 3002, 9)])
def test_matches_oracle(asg, dtype, tol, K, n, m):
    """Sizes walk the samples-per-CTA ladder (32 / 16 / 8 / 4 by shared-memory footprint) and ragged
    last blocks; shift and dot vectors on."""
    rng = np.random.default_rng(K * 1000 + n + m)
    laps = sparse_laplacians(rng, K, n) if n > 2 else np.array([[[1.0, -1.0], [-1.0, 1.0]]])
    means = rng.standard_normal((K, n))
    vec = rng.standard_normal((K, n))
    x = rng.standard_normal((m, n))
    with asg.QuadForm(laps, diag_add=0.3, dtype=dtype) as qf:
        q, d, lab = qf.apply(x, shift=means, vectors=vec, labels=True)
        q0 = qf.apply(x)
    q, d, lab, q0 = (t.cpu().numpy() for t in (q, d, lab, q0))
    ref = orc.smoothness(x, laps + 0.3 * np.eye(n), means)
    assert rel(q, ref) < tol
    assert rel(q0, orc.smoothness(x, laps + 0.3 * np.eye(n))) < tol
    dref = np.stack([(x - means[k]) @ vec[k] for k in range(K)], axis=1)
    assert np.linalg.norm(d - dref) / np.linalg.norm(dref) < tol
    assert np.array_equal(lab, np.argmin(q, axis=1))          # the argmin of the stored values


def test_too_many_nodes_for_shared_memory_raises(asg):
    with asg.QuadForm(torch.zeros((1, 8000, 8000), dtype=torch.float64, device="cuda"), dtype="float64") as qf:
        assert qf.nnz == 8000
        with pytest.raises(ValueError, match="does not fit"):
            qf.apply(torch.zeros((4, 8000), dtype=torch.float64, device="cuda"))


def test_argmin_ties_take_first(asg):
    lap = orc.laplacian_from_packed(np.ones((1, 10)))
    laps = np.concatenate([2 * lap, lap, lap, 3 * lap])
    x = np.random.default_rng(0).standard_normal((40, 5))
    _, lab = asg.QuadForm(laps).apply(x, labels=True)
    assert np.all(lab.cpu().numpy() == 1)


def test_strided_rows_and_empty(asg):
    rng = np.random.default_rng(1)
    laps = sparse_laplacians(rng, 2, 40)
    x = rng.standard_normal((50, 40))
    with asg.QuadForm(laps) as qf:
        assert qf.apply(x[:0]).shape == (0, 2)
        with pytest.raises(ValueError, match="n_samples"):
            qf.apply(x[:, :39])
        with pytest.raises(ValueError, match="shift"):
            qf.apply(x, shift=np.zeros((3, 40)))
    with pytest.raises(ValueError, match="K, n, n"):
        asg.QuadForm(np.zeros((2, 3, 4)))
    qf = asg.QuadForm(laps)
    qf.close()
    with pytest.raises(RuntimeError, match="closed"):
        qf.apply(x)


def test_c_abi_argument_errors(nat):
    import ctypes
    h = ctypes.c_void_p()
    assert nat.lib().gl_quadform_create(ctypes.byref(h), None, 0, 5, 0, 0.0, None) == nat.GL_E_SHAPE
    assert nat.lib().gl_quadform_create(ctypes.byref(h), None, 1, 5, 7, 0.0, None) == nat.GL_E_ARG
    assert nat.lib().gl_quadform_apply(None, None, 1, 1, None, None, None, None, None, None) == nat.GL_E_ARG
    assert nat.lib().gl_quadform_nnz(None) == -1
    assert nat.lib().gl_quadform_destroy(None) == nat.GL_OK


def test_kgraphs_device_loop_matches_host_formulation():
    """One outer iteration of KGraphs on the device equals the reference formulation evaluated with
    the oracle on the same Laplacians: labels, score."""
    from graph_learning_b200.clustering import KGraphs
    rng = np.random.default_rng(11)
    n, m, K = 30, 400, 3
    x = rng.standard_normal((m, n)) * rng.uniform(0.5, 2.0, size=(1, n))
    kg = KGraphs(n_clusters=K, avg_degree=4, max_iter=1, init_params="random", random_state=0)
    kg.fit(x)
    sm = orc.smoothness(x, kg.laplacians_)
    assert np.array_equal(kg.labels_, np.argmin(sm, axis=1))
    assert abs(kg.score_ - sm.min(axis=1).sum()) < 1e-9 * abs(kg.score_)
    assert rel(kg._smoothness(x), sm) < 1e-13
