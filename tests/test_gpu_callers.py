"""GPU parity tests of the batched callers of the hot path: the GLMM
M-step, k-Graphs, k-Graphs V2 and WindowLogModel, against the golden outputs of the unmodified
reference (tests/golden/make_golden.py; data of the reference's tests/test_clustering.py:15-49 and
tests/test_models.py:26-28)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def test_glmm_mstep_golden(golden):
    from graph_learning_b200.clustering.glmm import _estimate_gauss_laplacian_parameters as mstep
    g = golden("clustering_n25")
    w, mu, lap = mstep(g["x"], g["resp"], 2.0, avg_degree=3)
    assert rel(w, g["mstep_weights"]) < 1e-13
    assert rel(mu, g["mstep_means"]) < 1e-13
    assert rel(lap, g["mstep_laplacians"]) < 1e-9
    _, _, lap_t = mstep(g["x"], g["resp"], 2.0, theta=0.7)
    assert rel(lap_t, g["mstep_theta_laplacians"]) < 1e-9
    _, _, lap_b = mstep(g["x"], g["resp"], 2.0, avg_degree=2, blocks=g["blocks"])
    assert rel(lap_b, g["mstep_block_laplacians"]) < 1e-9
    with pytest.raises(ValueError, match="Exactly one"):
        mstep(g["x"], g["resp"], 2.0)


def test_weighted_distances_match_reference_formula():
    """glmm.py:67-68 verbatim (NumPy, small m) vs the row_scale/col_shift kernel."""
    from graph_learning_b200.clustering.glmm import weighted_sq_dists
    from oracle import logmodel_oracle as orc
    rng = np.random.default_rng(0)
    x = rng.standard_normal((30, 12))
    resp = rng.dirichlet(np.ones(3), size=30)
    means = resp.T @ x / resp.sum(0)[:, None]
    y = resp[:, :, None] * (x[:, None, :] - means[None])
    ref = np.sum((y[..., None] - y[..., None, :]) ** 2, axis=0)
    got = weighted_sq_dists(x, resp, means).cpu().numpy()
    assert rel(got, orc.to_packed(ref)) < 1e-13


def test_glmm_fit_predict_golden(golden):
    from graph_learning_b200.clustering import GLMM
    g = golden("clustering_n25")
    mdl = GLMM(n_components=3, avg_degree=3, random_state=7, max_iter=30)
    labels = mdl.fit_predict(g["x"])
    assert np.array_equal(labels, g["glmm_labels"])
    assert rel(mdl.laplacians_, g["glmm_laplacians"]) < 1e-7
    assert rel(mdl.means_, g["glmm_means"]) < 1e-9
    assert rel(mdl.weights_, g["glmm_weights"]) < 1e-9


def test_kgraphs_golden(golden):
    from graph_learning_b200.clustering import KGraphs, KGraphsV2
    g = golden("clustering_n25")
    kg = KGraphs(n_clusters=3, avg_degree=3, random_state=7, max_iter=30)
    labels = kg.fit_predict(g["x"])
    assert np.array_equal(labels, g["kgraphs_labels"])
    assert rel(kg.laplacians_, g["kgraphs_laplacians"]) < 1e-8
    assert abs(kg.score_ - float(g["kgraphs_score"])) < 1e-8 * abs(float(g["kgraphs_score"]))
    assert kg.predict(g["x"]).shape == labels.shape
    assert np.allclose(kg.predict_proba(g["x"]).sum(1), 1.0)
    kg2 = KGraphsV2(n_clusters=3, avg_degree=3, random_state=7, max_iter=30)
    labels2 = kg2.fit_predict(g["x"])
    assert np.array_equal(labels2, g["kgraphs2_labels"])
    assert rel(kg2.laplacians_, g["kgraphs2_laplacians"]) < 1e-8
    assert rel(kg2.means_, g["kgraphs2_means"]) < 1e-9


def test_clustering_beats_gaussian_mixture(golden):
    """The reference's own acceptance bar (tests/test_clustering.py:68-133): AMI >= GaussianMixture."""
    from sklearn.metrics import adjusted_mutual_info_score as ami
    from sklearn.mixture import GaussianMixture
    from graph_learning_b200.clustering import GLMM
    g = golden("clustering_n25")
    floor = ami(g["y_true"], GaussianMixture(n_components=3, random_state=0).fit_predict(g["x"]))
    got = ami(g["y_true"], GLMM(n_components=3, avg_degree=3, random_state=7, max_iter=30).fit_predict(g["x"]))
    assert got >= floor


@pytest.mark.parametrize("ws", [5, 10])
@pytest.mark.parametrize("k", [2, 5])
def test_window_logmodel_golden(golden, ws, k):
    from graph_learning_b200.temporal import WindowLogModel
    g = golden("window_n7")
    mdl = WindowLogModel(window_size=ws, avg_degree=k, random_state=240920).fit(g["x"])
    assert rel(np.asarray(mdl.theta_), g[f"theta_ws{ws}_k{k}"]) < 1e-12
    assert mdl.weights_.shape == g[f"weights_ws{ws}_k{k}"].shape
    assert rel(mdl.weights_, g[f"weights_ws{ws}_k{k}"]) < 1e-9


def test_glmm_random_laplacian_init_follows_the_reference_sampler(golden):
    """laplacian_init="random": one default_rng(random_state) per component (glmm.py:202-207,
    sampling/graphs.py:10-14) -- an integer seed repeats the draw, a Generator advances."""
    from graph_learning_b200.clustering.glmm import _random_laplacians
    g = golden("glmm_random_init_n7")
    n, k = int(g["n"]), int(g["k"])
    assert np.allclose(_random_laplacians(n, k, int(g["int_seed"])), g["laplacians_int_seed"], rtol=1e-13, atol=0)
    gen = np.random.default_rng(int(g["generator_seed"]))
    assert np.allclose(_random_laplacians(n, k, gen), g["laplacians_generator"], rtol=1e-13, atol=0)
