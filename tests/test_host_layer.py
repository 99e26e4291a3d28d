"""CPU tests of the Python host layer (no compute): the drop-in surface keeps the reference's names,
constructor arguments and error behaviour, and the product path FAILS LOUDLY without a CUDA device --
there is no CPU fallback and nothing under graph_learning_b200/ imports oracle/."""
import inspect
import os
import re

import numpy as np
import pytest

torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_no_gpu = pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour of a box without a GPU")


@pytest.fixture(scope="module")
def sl():
    import graph_learning_b200.smooth_learning as m
    return m


def test_ndim_errors_come_before_any_device_use(sl):
    # smooth_learning.py:130-131
    with pytest.raises(ValueError, match="Distances must be vector or matrix"):
        sl.gsp_learn_graph_log_degrees(np.zeros((2, 3, 3)), 1, 1)
    with pytest.raises(ValueError, match="Distances must be vector or matrix"):
        sl.gsp_learn_graph_log_degrees(np.float64(1.0), 1, 1)


def test_edge_count_must_be_triangular(sl):
    assert sl._n_from_edges(4950) == 100 and sl._n_from_edges(1) == 2
    with pytest.raises(ValueError, match="triangular"):
        sl._n_from_edges(4951)


def test_method_resolution(sl):
    assert sl._resolve_method("auto", torch.float32, sl.GRAM_MIN_SAMPLES) == "gram16"
    assert sl._resolve_method("auto", torch.float32, sl.GRAM_MIN_SAMPLES - 1) == "direct"
    assert sl._resolve_method("auto", torch.float64, 10 ** 6) == "direct"
    with pytest.raises(ValueError, match="method must be"):
        sl._resolve_method("fast", torch.float32, 10)
    with pytest.raises(ValueError):
        sl._resolve_method("gram", torch.float64, 1000)       # tensor-core form is float32 only
    with pytest.raises(ValueError, match="dtype must be"):
        sl._torch_dtype("float16")


@needs_no_gpu
def test_no_cpu_fallback(sl):
    x = np.random.default_rng(0).standard_normal((20, 6))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sl.LogModel(avg_degree=2).fit(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sl.gsp_learn_graph_log_degrees(np.ones(15), 1, 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sl.pdist_sq(x)
    from graph_learning_b200.clustering import GLMM, KGraphs
    from graph_learning_b200.clustering.assign import QuadForm
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        QuadForm(np.eye(3)[None])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        KGraphs(n_clusters=2, init_params="random", random_state=0).fit(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        GLMM(n_components=2, init_params="random", random_state=0).fit(x)
    from graph_learning_b200.knn import KNNLogModel
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        KNNLogModel(avg_degree=2, n_neighbors=3).fit(x)


def test_product_never_imports_the_oracle():
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|^\s*from\s+\.+oracle\b", re.M)
    for base, _dirs, files in os.walk(os.path.join(ROOT, "graph_learning_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(base, f), encoding="utf-8").read()
                assert not pat.search(src), f"{f} imports the oracle"
                assert "/root/reference" not in src, f"{f} reads the reference tree"


def test_estimators_keep_the_reference_constructor_surface(sl):
    """Same parameter names and defaults as the reference classes (smooth_learning.py:257-275,
    kgraphs.py:20-60, glmm.py:88-135, temporal/smooth.py:12-40); the extras (``dtype``, and for LogModel
    ``distributed`` / ``weights_on``) are keyword-only and default to the reference behaviour."""
    from sklearn.base import BaseEstimator, clone
    from sklearn.mixture._base import BaseMixture

    from graph_learning_b200.clustering import GLMM, KGraphs, KGraphsV2
    from graph_learning_b200.temporal import WindowLogModel

    def params(cls):
        return {k: v.default for k, v in inspect.signature(cls.__init__).parameters.items() if k != "self"}

    p = params(sl.LogModel)
    extras = {"dtype": "float64", "distributed": False, "weights_on": "all"}
    assert [k for k in p if k not in extras] == ["avg_degree", "edge_init", "maxit", "tol", "step_size", "edge_tol"]
    assert {k: p[k] for k in extras} == extras
    sig = inspect.signature(sl.LogModel.__init__).parameters
    assert all(sig[k].kind is inspect.Parameter.KEYWORD_ONLY for k in extras)
    assert (p["maxit"], p["tol"], p["step_size"], p["edge_tol"]) == (1000, 1e-5, 0.5, 1e-3)
    p = params(KGraphs)
    assert [k for k in p if k != "dtype"] == ["n_clusters", "avg_degree", "max_iter", "n_init", "init_params", "delta",
                                              "random_state"]
    assert (p["n_clusters"], p["avg_degree"], p["max_iter"], p["delta"], p["init_params"]) == (1, 2, 100, 1, "kmeans")
    assert "theta" in params(KGraphsV2)
    p = params(GLMM)
    assert (p["n_components"], p["avg_degree"], p["regul"], p["delta"], p["tol"]) == (1, 2, 0.15, 2, 1e-3)
    assert issubclass(GLMM, BaseMixture) and issubclass(sl.LogModel, BaseEstimator)
    assert issubclass(WindowLogModel, sl.LogModel)
    for est in (sl.LogModel(avg_degree=3, tol=1e-4), KGraphs(n_clusters=3), GLMM(n_components=2, regul=0.3)):
        c = clone(est)
        assert c.get_params() == est.get_params()


def test_kgraphs_rejects_unknown_init():
    from graph_learning_b200.clustering.utils import init_labels, one_hot
    from sklearn.utils import check_random_state
    with pytest.raises(ValueError, match="Invalid init_params"):
        init_labels(np.zeros((4, 2)), 2, "nope", check_random_state(0))
    lab = init_labels(np.zeros((10, 2)), 3, "random", check_random_state(0))
    assert sorted(np.bincount(lab, minlength=3)) == [3, 3, 4]
    assert np.array_equal(one_hot(np.array([0, 2, 1]), 3), np.eye(3)[[0, 2, 1]])
