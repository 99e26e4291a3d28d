"""Generate golden vectors by running the UNMODIFIED reference (LTS4/graph-learning).

Run in the build container only (the GPU box has no /root/reference):

    PYTHONPATH=/root/reference:/root/repo PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Writes tests/golden/*.npz.  Each file stores the seeded INPUTS next to the reference's
OUTPUTS so that the tests never have to regenerate inputs.  The reference is imported
read-only; nothing from it is copied into the repo.
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from graph_learn.operators import (  # noqa: E402  (reference)
    laplacian_squareform,
    laplacian_squareform_adj,
    laplacian_squareform_adj_vec,
    laplacian_squareform_vec,
    square_to_vec,
)
from graph_learn.smooth_learning import (  # noqa: E402  (reference)
    LogModel,
    get_theta,
    gsp_learn_graph_log_degrees,
    sum_squareform,
)
from graph_learn.clustering.glmm import GLMM, _estimate_gauss_laplacian_parameters  # noqa: E402
from graph_learn.clustering.kgraphs import KGraphs, KGraphsV2  # noqa: E402
from graph_learn.temporal.smooth import WindowLogModel  # noqa: E402
from graph_learn.sampling.graphs import sample_er_laplacian  # noqa: E402
from graph_learn.sampling.signals import sample_lgmrf  # noqa: E402
from scipy.spatial.distance import pdist, squareform  # noqa: E402

from graph_learning_b200.synthetic import rgg_signals_exact  # noqa: E402


def save(name, **arrays):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


class GLMMShim(GLMM):
    """scikit-learn 1.9 passes ``xp=`` to BaseMixture's private hooks; the reference overrides
    them without it (glmm.py:176,222,238,281,291).  Swallow the keyword, change nothing else."""

    def _check_parameters(self, X, xp=None):
        return super()._check_parameters(X)

    def _m_step(self, X, log_resp, xp=None):
        return super()._m_step(X, log_resp)

    def _estimate_log_prob(self, X, xp=None):
        return super()._estimate_log_prob(X)

    def _estimate_log_weights(self, xp=None):
        return super()._estimate_log_weights()

    def _set_parameters(self, params, xp=None):
        return super()._set_parameters(params)


def solver_states():
    """Raw (un-thresholded) iterates of the solver after a fixed number of iterations, and
    the converged result, config 1 (n=100, fp64)."""
    n, m, k = 100, 1000, 8
    x = rgg_signals_exact(n, m, seed=0)
    z = pdist(x.T) ** 2
    theta = get_theta(squareform(z), k)
    out = {"z": z, "theta": theta, "avg_degree": k, "n": n}
    for its in (1, 2, 3, 5, 50):
        out[f"w_raw_it{its}"] = gsp_learn_graph_log_degrees(
            z * theta, 1, 1, maxit=its, edge_tol=-np.inf
        )
    # converged, default parameters; iteration count is recovered by the oracle test
    out["w_final"] = gsp_learn_graph_log_degrees(z * theta, 1, 1)
    # non-default alpha / beta / step / tol
    out["w_alt"] = gsp_learn_graph_log_degrees(
        z * theta, 0.7, 1.9, maxit=400, tol=1e-6, step_size=0.8, edge_tol=1e-4
    )
    # warm start
    rng = np.random.default_rng(5)
    init = rng.uniform(0, 0.2, size=z.shape)
    out["edge_init"] = init
    out["w_warm"] = gsp_learn_graph_log_degrees(z * theta, 1, 1, edge_init=init.copy(), maxit=30)
    save("solver_n100", **out)


def logmodel_fit():
    """LogModel(...).fit(X): theta_ and weights_ (smooth_learning.py:286-300)."""
    for n, m, k, tag in ((60, 300, 6, "n60"), (33, 50, 4, "n33")):
        x = rgg_signals_exact(n, m, seed=11 + n)
        mdl = LogModel(avg_degree=k).fit(x)
        mdl_none = LogModel(avg_degree=None, maxit=40).fit(x)
        save(
            f"logmodel_{tag}",
            x=x,
            avg_degree=k,
            theta=mdl.theta_,
            weights=mdl.weights_,
            weights_none=mdl_none.weights_,
            pdist_sq=pdist(x.T) ** 2,
        )


def config1():
    """BASELINE.json configs[0]: LogModel on synthetic smooth signals, n = 100 nodes, m = 1,000 signals,
    fp64, with the unmodified reference.  The signals are rounded to float32 values first so that the
    float32 path sees the very same inputs (and the fixture compresses)."""
    x = rgg_signals_exact(100, 1000, seed=0).astype(np.float32).astype(np.float64)
    mdl = LogModel(avg_degree=8).fit(x)
    save("config1_n100", x=x.astype(np.float32), avg_degree=8, theta=mdl.theta_, weights=mdl.weights_)


def theta_cases():
    rng = np.random.default_rng(3)
    n = 40
    pts = rng.standard_normal((n, 5))
    zsq = squareform(pdist(pts) ** 2)
    ks = np.array([1, 2, 3, 5, 10, 20, 39])
    with np.errstate(divide="ignore", invalid="ignore"):
        thetas = np.array([get_theta(zsq, int(k)) for k in ks])
    blocks = np.repeat([0, 1, 2], [15, 12, 13])
    theta_blocks = get_theta(zsq, 3, blocks=blocks)
    theta_blocks_dict = get_theta(
        zsq, {(0, 0): 2, (0, 1): 3, (0, 2): 1 + 1, (1, 1): 4, (1, 2): 2, (2, 2): 3}, blocks=blocks
    )
    save("theta_n40", zsq=zsq, ks=ks, thetas=thetas, blocks=blocks, theta_blocks=theta_blocks,
         theta_blocks_dict=theta_blocks_dict)


def operators():
    rng = np.random.default_rng(17)
    n, kk = 23, 4
    e = n * (n - 1) // 2
    w = rng.standard_normal((kk, e))
    dual = rng.standard_normal((kk, n, n))
    s, st = sum_squareform(n)
    dn = rng.standard_normal((n, kk))
    save(
        "operators_n23",
        w=w,
        dual=dual,
        dn=dn,
        lap_vec=laplacian_squareform_vec(w),
        lap_single=laplacian_squareform(w[0]),
        adj_vec=laplacian_squareform_adj_vec(dual),
        adj_single=laplacian_squareform_adj(dual[0]),
        s_w=s @ w.T,
        st_d=st @ dn,
        s_dense=s.toarray(),
        sq2vec=square_to_vec(dual),
    )


def batched():
    """2-D (K, E) call: shared stop, spectral-norm residuals (smooth_learning.py:128-129,158-159)."""
    rng = np.random.default_rng(23)
    n, kk = 30, 3
    zs = np.stack([pdist(rgg_signals_exact(n, 80, seed=40 + c).T) ** 2 for c in range(kk)])
    th = np.array([[get_theta(squareform(z), 4)] for z in zs])
    w = gsp_learn_graph_log_degrees(zs * th, 1, 1)
    w_it = {f"w_raw_it{its}": gsp_learn_graph_log_degrees(zs * th, 1, 1, maxit=its, edge_tol=-np.inf)
            for its in (1, 2, 7, 40)}
    e = n * (n - 1) // 2
    init = rng.uniform(0, 0.3, size=(e, kk))  # internal (E, K) layout, cf. glmm.py:58-59
    w_warm = gsp_learn_graph_log_degrees(zs * th, 1, 1, edge_init=init.copy(), maxit=25)
    # each graph solved alone: differs from the batch (shared stop) -- kept to show that
    w_single = np.stack([squareform(gsp_learn_graph_log_degrees(z * t, 1, 1))
                         for z, t in zip(zs, th[:, 0])])
    save("batched_n30", z=zs, theta=th, w=w, edge_init=init, w_warm=w_warm, w_single=w_single, **w_it)


def batched_multitile():
    """A batch that spans several 128-node tiles and stops on the shared rule part-way (default tol 1e-5):
    K = 5 graphs (not a power of two), n = 140 nodes.  Inputs are stored as the point clouds the distances come
    from (the tests rebuild z from them); outputs are the reference's thresholded weights, which pin the stop
    iteration (one iteration more or less moves the weights by ~1e-5 relative)."""
    rng = np.random.default_rng(140)
    n, kk = 140, 5
    pts = rng.uniform(size=(kk, n, 3))
    zs = np.stack([pdist(p) ** 2 for p in pts])
    th = np.array([[get_theta(squareform(z), 6)] for z in zs])
    w = gsp_learn_graph_log_degrees(zs * th, 1, 1)
    w_it = gsp_learn_graph_log_degrees(zs * th, 1, 1, maxit=12)      # maxit-limited, thresholded at 1e-5
    save("batched_n140_k5", pts=pts, theta=th, w=w, w_it12=w_it)


def window_model():
    """tests/test_models.py:26-28 data; outputs stored (the reference test asserts nothing)."""
    data = np.random.default_rng(240920).standard_normal((23, 7))
    out = {"x": data}
    for ws in (5, 10):
        for k in (2, 5):
            mdl = WindowLogModel(window_size=ws, avg_degree=k, random_state=240920).fit(data)
            out[f"theta_ws{ws}_k{k}"] = np.asarray(mdl.theta_)
            out[f"weights_ws{ws}_k{k}"] = mdl.weights_
    save("window_n7", **out)


def clustering():
    """tests/test_clustering.py:15-49 data; M-step numbers and fitted labels."""
    rng = np.random.default_rng(2410241319)
    n_nodes, n_samples, n_clusters = 25, 100, 3
    y_true = rng.choice(n_clusters, size=n_samples, replace=True)
    laps = sample_er_laplacian(n_nodes, edge_p=0.2, edge_w_min=0.3, edge_w_max=2,
                               n_graphs=n_clusters, seed=rng)
    x = np.zeros((n_samples, n_nodes))
    for i, lap in enumerate(laps):
        mask = y_true == i
        x[mask] = sample_lgmrf(lap, n_samples=mask.sum(), seed=rng)

    resp = np.random.default_rng(1).dirichlet(np.ones(n_clusters), size=n_samples)
    wts, means, lapl = _estimate_gauss_laplacian_parameters(x, resp, 2.0, avg_degree=3)
    wts_t, means_t, lapl_t = _estimate_gauss_laplacian_parameters(x, resp, 2.0, theta=0.7)
    blocks = np.repeat([0, 1], [12, 13])
    _, _, lapl_b = _estimate_gauss_laplacian_parameters(x, resp, 2.0, avg_degree=2, blocks=blocks)

    glmm = GLMMShim(n_components=n_clusters, avg_degree=3, random_state=7, max_iter=30)
    lab_glmm = glmm.fit_predict(x)
    kg = KGraphs(n_clusters=n_clusters, avg_degree=3, random_state=7, max_iter=30)
    lab_kg = kg.fit_predict(x)
    kg2 = KGraphsV2(n_clusters=n_clusters, avg_degree=3, random_state=7, max_iter=30)
    lab_kg2 = kg2.fit_predict(x)
    save(
        "clustering_n25",
        x=x, y_true=y_true, resp=resp, blocks=blocks,
        mstep_weights=wts, mstep_means=means, mstep_laplacians=lapl,
        mstep_theta_laplacians=lapl_t, mstep_block_laplacians=lapl_b,
        glmm_labels=lab_glmm, glmm_laplacians=glmm.laplacians_, glmm_means=glmm.means_,
        glmm_weights=glmm.weights_,
        kgraphs_labels=lab_kg, kgraphs_laplacians=kg.laplacians_, kgraphs_score=kg.score_,
        kgraphs2_labels=lab_kg2, kgraphs2_laplacians=kg2.laplacians_, kgraphs2_means=kg2.means_,
    )


def assignment():
    """Assignment-step quantities of the unmodified reference on the clustering
    data and fitted parameters above: kgraphs.py:87-88, kgraphs.py:203-205, glmm.py:238-279."""
    g = np.load(os.path.join(HERE, "clustering_n25.npz"))
    x = g["x"]
    kg = KGraphs(n_clusters=3, avg_degree=3)
    kg.laplacians_ = g["kgraphs_laplacians"]
    smooth = kg._smoothness(x)
    kg2 = KGraphsV2(n_clusters=3, avg_degree=3)
    kg2.laplacians_, kg2.means_ = g["kgraphs2_laplacians"], g["kgraphs2_means"]
    csmooth = kg2._centered_smoothness(x)
    glmm = GLMMShim(n_components=3, avg_degree=3)
    glmm.laplacians_, glmm.means_, glmm.weights_ = g["glmm_laplacians"].copy(), g["glmm_means"], g["glmm_weights"]
    glmm.n_nodes_ = x.shape[1]
    log_prob = glmm._estimate_log_prob(x)
    # a dense, non-symmetric, non-Laplacian stack: the einsum is defined for any square matrices
    rng = np.random.default_rng(5)
    dense = rng.standard_normal((2, 25, 25))
    kg.laplacians_ = dense
    smooth_dense = kg._smoothness(x)
    save("assign_n25", x=x, kgraphs_laplacians=g["kgraphs_laplacians"], smoothness=smooth,
         labels=np.argmin(smooth, axis=1), kgraphs2_laplacians=g["kgraphs2_laplacians"],
         kgraphs2_means=g["kgraphs2_means"], centered_smoothness=csmooth,
         glmm_laplacians=g["glmm_laplacians"], glmm_means=g["glmm_means"], regul=np.float64(glmm.regul),
         log_prob=log_prob, dense=dense, smoothness_dense=smooth_dense)


def glmm_random_init():
    """GLMM(laplacian_init="random"): the reference draws every component with
    sample_uniform_laplacian(n, self.random_state) (glmm.py:202-207, sampling/graphs.py:10-14)."""
    from graph_learn.sampling.graphs import sample_uniform_laplacian
    n, k = 7, 3
    same = np.stack([sample_uniform_laplacian(n, 3) for _ in range(k)])          # integer seed: identical draws
    gen = np.random.default_rng(11)
    advancing = np.stack([sample_uniform_laplacian(n, gen) for _ in range(k)])   # a Generator advances
    save("glmm_random_init_n7", n=np.int64(n), k=np.int64(k), int_seed=np.int64(3), laplacians_int_seed=same,
         generator_seed=np.int64(11), laplacians_generator=advancing)


if __name__ == "__main__":
    if len(sys.argv) > 1:          # regenerate only the named sets
        for name in sys.argv[1:]:
            globals()[name]()
        sys.exit(0)
    solver_states()
    logmodel_fit()
    theta_cases()
    operators()
    batched()
    batched_multitile()
    window_model()
    clustering()
    assignment()
    glmm_random_init()
