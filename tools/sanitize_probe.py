"""Small solves that exercise every kernel with shared-memory reductions, float64 atomics, the hand-rolled
grid barrier or ticket logic, for `compute-sanitizer --tool memcheck|racecheck`:
the persistent iteration kernel (fp32 / fp64), the classic edge pass + node tail (stepwise and batched
K = 3 with the gated Gram kernel), the kNN one-launch kernel, theta, distances, layout kernels.

    compute-sanitizer --tool racecheck python tools/sanitize_probe.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200 import smooth_learning as sl  # noqa: E402
from graph_learning_b200.knn import KNNLogModel  # noqa: E402


def main():
    torch.cuda.set_device(0)
    rng = np.random.default_rng(0)
    n = 300
    x = rng.standard_normal((64, n))
    for dtype in ("float64", "float32"):
        mdl = sl.LogModel(avg_degree=5, maxit=12, dtype=dtype).fit(x)          # pdist, theta, persistent kernel, square
        assert np.isfinite(mdl.weights_).all()
    z = np.abs(rng.standard_normal(n * (n - 1) // 2)) + 0.1
    zt = torch.as_tensor(z, device="cuda")
    s = sl.Solver(n, 1, torch.float64)
    s.load(zt[None, :].contiguous())
    s.begin(maxit=6)
    for _ in range(6):                                                           # classic kernels, stepwise
        s.edge_pass()
        s.node_tail()
    assert s.status().n_iter == 6
    s.close()
    zb = np.stack([z, z * 1.2, z * 0.7])
    w = sl.gsp_learn_graph_log_degrees(zb, 1, 1, maxit=8, dtype="float32")       # batched path + gated Gram kernel
    assert np.isfinite(w).all()
    xk = rng.uniform(size=(3, 3000))
    k = KNNLogModel(avg_degree=4, n_neighbors=8, maxit=10).fit(xk)               # kNN search + one-launch iteration
    assert np.isfinite(k.weights_).all()
    print("sanitize probe ok")


if __name__ == "__main__":
    main()
