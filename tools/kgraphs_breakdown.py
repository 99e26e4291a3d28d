"""Stage times of one k-Graphs outer iteration at config 4 (K = 8, n = 1,000, m = 100,000, float32).
Measurement aid."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graph_learning_b200.smooth_learning as sl  # noqa: E402
from graph_learning_b200.clustering.assign import QuadForm  # noqa: E402
from graph_learning_b200.operators import laplacian_squareform_vec  # noqa: E402
from graph_learning_b200.synthetic import rgg_signals_filtered  # noqa: E402

K, n, m = 8, 1000, 100_000
x = torch.cat([rgg_signals_filtered(n, m // K, seed=10 + k, device="cuda", dtype=torch.float32) for k in range(K)])
labels = torch.arange(m, device="cuda") % K
rec = {}


def stage(name, fn, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    rec[name + "_ms"] = (time.perf_counter() - t0) / reps * 1e3
    return out


parts = stage("gather_rows", lambda: [x[labels == k] for k in range(K)])
sq = stage("distances", lambda: torch.stack([sl.pdist_sq(p, dtype=torch.float32) for p in parts]))
th = stage("thetas", lambda: [sl._theta_packed(sq[k], n, 10) for k in range(K)])
zs = sq * torch.tensor(th, device="cuda", dtype=torch.float32)[:, None]
w = stage("solve", lambda: sl.gsp_learn_graph_log_degrees(zs, 1, 1, dtype="float32"))
lap = stage("laplacians", lambda: laplacian_squareform_vec(w))


def assign():
    with QuadForm(lap, dtype=torch.float32) as qf:
        return qf.apply(x, labels=True)


stage("assign_total", assign)
print(json.dumps(rec))
