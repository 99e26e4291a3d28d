"""Assignment-step kernel at the config-4 shape (K = 8 sparse Laplacians of n = 1,000 nodes, about 10
edges per node, m = 100,000 float32 signals) for timing and for an ncu capture.  Measurement aid."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200.clustering.assign import QuadForm  # noqa: E402

K, n, m, deg = 8, 1000, 100_000, 10
g = torch.Generator(device="cuda").manual_seed(0)
iu = torch.triu_indices(n, n, 1, device="cuda")
lap = torch.zeros((K, n, n), device="cuda")
for k in range(K):
    keep = torch.rand(iu.shape[1], generator=g, device="cuda") < deg / (n - 1)
    w = torch.rand(iu.shape[1], generator=g, device="cuda") * keep
    lap[k][iu[0], iu[1]] = -w
    lap[k] = lap[k] + lap[k].t()
    lap[k] -= torch.diag(lap[k].sum(1))
x = torch.randn((m, n), generator=g, device="cuda")
with QuadForm(lap, dtype="float32") as qf:
    for _ in range(3):
        q, lab = qf.apply(x, labels=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    reps = 10
    for _ in range(reps):
        q, lab = qf.apply(x, labels=True)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    ref = torch.stack([((x[:2048] @ lap[k]) * x[:2048]).sum(1) for k in range(K)], 1)
    print(json.dumps({"K": K, "n": n, "m": m, "entries": qf.nnz, "kernel_ms": ms,
                      "entry_samples_per_s": qf.nnz * m / (ms * 1e-3),
                      "x_read_gbs": x.numel() * 4 / (ms * 1e-3) / 1e9,
                      "rel_vs_dense_first_2048": float((q[:2048] - ref).norm() / ref.norm())}))
