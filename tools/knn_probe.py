"""Measurement aid: where does the time of a kNN-support solve go?  Times, with CUDA events on one GPU at
n = 1e6 / k = 20, (a) the resident solver (ShardedKNNSolver in its one-launch single-GPU mode) and (b) the
one-shot C-ABI call for several iteration counts -- per-iteration time must not depend on the count.

    python tools/knn_probe.py [n] [k]
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200.knn import gsp_learn_graph_log_degrees_knn, knn_graph, locality_keys, theta_from_knn  # noqa: E402
from graph_learning_b200.sharded import ShardedKNNSolver  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    torch.cuda.set_device(0)
    x = torch.rand((3, n), generator=torch.Generator(device="cuda").manual_seed(0), device="cuda")
    ei, ej, d2, (_, d2l) = knn_graph(x, k)
    z = d2 * theta_from_knn(d2l, 10)
    keys = locality_keys(x)
    nnz = z.numel()
    out = {"n": n, "k": k, "edges": nnz, "resident_ms_per_it": {},
           "oneshot_ms": {}}
    sv = ShardedKNNSolver(ei, ej, z, n, dtype=torch.float32, order_key=keys)
    for its in (100, 100, 300, 900, 2000):
        sv.begin(maxit=its, tol=0.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sv.run(poll_every=50)
        e1.record()
        torch.cuda.synchronize()
        out["resident_ms_per_it"][str(its)] = e0.elapsed_time(e1) / its
    sv.close()
    for its in (100, 100, 300, 900):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gsp_learn_graph_log_degrees_knn(ei, ej, z, n, maxit=its, tol=0.0, order_key=keys)
        e1.record()
        torch.cuda.synchronize()
        out["oneshot_ms"][str(its)] = e0.elapsed_time(e1)
    o = out["oneshot_ms"]
    out["oneshot_ms_per_it_300_to_900"] = (o["900"] - o["300"]) / 600
    out["oneshot_ms_per_it_100_to_300"] = (o["300"] - o["100"]) / 200
    print(json.dumps(out))


if __name__ == "__main__":
    main()
