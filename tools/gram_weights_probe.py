"""Development probe: how the Z-build method moves the learned float32 weights (north_star: 1e-4
relative, identical support outside a band).  Reference = float64 direct distances + float64 solver
(parity-pinned to the oracle).  Run on the GPU box."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graph_learning_b200.smooth_learning as sl  # noqa: E402
from graph_learning_b200.synthetic import rgg_signals_filtered  # noqa: E402


def main():
    n, m, k = int(sys.argv[1]), int(sys.argv[2]), 8
    maxit = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    x = rgg_signals_filtered(n, m, seed=0, device="cuda", dtype=torch.float32)
    z64 = sl.pdist_sq(x.double(), method="direct")
    th64 = sl._theta_packed(z64, n, k)
    w64, info64 = sl.gsp_learn_graph_log_degrees(z64 * th64, 1, 1, maxit=maxit, return_info=True, edge_tol=-np.inf)
    w64t = sl.gsp_learn_graph_log_degrees(z64 * th64, 1, 1, maxit=maxit)
    scale = float((x.double() ** 2).sum(0).mean())
    print(json.dumps({"n": n, "m": m, "theta64": th64, "iters64": info64.n_iter, "mean_norm2": scale,
                      "z_median_over_norm2": float(z64.median()) / scale, "z_min_over_norm2": float(z64.min()) / scale,
                      "edges_kept": int((w64t > 0).sum().item() // 2)}), flush=True)
    for method, chunk, variant in (("direct", 0, 0), ("gram", 256, 0), ("gram16", 256, 0), ("gram16", 128, 0), ("gram", 64, 0),
                                   ("gram2", 256, 0), ("gram", 256, 2)):
        z = sl.pdist_sq(x, method=method, chunk=chunk, variant=variant)
        th = sl._theta_packed(z, n, k)
        w, info = sl.gsp_learn_graph_log_degrees(z * th, 1, 1, maxit=maxit, return_info=True, dtype="float32")
        d = w.double() - w64t
        away = (w64 - 1e-3).abs() > 1e-5
        mism = int((((w > 0) != (w64t > 0)) & away).sum().item())
        dz = z.double() - z64
        print(json.dumps({"method": method, "chunk": chunk, "variant": variant, "theta_rel": abs(th - th64) / th64, "iters": info.n_iter,
                          "w_rel": float(d.norm() / w64t.norm()), "w_max_abs": float(d.abs().max()),
                          "support_mismatch_outside_band": mism, "z_rel_fro": float(dz.norm() / z64.norm()),
                          "z_max_abs_over_norm2": float(dz.abs().max()) / scale}), flush=True)


if __name__ == "__main__":
    main()
