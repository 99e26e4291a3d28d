"""Measurement aid: where the time of LogModel.fit(X) goes at config 3 (host X -> host (n, n)).
Repeats the steps of LogModel._fit_resident one by one with a device synchronisation after each.

    python tools/fit_breakdown.py [n] [m] [iters]
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200 import _native as nat  # noqa: E402
from graph_learning_b200 import smooth_learning as sl  # noqa: E402
from graph_learning_b200.synthetic import rgg_signals_filtered  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 50000
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    torch.cuda.set_device(0)
    x = rgg_signals_filtered(n, m, seed=0, device="cuda", dtype=torch.float32)
    xh = torch.empty((m, n), dtype=torch.float32, pin_memory=True)
    xh.copy_(x)
    del x
    out = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
    rec = {"n": n, "m": m, "iters": iters, "runs": []}
    for rep in range(3):
        t = {}

        def lap(name, t0):
            torch.cuda.synchronize()
            t[name] = 1e3 * (time.perf_counter() - t0)
            return time.perf_counter()

        torch.cuda.synchronize()
        t_all = t0 = time.perf_counter()
        xd = xh.to("cuda", non_blocking=True)
        t0 = lap("h2d_x", t0)
        solver = sl.Solver(n, 1, torch.float32)
        t0 = lap("solver_create", t0)
        sl._pdist_into(xd, None, None, solver.z_view(0), nat.LAYOUT_TILED, 0, solver.tile_end, "gram")
        t0 = lap("z_build", t0)
        theta = sl._theta_tiles(solver.z_view(0), n, 8)
        solver.scale_z(theta)
        t0 = lap("theta_scale", t0)
        solver.load_w_packed(None)
        solver.begin(1, 1, 0.5, iters, 0.0)
        solver.run()
        t0 = lap("solve", t0)
        res = solver.square(1e-3)
        t0 = lap("tiles_to_square", t0)
        out.copy_(res, non_blocking=True)
        t0 = lap("d2h_w", t0)
        solver.close()
        del res, xd
        t0 = lap("free", t0)
        t["sum"] = 1e3 * (time.perf_counter() - t_all)
        # the estimator itself
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sl.LogModel(avg_degree=8, maxit=iters, tol=0.0, dtype="float32").fit(xh, out=out)
        torch.cuda.synchronize()
        t["LogModel.fit"] = 1e3 * (time.perf_counter() - t0)
        rec["runs"].append(t)
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
