"""Phase breakdown of the resident batched kernel (persistent_batch_kernel, K > 1) from its %globaltimer
stamps, CTA 0:  python tools/trace_batch.py   (TRACE_N, TRACE_K, TRACE_ITERS, TRACE_DTYPE from the environment).
Profiling aid, not part of the product path."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200 import _native as nat  # noqa: E402
from graph_learning_b200.smooth_learning import Solver, _wrap_device_memory  # noqa: E402


def main():
    n = int(os.environ.get("TRACE_N", "1000"))
    K = int(os.environ.get("TRACE_K", "8"))
    iters = int(os.environ.get("TRACE_ITERS", "600"))
    tdt = torch.float64 if os.environ.get("TRACE_DTYPE", "float32") == "float64" else torch.float32
    dev = torch.device("cuda", 0)
    lib = nat.lib()
    ntiles, tt = lib.gl_num_tiles(n), nat.TILE * nat.TILE
    gen = torch.Generator(device=dev).manual_seed(0)
    solver = Solver(n, K, tdt)
    # distances in packed order so that the masked slots of the tiles stay zero
    z = torch.rand((K, n * (n - 1) // 2), generator=gen, device=dev, dtype=tdt) * 3.0 + 0.05
    solver.load(z, None, None)
    solver.trace_enable()
    for rep in range(2):
        for k in range(K):
            solver.w_view(k).zero_()
        solver.begin(1.0, 1.0, 0.5, iters, 0.0)
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        solver.run()
        t1.record()
        torch.cuda.synchronize()
    st = solver.trace_read(iters)[50:iters - 1, :7].astype(np.float64)
    nxt = solver.trace_read(iters)[51:iters, :7].astype(np.float64)
    keep = (np.arange(50, iters - 1) % 50) != 49          # drop iterations that end a launch
    st, nxt = st[keep], nxt[keep]
    names = ["E sweep (CTA 0)", "barrier 1", "node chunk", "barrier 2", "folds", "decision"]
    d = np.diff(st, axis=1)
    rec = {"n": n, "K": K, "dtype": str(tdt), "tiles": int(ntiles), "items": int(ntiles) * K,
           "us_per_iter_events": 1e3 * t0.elapsed_time(t1) / iters,
           "us_per_iter_trace": float(np.median(nxt[:, 0] - st[:, 0]) / 1e3),
           "phases_us_median": {k: float(np.median(d[:, i]) / 1e3) for i, k in enumerate(names)},
           "loop_back_us": float(np.median(nxt[:, 0] - st[:, 6]) / 1e3)}
    print(json.dumps(rec), flush=True)
    solver.close()


if __name__ == "__main__":
    main()
