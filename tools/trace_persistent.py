"""Phase breakdown of the persistent iteration kernel (per rank), from its %globaltimer stamps.
    python tools/trace_persistent.py                      (1 GPU)
    torchrun --nproc-per-node 8 tools/trace_persistent.py (sharded, peer exchange)
Profiling aid, not part of the product path."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200 import _native as nat  # noqa: E402
from graph_learning_b200.sharded import ShardedLogModelSolver  # noqa: E402
from graph_learning_b200.smooth_learning import Solver, _wrap_device_memory  # noqa: E402


def main():
    n = int(os.environ.get("TRACE_N", "20000"))
    iters = int(os.environ.get("TRACE_ITERS", "600"))
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = nat.lib()
    ntiles, tt = lib.gl_num_tiles(n), nat.TILE * nat.TILE
    gen = torch.Generator(device=dev).manual_seed(0)
    if world == 1:
        solver = Solver(n, 1, torch.float32)
        sh = None
        b, e = 0, ntiles
    else:
        sh = ShardedLogModelSolver(n, torch.float32, exchange="peer")
        solver = sh.solver
        b, e = sh.tile_range
    z = _wrap_device_memory(solver.z_tiles_ptr(0), (e - b) * tt, torch.float32, dev, owner=solver)
    z.copy_(torch.rand((e - b) * tt, generator=gen, device=dev) * 3.0 + 0.05)
    solver.trace_enable()
    for rep in range(2):
        _wrap_device_memory(solver.w_tiles_ptr(0), (e - b) * tt, torch.float32, dev, owner=solver).zero_()
        (sh or solver).begin(1.0, 1.0, 0.5, iters, 0.0)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        (sh or solver).run()
        t1.record()
        torch.cuda.synchronize()
    st = solver.trace_read(iters)[50:iters - 1].astype(np.float64)
    nxt = solver.trace_read(iters)[51:iters].astype(np.float64)
    names = ["E sweep (CTA 0)", "barrier 1", "fold + publish", "peer flag wait", "node chunk", "barrier 2", "decision"]
    d = np.diff(st, axis=1)
    rec = {"rank": rank, "world": world, "ms_per_iter_events": t0.elapsed_time(t1) / iters,
           "us_per_iter_trace": float(np.median(nxt[:, 0] - st[:, 0]) / 1e3),
           "phases_us_median": {k: float(np.median(d[:, i]) / 1e3) for i, k in enumerate(names)},
           "phases_us_mean": {k: float(np.mean(d[:, i]) / 1e3) for i, k in enumerate(names)},
           "loop_back_us": float(np.median(nxt[:, 0] - st[:, 7]) / 1e3)}
    for r in range(world):
        if r == rank:
            print(json.dumps(rec), flush=True)
        if world > 1:
            dist.barrier()
    if sh is not None:
        sh.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
