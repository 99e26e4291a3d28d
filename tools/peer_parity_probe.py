"""Measurement aid (torchrun, >= 2 GPUs): repeats 20-iteration sharded fp32 solves at n nodes with the peer
exchange and reports, per repeat, the relative error of the gathered weights against the C oracle and
against a single-GPU solve on rank 0's device -- alternating GL_B200_EARLY_PUBLISH=1/0 -- plus the
per-iteration time.  A lost or early-read degree partial would show up as scatter between repeats."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200 import smooth_learning as sl  # noqa: E402
from graph_learning_b200.sharded import ShardedLogModelSolver  # noqa: E402
from oracle import c_oracle  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    pts = torch.rand((n, 2), generator=g, dtype=torch.float64).to(dev)
    iu = torch.triu_indices(n, n, 1, device=dev)
    z = ((pts[iu[0]] - pts[iu[1]]) ** 2).sum(1) * (0.15 * n)
    z32 = z.float().contiguous()
    E = z32.numel()
    out = {"n": n, "world": world, "runs": []}
    wo = None
    if rank == 0:
        c_oracle.set_threads(max(1, (os.cpu_count() or 2) - 2 * world))
        wo, _, _, _ = c_oracle.solve(z32.double().cpu().numpy(), maxit=20, tol=0.0, threshold=False)
    one = sl.Solver(n, 1, torch.float32)
    one.load(z32[None, :].contiguous())
    one.begin(maxit=20, tol=0.0)
    one.run()
    w1 = one.store(-np.inf)[0].double()
    one.close()
    var = os.environ.get("PROBE_TOGGLE", "GL_B200_EARLY_PUBLISH")   # which switch the repeats alternate
    sv = ShardedLogModelSolver(n, torch.float32, exchange="peer")
    sv.load_packed(z32[None, :].contiguous())
    for rep in range(2 * reps):
        mode = "1" if rep % 2 == 0 else "0"
        os.environ[var] = mode
        dist.barrier()
        sv.solver.w_view(0).zero_()
        sv.begin(1.0, 1.0, 0.5, 20, 0.0)
        sv.run()
        w = sv.gather_packed(-np.inf)[0].double()
        # timing: a 400-iteration run
        sv.solver.w_view(0).zero_()
        sv.begin(1.0, 1.0, 0.5, 400, 0.0)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sv.run()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rec = {var: mode, "vs_1gpu": float((w - w1).norm() / w1.norm()), "us_per_iteration": float(t.item()) * 1e3 / 400}
        if rank == 0:
            rec["vs_oracle"] = float(np.linalg.norm(w.cpu().numpy() - wo) / np.linalg.norm(wo))
        out["runs"].append(rec)
    sv.close()
    if rank == 0:
        out["one_gpu_vs_oracle"] = float(np.linalg.norm(w1.cpu().numpy() - wo) / np.linalg.norm(wo))
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
