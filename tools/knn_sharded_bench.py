"""Config 5, row-sharded over the ranks of a torchrun launch (or 1 GPU): kNN graph of n points in the
unit cube (built on rank 0's device data, replicated by seed), sharded solve, per-iteration time by
CUDA events (max over ranks).  Prints one JSON line on rank 0.  Measurement aid."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_learning_b200.knn import knn_graph, theta_from_knn  # noqa: E402
from graph_learning_b200.sharded import ShardedKNNSolver  # noqa: E402


def main():
    n = int(os.environ.get("KNN_N", "1000000"))
    k = int(os.environ.get("KNN_K", "20"))
    iters = int(os.environ.get("KNN_ITERS", "400"))
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    x = torch.rand((3, n), generator=torch.Generator(device=dev).manual_seed(0), device=dev)
    ei, ej, d2, (_, d2l) = knn_graph(x, k)       # every rank builds the same graph (same seed)
    theta = theta_from_knn(d2l, 10)
    sv = ShardedKNNSolver(ei, ej, d2 * theta, n, dtype=torch.float32)
    times = []
    for rep in range(3):
        sv.begin(maxit=iters, tol=0.0)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sv.run(poll_every=50)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    st = sv.status()
    w = sv.gather()
    nnz = int(w.numel())
    ms = min(times[1:]) / iters
    if rank == 0:
        print(json.dumps({"n": n, "k": k, "edges": nnz, "n_gpus": world, "iters": st.n_iter, "ms_per_iteration": ms,
                          "edge_iters_per_s": nnz / (ms * 1e-3), "algorithmic_GBps_20B_per_edge": 20 * nnz / (ms * 1e-3) / 1e9,
                          "local_slots_rank0": int(sv.lib.gl_knn_local_slots(sv._h)), "kept_edges": int((w > 0).sum()),
                          "weights_checksum": float(w.double().sum())}), flush=True)
    sv.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
