"""Worker of tests/test_gpu_multi.py: runs under torchrun with one process per GPU (NCCL) and checks the
multi-GPU product path on hardware -- the persistent kernel's PEER branch (flag publish, ld.acquire.sys
spin, peer loads; csrc/persistent.cu), the NCCL all-reduce schedule, and the distributed forms of
gsp_learn_graph_log_degrees and LogModel.fit -- against the single-GPU solver and the CPU oracle.
Rank 0 prints one JSON line; a failed assertion exits non-zero on that rank.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from graph_learning_b200 import sharded as shd  # noqa: E402
from graph_learning_b200 import smooth_learning as sl  # noqa: E402
from graph_learning_b200.synthetic import rgg_signals_exact  # noqa: E402
from oracle import logmodel_oracle as orc  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def same_on_all_ranks(t: torch.Tensor) -> bool:
    ref = t.clone()
    dist.broadcast(ref, 0)
    ok = torch.tensor([1.0 if torch.equal(t.view(torch.int64), ref.view(torch.int64)) else 0.0], device=t.device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return bool(ok.item() == 1.0)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    report = {"world": world}

    # ---- 1. solver: peer exchange + persistent kernel, and the NCCL schedule, fp64 and fp32 ---------------
    n = 700                                             # 6 x 6 tile grid: 21 tiles, uneven shards, ragged edge
    rng = np.random.default_rng(5)
    pts = rng.uniform(size=(n, 2))
    z_np = orc.to_packed(((pts[:, None] - pts[None]) ** 2).sum(-1)) * 60.0
    for dtype, tol in ((torch.float64, 1e-11), (torch.float32, 1e-4)):
        z = torch.as_tensor(z_np, device=dev).to(dtype)
        for exchange in ("peer", "nccl"):
            sv = shd.ShardedLogModelSolver(n, dtype, exchange=exchange)
            assert sv.exchange == exchange
            sv.load_packed(z[None, :].contiguous())
            for maxit, tol_stop in ((1, 0.0), (30, 0.0), (77, 0.0), (1000, 1e-5)):
                sv.solver.w_view(0).zero_()
                sv.begin(1.0, 1.0, 0.5, maxit, tol_stop)
                sv.run()
                st = sv.status()
                w = sv.gather_packed(-np.inf)[0]
                d = sv.solver.node_state()
                stat = torch.tensor([float(st.n_iter), st.rel_primal, st.rel_dual], dtype=torch.float64, device=dev)
                assert same_on_all_ranks(torch.cat([stat, d.reshape(-1)])), (exchange, maxit, "replicated state differs")
                # single-GPU solver on this rank's device: same kernels without the exchange
                one = sl.Solver(n, 1, dtype)
                one.load(z[None, :].contiguous())
                one.begin(1.0, 1.0, 0.5, maxit, tol_stop)
                one.run()
                st1 = one.status()
                w1 = one.store(-np.inf)[0]
                one.close()
                wo, oinfo = orc.solve_one_pass(z_np, maxit=maxit, tol=tol_stop, threshold=False)
                if dtype == torch.float64:
                    assert st.n_iter == st1.n_iter == oinfo.n_iter, (exchange, maxit, st, st1, oinfo)
                else:
                    assert abs(st.n_iter - oinfo.n_iter) <= 2 and abs(st1.n_iter - oinfo.n_iter) <= 2
                e_one = rel(w.cpu().numpy(), w1.cpu().numpy())
                e_orc = rel(w.cpu().numpy(), wo)
                assert e_one <= (1e-12 if dtype == torch.float64 else 1e-5), (exchange, maxit, e_one)
                assert e_orc <= tol, (exchange, maxit, e_orc)
                report[f"solver_{exchange}_{str(dtype)[6:]}_maxit{maxit}"] = {"n_iter": st.n_iter, "vs_1gpu": e_one, "vs_oracle": e_orc}
            sv.close()

    # ---- 2. the reference-facing call, distributed, host in -> host out ------------------------------------
    zs = z_np.copy()
    w_all, info = sl.gsp_learn_graph_log_degrees(zs, 1, 1, maxit=60, distributed=True, return_info=True)
    w_one = sl.gsp_learn_graph_log_degrees(zs, 1, 1, maxit=60)
    wo, _ = orc.solve_one_pass(z_np, maxit=60)
    assert isinstance(w_all, np.ndarray) and w_all.shape == (n, n) and info.n_iter == 60
    assert rel(w_all, w_one) <= 1e-12 and rel(w_all, orc.to_square(wo)) <= 1e-11
    w_root = sl.gsp_learn_graph_log_degrees(zs, 1, 1, maxit=60, distributed=True, result_on="root")
    assert (w_root is None) == (rank != 0)
    if rank == 0:   # two runs differ in the last bits (float64 atomics sum the degree partials in run order)
        assert rel(w_root, w_all) <= 1e-13
    # warm start + z_scale + pinned tensor in + out= on the root
    w0 = np.abs(np.random.default_rng(1).standard_normal(z_np.size)) * 0.01
    zt = torch.from_numpy(z_np / 3.0).pin_memory()
    out = torch.empty((n, n), dtype=torch.float64).pin_memory() if rank == 0 else None
    res = sl.gsp_learn_graph_log_degrees(zt, 1, 1, edge_init=torch.from_numpy(w0), maxit=25, z_scale=3.0,
                                         distributed=True, result_on="root", out=out)
    wo, _ = orc.solve_one_pass(z_np / 3.0 * 3.0, maxit=25, edge_init=w0.copy())
    if rank == 0:
        assert res is out and rel(out.numpy(), orc.to_square(wo)) <= 1e-11
    # tiny problem: one tile in all, the second rank owns none -> the collective form on every rank
    zt_small = z_np[:100 * 99 // 2].copy()
    w_s = sl.gsp_learn_graph_log_degrees(zt_small, 1, 1, maxit=40, distributed=True)
    wo_s, _ = orc.solve_one_pass(zt_small, maxit=40)
    assert rel(w_s, orc.to_square(wo_s)) <= 1e-11
    # a (K, E) batch is not sharded: every rank solves it (replicas), same answer as one GPU
    zb = np.stack([z_np, z_np * 1.3])
    wb = sl.gsp_learn_graph_log_degrees(zb, 1, 1, maxit=20, distributed=True)
    wb1 = sl.gsp_learn_graph_log_degrees(zb, 1, 1, maxit=20)
    assert rel(wb, wb1) <= 1e-13
    report["function_distributed"] = "ok"

    # ---- 3. LogModel.fit distributed: Z tiles per rank, theta from merged candidates -------------------------
    x = rgg_signals_exact(300, 600, seed=2)
    for dtype, tol in (("float64", 1e-9), ("float32", 1e-4)):
        one = sl.LogModel(avg_degree=6, maxit=400, dtype=dtype).fit(x)
        many = sl.LogModel(avg_degree=6, maxit=400, dtype=dtype, distributed=True).fit(x)
        assert many.theta_ == one.theta_, (many.theta_, one.theta_)
        assert many.n_iter_ == one.n_iter_
        e = rel(many.weights_, one.weights_)
        assert e <= (1e-12 if dtype == "float64" else 1e-5), e
        z = orc.pdist_sq(x)
        theta = orc.theta_for_degree(orc.to_square(z), 6)
        w_ref, _ = orc.solve_two_pass(z * theta, maxit=400)
        assert rel(many.weights_, orc.to_square(w_ref)) <= tol
        root = sl.LogModel(avg_degree=6, maxit=400, dtype=dtype, distributed=True, weights_on="root").fit(x)
        assert (root.weights_ is None) == (rank != 0)
        report[f"logmodel_fit_{dtype}"] = {"theta": many.theta_, "n_iter": many.n_iter_, "vs_1gpu": e}
    shd.release_workspaces()

    # ---- 4. kNN support, row-sharded -----------------------------------------------------------------------
    from graph_learning_b200.knn import knn_graph, theta_from_knn
    nn = 20000
    xk = torch.rand((3, nn), generator=torch.Generator(device=dev).manual_seed(0), device=dev)
    ei, ej, d2, (_, d2l) = knn_graph(xk, 10)
    theta = theta_from_knn(d2l, 5)
    zk = (d2 * theta).double()
    for exch in shd.KNN_EXCHANGES:
        sv = shd.ShardedKNNSolver(ei, ej, zk, nn, dtype=torch.float64, exchange=exch)
        sv.begin(maxit=40, tol=0.0)
        sv.run(poll_every=10)
        wk = sv.gather(-np.inf)
        stk = sv.status()
        sv.close()
        wo, oinfo = orc.solve_knn(ei.cpu().numpy(), ej.cpu().numpy(), zk.cpu().numpy(), nn, maxit=40, tol=0.0, threshold=False)
        assert stk.n_iter == oinfo.n_iter == 40
        ek = rel(wk.cpu().numpy(), wo)
        assert ek <= 1e-11, (exch, ek)
        report[f"knn_sharded_{exch}"] = {"edges": int(zk.numel()), "vs_oracle": ek}

    dist.barrier()
    if rank == 0:
        print(json.dumps(report), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
