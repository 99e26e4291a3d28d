#!/usr/bin/env python
"""bench.py -- LogModel edges*iterations/s on B200 (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W            # our arm (config 3)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference --steps K --warmup W    # the reference's own CPU code
    python bench.py --config 2 [--dtype float64]             # config 2 (n = 2,000, m = 10,000), converged parity
    python bench.py --config 4 | --config 5                  # batched K = 8 M-step | kNN support, n = 1e6

Workload (config 3 of BASELINE.json): LogModel dense, n = 20,000 nodes (E = 199,990,000 edges),
m = 50,000 synthetic smooth signals (random-geometric-graph low-pass filtered noise), fp32.
One "step" = one solver call of ITERS iterations (reference default maxit = 1000, tol = 0 so the
count is fixed) over all E edges.  value = E * ITERS * steps / device time, Z already resident in
HBM; e2e = the same through the reference-facing call gsp_learn_graph_log_degrees with HOST buffers
(H2D of the packed Z -- each rank its own rows --, solve, gather, D2H of the dense (n, n) weights the
reference returns) at every N.  At N > 1 the triangle tiles are sharded over the ranks (total work
fixed -> "strong").  Before anything is timed, every N runs a PARITY leg: 10 and 20 iterations from
the zero start on the benchmark's own Z, gathered and compared with the C oracle on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CONFIGS = {
    2: dict(n=2000, m=10000, name="config2"),
    3: dict(n=20000, m=50000, name="config3"),
}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--config", type=int, default=3, choices=[2, 3, 4, 5],
                   help="BASELINE.json config: 3 = the headline (default), 2 = n 2,000 converged parity, "
                        "4 = K 8 batched M-step, 5 = kNN support n 1e6")
    p.add_argument("--n", type=int, default=None, help="nodes (overrides the config's)")
    p.add_argument("--m", type=int, default=None, help="signals (overrides the config's)")
    p.add_argument("--iters", type=int, default=1000, help="solver iterations per step (maxit)")
    p.add_argument("--avg-degree", type=int, default=8)
    p.add_argument("--dtype", default="float32", choices=["float32", "float64"])
    p.add_argument("--z-method", default="gram16", choices=["gram", "gram16", "gram2", "direct"],
                   help="Z build (setup): tcgen05 Gram GEMM on 3-slice bf16 / 2-slice fp16 / 2-slice bf16 splits, or exact differences")
    p.add_argument("--z-variant", type=int, default=-1, help="kernel variant of the Gram GEMM (-1: the method's default)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-parity", action="store_true")
    p.add_argument("--no-fit-e2e", action="store_true")
    p.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl"],
                   help="multi-GPU exchange: peer-memory reads in the node tail, or ncclAllReduce")
    p.add_argument("--cpu-n", type=int, default=2000, help="nodes of the bounded CPU sample (config 2's n)")
    p.add_argument("--cpu-iters", type=int, default=8, help="iterations per CPU step")
    a = p.parse_args()
    if a.config in CONFIGS:
        a.n = a.n or CONFIGS[a.config]["n"]
        a.m = a.m or CONFIGS[a.config]["m"]
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own code (graph_learn, installed unmodified under baseline/_ref by
# __graft_entry__.build(); /root/reference when this runs in the build container), else the NumPy/SciPy
# restatement in oracle/ (kind "port").
# ---------------------------------------------------------------------------------------------------
def load_reference():
    """(module graph_learn.smooth_learning, where) or (None, why)."""
    for where in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(where, "graph_learn")):
            if where not in sys.path:
                sys.path.insert(0, where)
            try:
                sys.dont_write_bytecode = True
                import graph_learn.smooth_learning as ref  # noqa: PLC0415
                return ref, where
            except Exception as exc:  # noqa: BLE001
                return None, f"import failed: {exc!r}"
    return None, "graph_learn not found under baseline/_ref or /root/reference"


def cpu_sample_inputs(n_s: int, seed: int = 0, m_s: int = 64):
    """Bounded sample of the workload: the same kind of synthetic smooth-signal distances on n_s nodes
    (CPU throughput in edges*iters/s does not depend on n once the edge vectors leave the caches).
    Returns (signals (m_s, n_s), theta-scaled packed distances)."""
    from oracle import logmodel_oracle as orc
    rng = np.random.default_rng(seed)
    pts = rng.uniform(size=(n_s, 2))
    centres = rng.uniform(size=(m_s, 2))
    # smooth signals: low-pass kernel features of the point cloud, cheap to build
    x = np.exp(-((pts[None, :, :] - centres[:, None, :]) ** 2).sum(-1) / 0.05) + 0.01 * rng.standard_normal((m_s, n_s))
    z = orc.pdist_sq(x)
    theta = orc.theta_for_degree(orc.to_square(z), 8)
    return x, z * theta


def reference_step_times(ref, z, iters: int, steps: int, warmup: int):
    """Times `steps` calls of the reference solver (fixed `iters` iterations: tol = 0 never stops) and,
    separately, the O(n^2) pure-Python operator build it repeats on every call
    (smooth_learning.py:133 -> :13-44).  Returns (sum of call times, sum of build times)."""
    n = int(round((1 + np.sqrt(1 + 8 * z.size)) / 2))
    for _ in range(warmup):
        ref.gsp_learn_graph_log_degrees(z, 1, 1, maxit=max(1, iters // 4), tol=0.0)
    t_call = t_build = 0.0
    for _ in range(steps):
        t0 = time.perf_counter()
        ref.gsp_learn_graph_log_degrees(z, 1, 1, maxit=iters, tol=0.0)
        t1 = time.perf_counter()
        ref.sum_squareform(n)
        t2 = time.perf_counter()
        t_call += t1 - t0
        t_build += t2 - t1
    return t_call, t_build


def port_step_times(z, iters: int, steps: int, warmup: int):
    from oracle import logmodel_oracle as orc
    n = int(round((1 + np.sqrt(1 + 8 * z.size)) / 2))
    op = orc.sum_operator(n)
    for _ in range(warmup):
        orc.solve_two_pass(z, maxit=max(1, iters // 4), tol=0.0, operator=op)
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.solve_two_pass(z, maxit=iters, tol=0.0, operator=op)
    return time.perf_counter() - t0, 0.0


def c_oracle_rate(z, iters: int):
    """The OpenMP C restatement (one sweep, all host threads) -- a stronger CPU baseline than the
    reference's own schedule; reported next to it, not instead of it."""
    try:
        from oracle import c_oracle
        c_oracle.solve(z, maxit=2, tol=0.0, threshold=False)
        t0 = time.perf_counter()
        c_oracle.solve(z, maxit=iters, tol=0.0, threshold=False)
        return z.size * iters / (time.perf_counter() - t0), c_oracle.threads()
    except Exception:  # noqa: BLE001
        return None, 0


def cpu_rate(a, steps: int, warmup: int):
    """Per-iteration rate of the CPU path on the bounded sample: (value, block for cpu_baseline, wall s)."""
    _, z = cpu_sample_inputs(a.cpu_n)
    ref, where = load_reference()
    e_s = z.size
    if ref is not None:
        t_call, t_build = reference_step_times(ref, z, a.cpu_iters, steps, warmup)
        value = e_s * a.cpu_iters * steps / max(t_call - t_build, 1e-9)
        block = {
            "value": value, "unit": "edge*iters/s", "cores": 1, "kind": "reference", "host_cores": os.cpu_count(),
            "reference_from": where,
            "sample": f"n={a.cpu_n} nodes (E={e_s}; config 2's size -- the benchmarked config 3 has n=20000, where the "
                      f"reference needs ~39 GB and ~10 s per iteration), {a.cpu_iters} iterations per call, {steps} calls; "
                      "the UNMODIFIED reference graph_learn.smooth_learning.gsp_learn_graph_log_degrees (fp64, four SciPy CSR "
                      "mat-vecs per iteration, single-threaded), time of the call minus the sum_squareform operator build "
                      "it repeats on every call (BASELINE.md section 4 item 2)",
            "ms_per_iteration": 1e3 * (t_call - t_build) / (a.cpu_iters * steps),
            "operator_build_s_per_call": t_build / steps,
            "value_including_operator_build": e_s * a.cpu_iters * steps / t_call,
        }
        return value, block, t_call + t_build
    t_call, _ = port_step_times(z, a.cpu_iters, steps, warmup)
    value = e_s * a.cpu_iters * steps / t_call
    block = {
        "value": value, "unit": "edge*iters/s", "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
        "reference_unavailable": where,
        "sample": f"n={a.cpu_n} nodes (E={e_s}), {a.cpu_iters} iterations per call, {steps} calls, operator construction "
                  "excluded; NumPy/SciPy restatement of smooth_learning.py:146-167 (oracle/logmodel_oracle.py, fp64, four "
                  "CSR mat-vecs per iteration, single-threaded like the reference)",
        "ms_per_iteration": 1e3 * t_call / (a.cpu_iters * steps),
    }
    return value, block, t_call


def reference_stage_timings(a):
    """BASELINE.md section 4 items 1, 3, 4, 5 with the reference's own functions on this box's host cores
    (item 2 is cpu_rate).  Bounded: n = cpu_n (config 2's size), pdist on an m-subsample."""
    ref, where = load_reference()
    if ref is None:
        return {"unavailable": where}
    from scipy.spatial.distance import pdist, squareform
    n = a.cpu_n
    m_s = 1000
    x, _ = cpu_sample_inputs(n, m_s=m_s)
    out = {"n": n, "reference_from": where}
    t0 = time.perf_counter()
    ref.sum_squareform(n)
    out["sum_squareform_build_s"] = time.perf_counter() - t0                       # smooth_learning.py:13-44
    t0 = time.perf_counter()
    z = pdist(x.T) ** 2                                                             # :278
    out["pdist_s"] = time.perf_counter() - t0
    out["pdist_m"] = m_s
    out["pdist_s_extrapolated_to_m10000"] = out["pdist_s"] * 10000 / m_s            # linear in m, labelled
    t0 = time.perf_counter()
    theta = ref.get_theta(squareform(z), a.avg_degree)                              # :177-235
    out["get_theta_s"] = time.perf_counter() - t0
    out["theta"] = float(theta)
    x1, _ = cpu_sample_inputs(100, seed=1, m_s=1000)
    t0 = time.perf_counter()
    mdl = ref.LogModel(avg_degree=a.avg_degree).fit(x1)                             # :286-300, config 1, converged
    out["logmodel_fit_n100_m1000_s"] = time.perf_counter() - t0
    out["logmodel_fit_n100_edges"] = int((np.asarray(mdl.weights_) > 0).sum() // 2)
    t0 = time.perf_counter()
    ref.LogModel(avg_degree=a.avg_degree, maxit=10).fit(x)
    out[f"logmodel_fit_n{n}_m{m_s}_maxit10_s"] = time.perf_counter() - t0
    return out


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    value, block, wall = cpu_rate(a, a.steps, min(a.warmup, 1))
    line = {
        "impl": "reference",
        "metric": "LogModel edges*iters/sec",
        "value": value,
        "unit": "edge*iters/s",
        "n_gpus": a.gpus,
        "steps": a.steps,
        "warmup": a.warmup,
        "ms_per_step": 1e3 * wall / a.steps,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(a, cpu=True),
        "cpu_baseline": block,
        "e2e": {"value": value, "unit": "edge*iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(a, cpu=False):
    name = CONFIGS.get(a.config, {}).get("name", f"config{a.config}")
    esz = 4 if a.dtype == "float32" else 8
    state_gb = 2 * esz * a.n * (a.n - 1) / 2 / 1e9
    l2 = (f"inputs (w+z = {state_gb:.1f} GB) exceed the 126 MB L2; no flush needed" if state_gb > 0.5 else
          f"L2-RESIDENT: w+z = {state_gb * 1e3:.0f} MB fits the 126 MB L2, so the 'hbm' roofline figure is an L2 rate (labelled)")
    return {
        "workload": f"{name}: LogModel dense n={a.n} (E={a.n * (a.n - 1) // 2}), m={a.m}, "
                    f"{a.dtype}, maxit={a.iters}, tol=0 (fixed iteration count), avg_degree={a.avg_degree}",
        "n_nodes": a.n, "n_edges": a.n * (a.n - 1) // 2, "n_signals": a.m, "iters_per_step": a.iters,
        "l2": l2,
        "parallelism": f"tile-shard x{a.gpus}" + (f", exchange={a.exchange}" if a.gpus > 1 else ""),
        **({"cpu_sample_nodes": a.cpu_n, "cpu_sample_iters": a.cpu_iters,
            "cpu_sample_note": f"the CPU arm runs a bounded n={a.cpu_n} sample of this workload, not n={a.n}"} if cpu else {}),
    }


# ---------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            # warm the two queries of the loop here: the first NVML clock / reasons query on a fresh process
            # takes milliseconds and holds driver locks, which stalls kernel launches -- visible as a +30 %
            # outlier when the whole timed region is 30 ms (config 2)
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            try:
                pynvml.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:  # noqa: BLE001
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


class Env:
    """Rank / device / process-group plumbing shared by every GPU leg."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        # stdout carries exactly ONE JSON line: everything libraries print (e.g. NCCL's version banner)
        # is routed to stderr, the line itself is written to the saved descriptor at the end.
        sys.stdout.flush()
        self.real_stdout = os.dup(1)
        os.dup2(2, 1)
        if self.world != a.gpus and self.world > 1:
            raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={self.world}")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def emit(self, line):
        if self.rank == 0:
            os.write(self.real_stdout, (json.dumps(line) + "\n").encode())

    def finish(self, cleanup=(), rc=0):
        """Orderly teardown; the exit status is that of the teardown itself (a fault while unmapping peer
        memory or destroying the communicator must not be masked)."""
        try:
            for fn in cleanup:
                fn()
            self.torch.cuda.synchronize()
            if self.world > 1:
                self.dist.barrier()
                self.torch.cuda.synchronize()
                self.dist.destroy_process_group()
        except Exception as exc:  # noqa: BLE001
            print(f"bench teardown failed on rank {self.rank}: {exc!r}", file=sys.stderr)
            rc = 1
        sys.stdout.flush()
        sys.stderr.flush()
        # leave without interpreter finalisers (they may wait on an already destroyed communicator);
        # rc reports whether the explicit teardown above went through
        os._exit(rc)


# ---------------------------------------------------------------------------------------------------
# GPU arm, dense configs 2 and 3
# ---------------------------------------------------------------------------------------------------
def run_dense(a):
    import ctypes

    env = Env(a)
    torch, dist = env.torch, env.dist
    from graph_learning_b200 import _native as nat
    from graph_learning_b200 import sharded as shd
    from graph_learning_b200.sharded import ShardedLogModelSolver
    from graph_learning_b200.smooth_learning import (LogModel, Solver, _gram_workspace, gsp_learn_graph_log_degrees,
                                                     release_gram_workspace)
    from graph_learning_b200.synthetic import rgg_signals_filtered

    rank, world, dev = env.rank, env.world, env.dev
    lib = nat.lib()
    tdt = torch.float32 if a.dtype == "float32" else torch.float64
    gdt = nat.GL_F32 if a.dtype == "float32" else nat.GL_F64
    esz = 4 if a.dtype == "float32" else 8
    n, m = a.n, a.m
    E = n * (n - 1) // 2
    ntiles = lib.gl_num_tiles(n)
    tt = nat.TILE * nat.TILE
    tol_rel = 1e-4 if a.dtype == "float32" else 1e-9

    # ---- synthetic inputs (setup, untimed except for the reported Z build) ----------------------
    x = rgg_signals_filtered(n, m, seed=0, device=dev, dtype=tdt)
    if world > 1:
        # one realisation for all ranks: the generator's sparse products (cuSPARSE) are not bit-reproducible
        # from device to device, and the parity leg compares the shards with rank 0's copy of Z
        dist.broadcast(x, 0)
    z_full = torch.empty(ntiles * tt, dtype=tdt, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    use_gram = a.dtype == "float32" and a.z_method != "direct"
    terms = 2 if a.z_method in ("gram2", "gram16") else 3
    z_variant = a.z_variant if a.z_variant >= 0 else (7 if a.z_method == "gram16" else 0)

    def build_z():
        if use_gram:
            ws = _gram_workspace(m, n, terms, dev)
            nat.check(lib.gl_pdist_sq_gram(x.data_ptr(), m, n, x.stride(0), None, None, z_full.data_ptr(),
                                           nat.LAYOUT_TILED, 0, ntiles, terms, 0, z_variant, ws.data_ptr(), ws.numel(), None))
        else:
            nat.check(lib.gl_pdist_sq(x.data_ptr(), m, n, x.stride(0), None, None, z_full.data_ptr(),
                                      nat.LAYOUT_TILED, 0, ntiles, gdt, None))

    if use_gram:
        build_z()  # warm-up (first-use module load, workspace allocation)
    torch.cuda.synchronize()
    ev[0].record()
    build_z()
    ev[1].record()
    torch.cuda.synchronize()
    z_ms = ev[0].elapsed_time(ev[1])
    release_gram_workspace()
    x_host = None
    if not a.no_fit_e2e:
        x_host = torch.empty((m, n), dtype=tdt, pin_memory=True)
        x_host.copy_(x)
    del x
    th = ctypes.c_double()
    t0 = time.perf_counter()
    nat.check(lib.gl_theta(z_full.data_ptr(), nat.LAYOUT_TILED, n, a.avg_degree, gdt, ctypes.byref(th), None))
    theta_ms = 1e3 * (time.perf_counter() - t0)
    theta = th.value

    # ---- resident solver state ---------------------------------------------------------------------
    if world == 1:
        solver = Solver(n, 1, tdt)
        solver.z_view(0).copy_(z_full)
        sharded = None
    else:
        sharded = ShardedLogModelSolver(n, tdt, exchange=a.exchange)
        sharded.load_tiles(z_full)
        solver = sharded.solver
    solver.scale_z(theta)          # theta multiplied into the resident tiles in place (smooth_learning.py:289)
    # the same scaling on the setup copy used by the parity / e2e legs: product in float64, rounded once,
    # exactly what gl_solver_scale_z does
    z_full.copy_((z_full.double() * theta).to(tdt))

    def solve(iters, tol=0.0):
        solver.w_view(0).zero_()
        if sharded is None:
            solver.begin(1.0, 1.0, 0.5, iters, tol)
            solver.run()
        else:
            sharded.begin(1.0, 1.0, 0.5, iters, tol)
            sharded.run(20)

    def gathered_packed():
        """untresholded packed weights of all shards on every rank"""
        if sharded is None:
            return solver.store(-np.inf)[0]
        return sharded.gather_packed(-np.inf)[0]

    # ---- parity leg: the benchmark's own Z, 10 and 20 iterations, against the C oracle --------------
    parity = None
    if not a.no_parity:
        parity = parity_leg(a, env, lib, nat, solver, sharded, solve, gathered_packed, z_full, n, E, ntiles, tdt, gdt, tol_rel)

    for _ in range(a.warmup):
        solve(a.iters)
    assert solver.status().n_iter == a.iters, solver.status()
    clocks = ClockSampler(env.local)
    env.barrier()
    l0 = nat.launch_count()
    sh0 = getattr(sharded, "launches", 0) if sharded else 0
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        solve(a.iters)
    e1.record()
    env.barrier()
    clk = clocks.stop()
    ms = env.max_over_ranks(e0.elapsed_time(e1))
    launches = (nat.launch_count() - l0) + ((sharded.launches - sh0) if sharded else 0)
    value = E * a.iters * a.steps / (ms * 1e-3)

    # ---- dominant kernel alone, CUDA events on its stream ------------------------------------------
    local_edges = E if world == 1 else (sharded.tile_range[1] - sharded.tile_range[0]) * tt
    hbm_peak, peak_src = peaks()
    persistent = (world == 1 or sharded.exchange == "peer") and os.environ.get("GL_B200_PERSISTENT", "1") != "0"
    if persistent:
        # gl::persistent_kernel: one cooperative launch runs a chunk of whole iterations
        kname = "gl::persistent_kernel"
        if world == 1:
            solver.begin(1.0, 1.0, 0.5, a.iters, 0.0)
        else:
            sharded.begin(1.0, 1.0, 0.5, a.iters, 0.0)
        env.barrier()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lk = nat.launch_count()
        k0.record()
        solver.run()
        k1.record()
        torch.cuda.synchronize()
        n_launch = max(1, nat.launch_count() - lk)
        kernel_ms = k0.elapsed_time(k1) / n_launch
        bytes_per_launch = 3 * esz * local_edges * a.iters / n_launch
        iters_per_launch = a.iters / n_launch
    else:
        kname = "gl::edge_pass_kernel"
        sharded.begin(1.0, 1.0, 0.5, 10 ** 6, 0.0) if world > 1 else solver.begin(1.0, 1.0, 0.5, 10 ** 6, 0.0)
        reps = 30
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for i in range(reps + 3):
            if i >= 3:
                evs[i - 3][0].record()
            solver.edge_pass()
            if i >= 3:
                evs[i - 3][1].record()
            if world > 1:
                sharded._allreduce()
            solver.node_tail()
        torch.cuda.synchronize()
        kernel_ms = float(np.mean([s.elapsed_time(e) for s, e in evs]))
        bytes_per_launch = 3 * esz * local_edges
        iters_per_launch = 1
    achieved = bytes_per_launch / (kernel_ms * 1e-3) / 1e9
    traffic, traffic_src = read_traffic((1.0 if world == 1 else local_edges / float(ntiles * tt)) * iters_per_launch, world, n)

    line = {
        "metric": "LogModel edges*iters/sec",
        "value": value,
        "unit": "edge*iters/s",
        "n_gpus": world,
        "steps": a.steps,
        "warmup": a.warmup,
        "ms_per_step": ms / a.steps,
        "ms_per_iteration": ms / a.steps / a.iters,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f32" if a.dtype == "float32" else "f64",
        "data": "synthetic",
        "config": workload_config(a),
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": {
            "bound": "hbm", "kernel": kname, "achieved": achieved, "peak": hbm_peak,
            "unit": "GB/s", "frac": achieved / hbm_peak, "peak_source": peak_src,
            "frac_of_nominal_8TBps": achieved / 8000.0,
            "bytes_per_edge_per_iteration": 3 * esz, "edges_per_launch": int(local_edges),
            "iterations_per_launch": iters_per_launch, "bytes_per_launch": bytes_per_launch, "kernel_ms": kernel_ms,
            "whole_iteration_frac": (3 * esz * E / world) / (ms * 1e-3 / a.steps / a.iters) / 1e9 / hbm_peak,
            "traffic": traffic, "traffic_source": traffic_src,
        },
        "setup": z_setup(a, z_ms, use_gram, terms, E, m, n, theta_ms, theta),
    }
    if parity is not None:
        line["parity"] = parity
    if a.config == 2 and not a.no_parity:
        line["converged_parity"] = converged_parity(a, env, solver, sharded, solve, gathered_packed, z_full, n, tdt, gdt, lib, nat,
                                                    ntiles, tol_rel)

    # free the resident benchmark state before the API legs allocate their own workspaces
    z_packed = torch.empty(E, dtype=tdt, device=dev)
    nat.check(lib.gl_tiles_to_packed(z_full.data_ptr(), z_packed.data_ptr(), n, 1, gdt, 0, ntiles, None))
    del z_full
    if sharded is not None:
        sharded.close()
    else:
        solver.close()
    sharded = solver = None

    # ---- end to end through the reference-facing call, host buffers, every N ---------------------------
    if not a.no_e2e:
        z_host = torch.empty(E, dtype=tdt, pin_memory=True)
        z_host.copy_(z_packed)
        del z_packed
        w_host = torch.empty((n, n), dtype=tdt, pin_memory=True) if rank == 0 else None
        torch.cuda.synchronize()
        dist_arg = world > 1

        def e2e_step():
            return gsp_learn_graph_log_degrees(z_host, 1, 1, maxit=a.iters, tol=0.0, dtype=a.dtype, out=w_host,
                                               return_info=True, distributed=dist_arg, result_on="root")

        for _ in range(min(a.warmup, 2)):
            e2e_step()
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            _, info = e2e_step()
        env.barrier()
        dt = env.max_over_ranks(time.perf_counter() - t0)
        assert info.n_iter == a.iters
        if world > 1:
            _, _, eb, ee = shd.sharded_workspace(n, tdt).solver.shard_rows()
            h2d = env.sum_over_ranks(float((ee - eb) * esz))
        else:
            h2d = float(E * esz)
        line["e2e"] = {"value": E * a.iters * a.steps / dt, "unit": "edge*iters/s",
                       "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(n * n * esz),
                       "ms_per_step": 1e3 * dt / a.steps,
                       "api": "gsp_learn_graph_log_degrees(pinned host packed (E,) on every rank -> pinned host (n,n) on rank 0"
                              + (f"; distributed=True over {world} ranks, each rank uploads the rows of its own tile shard, "
                                 "h2d = sum over ranks)" if world > 1 else ")")}
        del z_host
    else:
        del z_packed

    # ---- LogModel.fit(X) end to end: host X -> Z tiles per rank -> theta -> solve -> host (n, n) ------------
    if x_host is not None:
        mdl = LogModel(avg_degree=a.avg_degree, maxit=a.iters, tol=0.0, dtype=a.dtype, distributed=world > 1,
                       weights_on="root")
        w_host = torch.empty((n, n), dtype=tdt, pin_memory=True) if rank == 0 else None
        for _ in range(2):            # warm-up: workspace allocations, first-use kernel loads
            mdl.fit(x_host, out=w_host)
        env.barrier()
        reps = 3
        t0 = time.perf_counter()
        for _ in range(reps):
            mdl.fit(x_host, out=w_host)
        env.barrier()
        dt = env.max_over_ranks(time.perf_counter() - t0) / reps
        line["fit_e2e"] = {
            "api": "LogModel(avg_degree, maxit, tol=0, dtype" + (", distributed=True, weights_on='root'" if world > 1 else "")
                   + ").fit(pinned host X (m, n)) -> theta_, weights_ (n, n) on the host",
            "ms_per_fit": 1e3 * dt, "value": E * a.iters / dt, "unit": "edge*iters/s",
            "h2d_bytes_per_fit": int(m * n * esz) * world, "d2h_bytes_per_fit": int(n * n * esz),
            "n_iter": int(mdl.n_iter_), "theta": float(mdl.theta_),
            "theta_matches_setup": bool(abs(float(mdl.theta_) - theta) <= 1e-6 * abs(theta)),
        }
        del x_host, mdl
        release_gram_workspace()

    # ---- CPU baseline: the reference's own code on this box's host cores (rank 0, N = 1) -----------
    if world == 1 and rank == 0 and not a.no_cpu_baseline:
        _, block, _ = cpu_rate(a, 1, 1)
        _, zc = cpu_sample_inputs(a.cpu_n)
        vc, nthr = c_oracle_rate(zc, 4 * a.cpu_iters)
        block["c_openmp_oracle"] = {"value": vc, "threads": nthr,
                                    "note": "one-sweep C restatement with OpenMP, all host threads (stronger than the reference schedule)"}
        block["stages"] = reference_stage_timings(a)
        line["cpu_baseline"] = block
    env.emit(line)
    env.finish([shd.release_workspaces])


def parity_leg(a, env, lib, nat, solver, sharded, solve, gathered_packed, z_full, n, E, ntiles, tdt, gdt, tol_rel):
    """10 and 20 iterations from the zero start on the benchmark's own (theta-scaled) Z: the gathered
    weights of all shards against oracle/logmodel_oracle.c on rank 0 (maxit-limited run at
    config 3), the replicated node state and the solver status bit-identical on every rank."""
    torch, dist = env.torch, env.dist
    z64 = None
    if env.rank == 0:
        zp = torch.empty(E, dtype=tdt, device=env.dev)
        nat.check(lib.gl_tiles_to_packed(z_full.data_ptr(), zp.data_ptr(), n, 1, gdt, 0, ntiles, None))
        z64 = zp.cpu().numpy().astype(np.float64)
        del zp
    out = {"against": "oracle/logmodel_oracle.c (float64 one-sweep restatement of smooth_learning.py:125-167, pinned to "
                      "the reference's goldens) on the benchmark's own Z", "tolerance": tol_rel, "ranks": env.world,
           "iters": [], "rel_err": [], "oracle_s": 0.0}
    ok = True
    for its in (10, 20):
        solve(its)
        st = solver.status()
        w = gathered_packed()
        d = solver.node_state()
        same_state = True
        if env.world > 1:
            stat = torch.tensor([float(st.n_iter), st.rel_primal, st.rel_dual], dtype=torch.float64, device=env.dev)
            both = torch.cat([stat, d.reshape(-1)])
            ref0 = both.clone()
            dist.broadcast(ref0, 0)
            same = torch.tensor([1.0 if torch.equal(both.view(torch.int64), ref0.view(torch.int64)) else 0.0], device=env.dev)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            same_state = bool(same.item() == 1.0)
        rel = None
        if env.rank == 0:
            from oracle import c_oracle
            if env.world > 1:   # torchrun pins OMP_NUM_THREADS=1; the other ranks only wait here
                out["oracle_threads"] = c_oracle.set_threads(max(1, (os.cpu_count() or 2) - 2 * env.world))
            t0 = time.perf_counter()
            wo, it_o, _, _ = c_oracle.solve(z64, maxit=its, tol=0.0, threshold=False)
            out["oracle_s"] += time.perf_counter() - t0
            wg = w.cpu().numpy().astype(np.float64)
            rel = float(np.linalg.norm(wg - wo) / np.linalg.norm(wo))
            assert it_o == its == st.n_iter, (it_o, its, st.n_iter)
        rel = env.max_over_ranks(rel if rel is not None else 0.0)
        out["iters"].append(its)
        out["rel_err"].append(rel)
        ok = ok and rel <= tol_rel and same_state and st.n_iter == its
        out["node_state_and_status_bit_identical_on_all_ranks"] = same_state and out.get(
            "node_state_and_status_bit_identical_on_all_ranks", True)
        del w
    out["ok"] = bool(ok)
    if not ok:
        env.emit({"error": "parity leg failed", "parity": out})
        env.finish(rc=1)
    return out


def converged_parity(a, env, solver, sharded, solve, gathered_packed, z_full, n, tdt, gdt, lib, nat, ntiles, tol_rel):
    """Config 2: run to convergence (tol 1e-5, maxit 1000 -- the reference defaults) and compare weights,
    support and iteration count with the C oracle on the same Z."""
    torch = env.torch
    E = n * (n - 1) // 2
    solve(1000, 1e-5)
    st = solver.status()
    w = gathered_packed()
    res = {"n_iter": int(st.n_iter), "converged": bool(st.converged)}
    if env.rank == 0:
        from oracle import c_oracle
        zp = torch.empty(E, dtype=tdt, device=env.dev)
        nat.check(lib.gl_tiles_to_packed(z_full.data_ptr(), zp.data_ptr(), n, 1, gdt, 0, ntiles, None))
        z64 = zp.cpu().numpy().astype(np.float64)
        wo, it_o, _, _ = c_oracle.solve(z64, maxit=1000, tol=1e-5, threshold=False)
        wg = w.cpu().numpy().astype(np.float64)
        rel = float(np.linalg.norm(wg - wo) / np.linalg.norm(wo))
        band = np.abs(wo - 1e-3) < 1e-5
        res.update({"oracle_n_iter": int(it_o), "rel_err": rel, "tolerance": tol_rel,
                    "support_mismatches_outside_band": int(((wg >= 1e-3) != (wo >= 1e-3))[~band].sum()),
                    "ok": bool(rel <= tol_rel and abs(int(st.n_iter) - int(it_o)) <= (2 if a.dtype == "float32" else 0))})
    return res


def z_setup(a, z_ms, use_gram, terms, E, m, n, theta_ms, theta):
    """Z build (K1) report: algorithmic flops 2*E*m (triangle only) and, for the
    tensor-core form, the bf16 MMA flops actually executed (full 128 x 128 units, 3 or 6 cross
    terms per k-step) against the measured cuBLAS bf16 peaks."""
    out = {"z_build_ms": z_ms, "z_build_algorithmic_tflops": 2.0 * E * m / (z_ms * 1e-3) / 1e12,
           "theta_ms": theta_ms, "theta": theta}
    if not use_gram:
        out["z_kernel"] = "gl::pdist_kernel (direct differences, FP32/FP64 FMA pipe)"
        return out
    nt = -(-n // 128)
    units = nt * (nt + 1) // 2                                       # default variant: 128 x 128 units
    m_pad = -(-m // 64) * 64
    pairs = 4 if a.z_method == "gram16" else (3 if terms == 2 else 6)
    mma_flops = units * 2.0 * 128 * 128 * m_pad * pairs
    achieved = mma_flops / (z_ms * 1e-3) / 1e12
    burst = sustained = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        burst, sustained = pk.get("bf16_tflops"), pk.get("bf16_tflops_sustained")
    except OSError:
        burst, sustained = 1590.0, 1400.0
    out.update({
        "z_method": a.z_method, "mma_products_per_k_step": pairs,
        "z_kernel": f"gl::gram_kernel<{terms},128,32,split> (tcgen05.mma kind::f16 on {terms}-slice {'fp16' if a.z_method == 'gram16' else 'bf16'} operands with "
                    "an exact leading slice, TMA SWIZZLE_64B, leading/small TMEM accumulators folded into float64 "
                    "registers every 256 samples) + gl::gram_colmax_kernel + gl::gram_split_kernel",
        "z_roofline": {"bound": "tensor", "achieved": achieved, "unit": "TFLOP/s (bf16 MMA flops executed, "
                       "split pre-pass included in the time)", "peak": burst, "frac": achieved / burst if burst else None,
                       "peak_sustained": sustained, "frac_of_sustained": achieved / sustained if sustained else None,
                       "mma_flops": mma_flops, "algorithmic_flops": 2.0 * E * m},
    })
    return out


def read_traffic(fraction=1.0, world=1, n=20000):
    """dram bytes per launch of the dominant kernel from the committed ncu captures (taken at N = 1,
    n = 20,000, all tiles): per-iteration traffic x `fraction` (= share of the tiles this rank owns x
    iterations per launch).  Returns (bytes or None, where the figure comes from)."""
    if n != 20000:
        return None, "no ncu capture at this size"
    for name, key in (("persistent_traffic.json", "dram_bytes_per_iteration"),
                      ("edge_pass_traffic.json", "dram_bytes_per_launch")):
        path = os.path.join(ROOT, "profiles", name)
        if os.path.exists(path):
            try:
                with open(path) as f:
                    v = json.load(f).get(key)
                if v is not None:
                    src = f"ncu --set full capture at N=1 (profiles/{name})"
                    if world > 1:
                        src += f", scaled by this rank's tile share (not measured at N={world})"
                    return v * fraction, src
            except Exception:  # noqa: BLE001
                continue
    return None, "no capture found"


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.config in (2, 3):
        run_dense(args)
    else:
        import bench_configs
        bench_configs.run(args)
