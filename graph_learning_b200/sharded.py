"""Multi-GPU LogModel solve: the triangle's 128x128 tiles are split into contiguous,
work-balanced shards, one per rank (one process per GPU, ``torch.distributed``); every rank keeps
its shard of w and z resident and the node vectors replicated.  Per iteration the only exchange is
the sum of the float64 exchange buffer [S p | S w' | residual sums] (2n + 2 values, 320 KB at
n = 20,000) between the edge pass and the node tail: either ONE NCCL all-reduce
inside a captured graph (``exchange="nccl"``) or -- the default for 2..8 ranks -- reads of every rank's
partials from peer-mapped memory inside the persistent iteration kernel (``exchange="peer"``).

:func:`solve_distributed` and :func:`sharded_workspace` put that path behind the reference-facing calls
(``gsp_learn_graph_log_degrees(..., distributed=True)``, ``LogModel(..., distributed=True)``);
:class:`ShardedKNNSolver` is the row-sharded kNN-support solve (config 5).

The dense schedule is written against a tiny backend protocol so that the same driver runs the CUDA
``Solver`` over NCCL and, in the CPU test-suite, a host emulation over gloo.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Protocol

import numpy as np
import torch
import torch.distributed as dist

from . import _native as nat

__all__ = ["shard_tiles", "ShardBackend", "sharded_schedule", "ShardedLogModelSolver", "sharded_workspace",
           "solve_distributed", "release_workspaces", "shard_rows", "knn_shard_labels", "KnnShardBackend",
           "knn_exchange", "knn_sharded_schedule", "ShardedKNNSolver"]


def shard_tiles(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous tile range [begin, end) of ``rank`` (diagonal tiles weigh half)."""
    b, e = ctypes.c_int(), ctypes.c_int()
    nat.check(nat.lib().gl_shard_tiles(int(n), int(rank), int(world), ctypes.byref(b), ctypes.byref(e)))
    return b.value, e.value


class ShardBackend(Protocol):
    def begin(self) -> None: ...            # local partial of S w0 into the exchange buffer
    def begin_finish(self) -> None: ...     # replicated node initialisation
    def edge_pass(self) -> None: ...        # local edge sweep, partials into the exchange buffer
    def node_tail(self) -> None: ...        # replicated node tail + stop rule
    def exchange(self): ...                 # tensor to all-reduce (sum)
    def done(self) -> bool: ...             # stop flag (may synchronise)


def sharded_schedule(backend: ShardBackend, maxit: int, group=None, poll_every: int = 1) -> None:
    """begin -> all-reduce -> finish -> [edge pass -> all-reduce -> node tail]*  (eager form)."""
    backend.begin()
    dist.all_reduce(backend.exchange(), group=group)
    backend.begin_finish()
    it = 0
    while it < maxit:
        for _ in range(min(poll_every, maxit - it)):
            backend.edge_pass()
            dist.all_reduce(backend.exchange(), group=group)
            backend.node_tail()
            it += 1
        if backend.done():
            break


class ShardedLogModelSolver:
    """CUDA backend: one :class:`~graph_learning_b200.smooth_learning.Solver` per rank over its tile
    shard, iterations captured in a CUDA graph together with the NCCL all-reduce."""

    def __init__(self, n: int, dtype=torch.float32, n_graphs: int = 1, group=None, exchange: str = "auto"):
        """``exchange``: "peer" = the node tail reads all ranks' partials over NVLink from
        peer-mapped memory (no collective call; single graph only), "nccl" = ncclAllReduce inside
        the captured graph, "auto" = peer when possible."""
        from .smooth_learning import Solver

        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n = n
        if n_graphs > 1 and self.world > 1:
            # the shared stop rule of a batch needs spectral norms over ALL edges (smooth_learning.py:158-159)
            raise ValueError("a (K, E) batch is not tile-sharded: solve it on one GPU (replicas only)")
        self.tile_range = shard_tiles(n, self.rank, self.world)
        self.solver = Solver(n, n_graphs, dtype, tile_range=self.tile_range)
        self._xch = self.solver.exchange()
        # a rank without tiles (tiny n, many ranks) cannot run the persistent kernel the peer exchange lives
        # in: such a split uses the collective form on every rank
        every_rank_has_tiles = all(e > b for b, e in (shard_tiles(n, r, self.world) for r in range(self.world)))
        if exchange == "auto":
            exchange = "peer" if (self.world > 1 and n_graphs == 1 and self.world <= 8 and every_rank_has_tiles) else "nccl"
        elif exchange == "peer" and not every_rank_has_tiles:
            raise ValueError(f"n={n} leaves some of the {self.world} ranks without a tile: use exchange='nccl' (or fewer ranks)")
        self.exchange = exchange if self.world > 1 else "none"
        if self.exchange == "peer":
            handles = [None] * self.world
            dist.all_gather_object(handles, self.solver.comm_handle(), group=group)
            self.solver.comm_connect(self.rank, self.world, b"".join(handles))
            # rank r's tiles start in tile row I0(r): its partials of the nodes above are zeros, peers skip them
            firsts = (ctypes.c_int * self.world)()
            for r in range(self.world):
                rb, re = shard_tiles(n, r, self.world)
                row0 = ctypes.c_int()
                nat.check(nat.lib().gl_shard_rows(n, rb, re, ctypes.byref(row0), None, None, None))
                firsts[r] = row0.value if re > rb else 0
            nat.check(nat.lib().gl_solver_comm_first_nodes(self.solver._h, firsts))
            dist.barrier(group=group)
        self.launches = 0
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._graph_iters = 0
        self.maxit = 0

    # -- data ---------------------------------------------------------------------------------------
    def load_packed(self, z_packed: torch.Tensor, z_scale=None, w_packed: Optional[torch.Tensor] = None):
        self.solver.load(z_packed, z_scale, w_packed)

    def load_tiles(self, z_tiles_full: torch.Tensor):
        """Copy this rank's contiguous slice out of a full tiled Z buffer (graph 0)."""
        tt = nat.TILE * nat.TILE
        b, e = self.tile_range
        from .smooth_learning import _wrap_device_memory

        dst = _wrap_device_memory(self.solver.z_tiles_ptr(0), (e - b) * tt, self.solver.dtype,
                                  self.solver.device, owner=self.solver)
        dst.copy_(z_tiles_full[b * tt:e * tt])
        w = _wrap_device_memory(self.solver.w_tiles_ptr(0), (e - b) * tt, self.solver.dtype,
                                self.solver.device, owner=self.solver)
        w.zero_()

    # -- schedule -------------------------------------------------------------------------------------
    def _allreduce(self):
        if self.exchange == "nccl":
            dist.all_reduce(self._xch, group=self.group)

    def begin(self, alpha=1.0, beta=1.0, step_size=0.5, maxit=1000, tol=1e-5):
        self.maxit = int(maxit)
        self.solver.begin(alpha, beta, step_size, maxit, tol, finish=False)
        self._allreduce()
        self.solver.begin_finish()
        self._graph = None  # parameters are baked into captured launches

    def _capture(self, iters: int):
        g = torch.cuda.CUDAGraph()
        # warm NCCL / lazy init outside capture
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            for _ in range(iters):
                self.solver.edge_pass()
                self._allreduce()
                self.solver.node_tail()
        self._graph, self._graph_iters = g, iters

    def run(self, poll_every: int = 0) -> None:
        """Replay the iteration graph until the device stop flag fires (identical on all ranks,
        since every rank evaluates the same all-reduced sums)."""
        if self.maxit <= 0:
            return
        if self.exchange != "nccl":
            # no collective call in the loop: the library's own graph + pipelined polling drives it
            before = nat.launch_count()
            self.solver.run(poll_every)
            self.launches += nat.launch_count() - before
            return
        if poll_every <= 0:
            poll_every = 10
        poll_every = min(poll_every, self.maxit)
        if self._graph is None or self._graph_iters != poll_every:
            self._capture(poll_every)
        per_iter = 2 + (1 if self.solver.K > 1 else 0)
        done_it = 0
        while done_it < self.maxit:
            self._graph.replay()
            done_it += poll_every
            self.launches += per_iter * poll_every
            st = self.solver.status()
            if st.n_iter >= self.maxit or st.converged:
                break

    def status(self):
        return self.solver.status()

    def gather_square(self, edge_tol: float = 1e-3, result_on: str = "all") -> Optional[torch.Tensor]:
        """The reference's return value of the 1-D call (:171-172): the dense symmetric (n, n) matrix,
        on every rank ("all") or on rank 0 only ("root": the others return None).  Tile shards are
        contiguous slices of one tiled buffer, so the gather is one NCCL transfer per rank (straight
        into the root's buffer, or an all-reduce of disjoint slices); the square form is written by
        ``gl_tiles_to_square`` with the edge_tol threshold (:169)."""
        if result_on not in ("all", "root"):
            raise ValueError("result_on must be 'all' or 'root'")
        lib = nat.lib()
        s = self.solver
        tt = nat.TILE * nat.TILE
        if self.world == 1:
            return s.square(edge_tol)
        total = lib.gl_num_tiles(self.n)
        mine = s.w_view(0)
        b, e = self.tile_range
        full = None
        if result_on == "all":
            full = torch.zeros(total * tt, dtype=s.dtype, device=s.device)
            full[b * tt:e * tt].copy_(mine)
            dist.all_reduce(full, group=self.group)
        else:
            root = dist.get_global_rank(self.group, 0) if self.group is not None else 0
            if self.rank == 0:
                full = torch.empty(total * tt, dtype=s.dtype, device=s.device)
                full[b * tt:e * tt].copy_(mine)
                ops = []
                for r in range(1, self.world):
                    rb, re = shard_tiles(self.n, r, self.world)
                    if re > rb:
                        peer = dist.get_global_rank(self.group, r) if self.group is not None else r
                        ops.append(dist.P2POp(dist.irecv, full[rb * tt:re * tt], peer, group=self.group))
            else:
                ops = [dist.P2POp(dist.isend, mine, root, group=self.group)] if e > b else []
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
        if full is None:
            return None
        out = torch.empty((self.n, self.n), dtype=s.dtype, device=s.device)
        nat.check(lib.gl_tiles_to_square(ctypes.c_void_p(full.data_ptr()), ctypes.c_void_p(out.data_ptr()), self.n,
                                         nat.GL_F32 if s.dtype == torch.float32 else nat.GL_F64, float(edge_tol),
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out

    def gather_packed(self, edge_tol: float = 1e-3) -> torch.Tensor:
        """All shards' weights as one packed (K, E) tensor on every rank (sum of disjoint stores)."""
        out = self.solver.store(edge_tol)
        if self.world > 1:
            dist.all_reduce(out, group=self.group)
        return out

    def close(self):
        self._graph = None
        if self.exchange == "peer":
            torch.cuda.synchronize()
            dist.barrier(group=self.group)  # nobody unmaps memory a peer may still read
        self.solver.close()


# ----------------------------------------------------------------------------------------------------
# the reference-facing distributed call (gsp_learn_graph_log_degrees(..., distributed=...))
# ----------------------------------------------------------------------------------------------------
_workspaces: dict = {}


def sharded_workspace(n: int, dtype, group=None) -> ShardedLogModelSolver:
    """One cached :class:`ShardedLogModelSolver` per (n, dtype, group): creating one costs device
    allocations plus, in peer mode, a CUDA IPC handle exchange and a barrier, which a caller in a loop
    (bench, GLMM-style outer iterations) should not pay per solve.  Collective: every rank of the group
    must ask for the same workspace in the same order."""
    key = (int(n), dtype, id(group) if group is not None else None)
    ws = _workspaces.get(key)
    if ws is None:
        for old in list(_workspaces):          # one resident workspace at a time (they are n^2-sized)
            _workspaces.pop(old).close()
        ws = ShardedLogModelSolver(n, dtype, group=group)
        _workspaces[key] = ws
    return ws


def release_workspaces() -> None:
    """Free the cached distributed workspace (collective in peer mode: ends with a barrier)."""
    for key in list(_workspaces):
        _workspaces.pop(key).close()


def solve_distributed(distances, alpha, beta, edge_init, maxit, tol, step_size, edge_tol, dtype, group,
                      result_on, out, return_info, z_scale=None):
    """1-D ``gsp_learn_graph_log_degrees`` over all ranks of ``group`` (see its docstring).  Every rank
    passes the same packed (E,) distances (host array, CPU tensor -- copied asynchronously when pinned --
    or CUDA tensor) and moves only the rows of its own tile shard to its device."""
    from .smooth_learning import _n_from_edges, _to_device
    if not dist.is_initialized():
        raise RuntimeError("distributed=True needs an initialised torch.distributed process group "
                           "(one process per GPU, e.g. under torchrun)")
    is_np = not isinstance(distances, torch.Tensor)
    is_host_tensor = isinstance(distances, torch.Tensor) and not distances.is_cuda
    n = _n_from_edges(int(distances.shape[0]))
    ws = sharded_workspace(n, dtype, group)
    _, _, e0, e1 = ws.solver.shard_rows()
    z_rows = _to_device(distances[e0:e1], dtype)
    w_rows = None
    if edge_init is not None:
        if tuple(edge_init.shape) != tuple(distances.shape):
            raise ValueError(f"edge_init has shape {tuple(edge_init.shape)}, expected {tuple(distances.shape)}")
        w_rows = _to_device(edge_init[e0:e1], dtype)
    if z_scale is not None:
        z_scale = float(np.asarray(z_scale, dtype=np.float64).ravel()[0])
    ws.solver.load_rows(z_rows, z_scale, w_rows)
    ws.begin(alpha, beta, step_size, maxit, tol)
    ws.run()
    info = ws.status()
    result = ws.gather_square(edge_tol, result_on)
    res = result
    if result is not None:
        if out is not None:
            out_t = torch.from_numpy(out) if isinstance(out, np.ndarray) else out
            out_t.copy_(result, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            res = out
        elif is_np:
            res = result.cpu().numpy()
        elif is_host_tensor:
            res = result.cpu()
    return (res, info) if return_info else res


# ----------------------------------------------------------------------------------------------------
# kNN-support variant (config 5): rows sharded, (a, b) slices all-gathered, 4 scalars all-reduced
# ----------------------------------------------------------------------------------------------------
def shard_rows(n: int, rank: int, world: int) -> tuple[int, int, int]:
    """Equal contiguous node ranges: (row_begin, row_end, rows_per_rank); the last ranks' ranges
    may run past n (padding) so that every slice of the all-gather has the same length."""
    per = -(-int(n) // int(world))
    per = -(-per // 128) * 128   # whole 128-node blocks keep the slices aligned with the node padding
    return rank * per, (rank + 1) * per, per


class KnnShardBackend(Protocol):
    def begin(self) -> None: ...        # initial degrees + node step of the own rows
    def edge_step(self) -> None: ...    # slot sweep over the own rows' slots
    def node_step(self) -> None: ...    # node update of the own rows, own (a, b) slice, scalar partials
    def decide(self, init: bool) -> None: ...
    def ab(self): ...                   # full (world * rows_per_rank, 2) array; own slice is current
    def scalars(self): ...              # (4,) float64 partial sums
    def done(self) -> bool: ...


def knn_exchange(backend: KnnShardBackend, rank: int, world: int, group=None) -> None:
    """What crosses ranks per iteration: the in-place all-gather of the (a, b) slices and the
    all-reduce of the four residual scalars.  Nothing edge-sized moves."""
    if world == 1:
        return
    full = backend.ab()
    per = full.shape[0] // world
    dist.all_gather_into_tensor(full, full[rank * per:(rank + 1) * per], group=group)
    dist.all_reduce(backend.scalars(), group=group)


def knn_sharded_schedule(backend: KnnShardBackend, maxit: int, rank: int, world: int, group=None,
                         poll_every: int = 1) -> None:
    """begin -> exchange -> decide(init) -> [edge step -> node step -> exchange -> decide]*"""
    backend.begin()
    knn_exchange(backend, rank, world, group)
    backend.decide(True)
    it = 0
    while it < maxit:
        for _ in range(min(poll_every, maxit - it)):
            backend.edge_step()
            backend.node_step()
            knn_exchange(backend, rank, world, group)
            backend.decide(False)
            it += 1
        if backend.done():
            break


KNN_EXCHANGES = ("peer", "nccl")


def knn_shard_labels(key: torch.Tensor, ei: torch.Tensor, ej: torch.Tensor, n: int, world: int) -> torch.Tensor:
    """New node numbers (int32, old -> new) for a row-sharded kNN solve: ascending ``key`` first (Morton
    codes: spatially compact shards, so that most far ends of a rank's slots are its own rows), then inside
    every rank's row range a stable sort by decreasing degree (rows of one 32-row slice get similar lengths:
    no padding; rows of equal degree keep the key order) -- what the library does itself for a single
    full-range shard.  Plain torch ops on the tensors' device (setup-time glue)."""
    dev = key.device
    arrangement = torch.argsort(key.to(torch.int64), stable=True)
    deg = torch.bincount(ei.long(), minlength=n) + torch.bincount(ej.long(), minlength=n)
    per = shard_rows(n, 0, world)[2]
    pos = torch.arange(n, device=dev, dtype=torch.int64)
    dmax = int(deg.max()) if n else 0
    composite = (pos // per) * (dmax + 1) + (dmax - deg[arrangement])
    arrangement = arrangement[torch.argsort(composite, stable=True)]
    lab = torch.empty(n, dtype=torch.int32, device=dev)
    lab[arrangement] = torch.arange(n, dtype=torch.int32, device=dev)
    return lab


class ShardedKNNSolver:
    """CUDA backend of the row-sharded kNN-support solve (``gl_knn_*``): one process per GPU, each
    rank builds the directed slots of its own rows from the edges touching them; iterations
    (slot sweep, node step, NCCL all-gather + all-reduce, decision) are captured in a CUDA graph."""

    def __init__(self, ei, ej, z, n_nodes: int, dtype=torch.float32, edge_init=None, group=None, rank=None,
                 world=None, exchange: str = "auto", order_key=None):
        """``rank`` / ``world`` override the process group's (single-process emulation of several
        shards in the tests; the caller then performs the exchange itself).

        ``exchange``: "peer" = the whole iteration is ONE launch per rank and the (gamma d, gamma pb)
        slices, the residual partials and the sequence flags travel by stores into peer-mapped memory
        over NVLink (``gl_knn_comm_*``; NCCL only carries the one initial all-gather of a solve);
        "nccl" = slot sweep, node step, in-place ``ncclAllGather`` + ``ncclAllReduce`` and a decision
        launch per iteration, all inside a captured graph; "auto" = peer for 2..8 ranks.

        ``order_key`` (per-node non-negative integers, e.g. :func:`~graph_learning_b200.knn.locality_keys`):
        the nodes are renumbered by ascending key before the rows are cut into contiguous shards, so that
        a rank's far ends are mostly its own rows (setup-time torch glue; results stay per edge of the
        caller's list)."""
        from .smooth_learning import _device, _gl_dtype, _to_device
        self.lib = nat.lib()
        self.group = group
        if exchange not in KNN_EXCHANGES + ("auto",):
            raise ValueError(f"exchange must be one of {KNN_EXCHANGES + ('auto',)}")
        self.emulated = world is not None
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.rank = rank if rank is not None else (dist.get_rank(group) if dist.is_initialized() else 0)
        self.n, self.dtype = int(n_nodes), dtype
        if exchange == "auto":
            exchange = "peer" if (not self.emulated and 2 <= self.world <= 8) else "nccl"
        if exchange == "peer" and (self.emulated or self.world > 8):
            raise ValueError("the peer exchange needs one process per GPU and at most 8 ranks")
        self.exchange = exchange if self.world > 1 else ("nccl" if self.emulated else "single")
        dev = _device()
        ei = torch.as_tensor(ei, device=dev).to(torch.int32).contiguous()
        ej = torch.as_tensor(ej, device=dev).to(torch.int32).contiguous()
        key32 = None
        if order_key is not None and self.exchange == "single":
            # one GPU: the library arranges its rows itself (Morton key, stable global degree sort)
            key32 = torch.as_tensor(order_key, device=dev).to(torch.int32).contiguous()
        elif order_key is not None:
            key = torch.as_tensor(order_key, device=dev).to(torch.int64)
            if key.numel() != self.n:
                raise ValueError("order_key must hold one key per node")
            lab = knn_shard_labels(key, ei, ej, self.n, self.world)
            a, b = lab[ei.long()], lab[ej.long()]
            ei, ej = torch.minimum(a, b).contiguous(), torch.maximum(a, b).contiguous()
        z = _to_device(z, dtype)
        self.nnz_global = int(z.numel())
        dm = ctypes.c_int()
        nat.check(self.lib.gl_knn_max_degree(ei.data_ptr(), ej.data_ptr(), self.nnz_global, self.n, ctypes.byref(dm), None))
        self.dmax = dm.value
        self.row_begin, self.row_end, self.per = shard_rows(self.n, self.rank, self.world)
        # only the edges touching this rank's rows are kept (torch glue; setup)
        touch = ((ei >= self.row_begin) & (ei < self.row_end)) | ((ej >= self.row_begin) & (ej < self.row_end))
        self.local_ids = touch.nonzero(as_tuple=True)[0]
        self._ei, self._ej = ei[touch].contiguous(), ej[touch].contiguous()
        self._z = z[touch].contiguous()
        self._w0 = None if edge_init is None else _to_device(edge_init, dtype)[touch].contiguous()
        h = ctypes.c_void_p()
        nat.check(self.lib.gl_knn_create_ordered(
            ctypes.byref(h), self._ei.data_ptr(), self._ej.data_ptr(), self._z.data_ptr(),
            None if self._w0 is None else self._w0.data_ptr(), int(self._z.numel()), self.n, _gl_dtype(dtype),
            self.row_begin, self.row_end, self.per * self.world, None if key32 is None else key32.data_ptr(), 0, 0, None))
        self._h = h
        if self.exchange == "single":
            nat.check(self.lib.gl_knn_set_fused(self._h, 2))   # whole iteration in one launch (knn_fused_kernel)
        from .smooth_learning import _wrap_device_memory
        cnt = ctypes.c_int64()
        ab_ptr = self.lib.gl_knn_ab_ptr(self._h, ctypes.byref(cnt))
        self._ab = _wrap_device_memory(ab_ptr, 2 * cnt.value, dtype, dev, owner=self).view(cnt.value, 2)[:self.per * self.world]
        self._scal = _wrap_device_memory(self.lib.gl_knn_scalars_ptr(self._h), 4, torch.float64, dev, owner=self)
        self._graph = None
        self.launches = 0
        if self.exchange == "peer":
            buf = ctypes.create_string_buffer(64)
            nat.check(self.lib.gl_knn_comm_handle(self._h, buf))
            handles = [None] * self.world
            dist.all_gather_object(handles, buf.raw, group=group)
            nat.check(self.lib.gl_knn_comm_connect(self._h, self.rank, self.world, b"".join(handles)))
            # who reads which node: a pair is pushed only to the ranks whose slots reference it
            need = torch.empty(self.n, dtype=torch.int32, device=dev)
            nat.check(self.lib.gl_knn_mark_far_ends(self._h, need.data_ptr(), None))
            need <<= self.rank
            dist.all_reduce(need, group=group)
            nat.check(self.lib.gl_knn_comm_need(self._h, need.data_ptr(), None))
            torch.cuda.synchronize()
            self.pushed_fraction = float(((need & ~(1 << self.rank))[self.row_begin:min(self.row_end, self.n)] != 0)
                                         .double().mean()) if self.row_begin < self.n else 0.0
            dist.barrier(group=group)

    # -- backend protocol ---------------------------------------------------------------------------------
    def begin(self, alpha=1.0, beta=1.0, step_size=0.5, maxit=1000, tol=1e-5):
        self.maxit = int(maxit)
        nat.check(self.lib.gl_knn_begin(self._h, float(alpha), float(beta), float(step_size), int(maxit), float(tol),
                                        self.dmax, _cur_stream()))
        params = (float(alpha), float(beta), float(step_size), int(maxit), float(tol))
        if getattr(self, "_graph_params", None) != params:
            self._graph = None            # the parameters are baked into the captured launches
            self._graph_params = params
        if self.emulated:
            return   # the caller exchanges and calls decide(True)
        knn_exchange(self, self.rank, self.world, self.group)
        self.decide(True)

    def edge_step(self):
        nat.check(self.lib.gl_knn_edge_step(self._h, _cur_stream()))

    def node_step(self):
        nat.check(self.lib.gl_knn_node_step(self._h, _cur_stream()))

    def decide(self, init: bool):
        nat.check(self.lib.gl_knn_decide(self._h, 1 if init else 0, _cur_stream()))

    def ab(self):
        return self._ab

    def scalars(self):
        return self._scal

    def status(self):
        from .smooth_learning import SolveInfo
        it, cv = ctypes.c_int(), ctypes.c_int()
        rp, rd = ctypes.c_double(), ctypes.c_double()
        nat.check(self.lib.gl_knn_status(self._h, ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd), ctypes.byref(cv),
                                         _cur_stream()))
        return SolveInfo(it.value, rp.value, rd.value, bool(cv.value))

    def done(self) -> bool:
        st = self.status()
        return st.converged or st.n_iter >= self.maxit

    # -- driver ----------------------------------------------------------------------------------------------
    def _capture(self, iters: int):
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            for _ in range(iters):
                self.edge_step()
                if self.exchange in ("peer", "single"):
                    continue          # the launch above was the whole iteration (exchange included)
                self.node_step()
                knn_exchange(self, self.rank, self.world, self.group)
                self.decide(False)
        self._graph, self._graph_iters = g, iters

    def run(self, poll_every: int = 20) -> None:
        if self.maxit <= 0:
            return
        peer = self.exchange == "peer"
        one_launch = self.exchange in ("peer", "single")
        poll_every = max(1, min(poll_every, self.maxit))
        if one_launch:
            poll_every += poll_every & 1      # the two (a, b) buffers swap roles every launch: even count per graph
        if self._graph is None or self._graph_iters != poll_every:
            self._capture(poll_every)
        # peer mode: a launch applies the stop rule to the iteration before it, so one more launch than
        # iterations is enqueued (it only decides; launches after the stop are no-ops on the device)
        total = self.maxit + 1 if peer else self.maxit
        n_graphs = -(-total // poll_every)
        # pipelined polling: replay g + 1 is enqueued before the status words of replay g are looked at
        # (asynchronous copy into pinned memory + an event), so the device never waits for the host
        if getattr(self, "_flags_host", None) is None:
            from .smooth_learning import _wrap_device_memory
            self._flags_host = torch.zeros((2, 8), dtype=torch.int32).pin_memory()
            self._flags_dev = _wrap_device_memory(self.lib.gl_knn_flags_ptr(self._h), 8, torch.int32, self._z.device,
                                                  owner=self)
            self._poll_ev = [torch.cuda.Event(), torch.cuda.Event()]
        launched = checked = 0
        while True:
            while launched < n_graphs and launched - checked < 2:
                self._graph.replay()
                self._flags_host[launched & 1].copy_(self._flags_dev, non_blocking=True)
                self._poll_ev[launched & 1].record()
                launched += 1
                self.launches += (1 if one_launch else 3) * poll_every
            if checked >= launched:
                break
            self._poll_ev[checked & 1].synchronize()
            done = bool(self._flags_host[checked & 1, 0].item())
            checked += 1
            if done:
                break
        torch.cuda.current_stream().synchronize()

    def gather(self, edge_tol: float = 1e-3) -> torch.Tensor:
        """Weights over the GLOBAL edge list on every rank (each edge is written by the owner of its
        first end; sum of disjoint stores)."""
        loc = torch.zeros_like(self._z)
        nat.check(self.lib.gl_knn_store(self._h, loc.data_ptr(), float(edge_tol), _cur_stream()))
        out = torch.zeros(self.nnz_global, dtype=self._z.dtype, device=self._z.device)
        out[self.local_ids] = loc
        if self.world > 1 and not self.emulated:
            dist.all_reduce(out, group=self.group)
        return out

    def close(self):
        self._graph = None
        if getattr(self, "_h", None):
            torch.cuda.synchronize()
            if self.exchange == "peer":
                dist.barrier(group=self.group)   # nobody unmaps memory a peer may still write
            self.lib.gl_knn_destroy(self._h)
            self._h = None


def _cur_stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
