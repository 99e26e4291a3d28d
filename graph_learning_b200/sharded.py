"""Multi-GPU LogModel solve: the triangle's 128x128 tiles are split into contiguous,
work-balanced shards, one per rank (one process per GPU, ``torch.distributed``); every rank keeps
its shard of w and z resident and the node vectors replicated.  Per iteration the only exchange is
ONE all-reduce of the float64 exchange buffer [S p | S w' | residual Grams] (2nK + 2K^2 values,
320 KB at n = 20,000) between the edge pass and the node tail.

The schedule is written against a tiny backend protocol so that the same driver runs the CUDA
``Solver`` over NCCL and, in the CPU test-suite, a host emulation over gloo.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Protocol

import numpy as np
import torch
import torch.distributed as dist

from . import _native as nat

__all__ = ["shard_tiles", "ShardBackend", "sharded_schedule", "ShardedLogModelSolver"]


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
        self.tile_range = shard_tiles(n, self.rank, self.world)
        self.solver = Solver(n, n_graphs, dtype, tile_range=self.tile_range)
        self._xch = self.solver.exchange()
        if exchange == "auto":
            exchange = "peer" if (self.world > 1 and n_graphs == 1 and self.world <= 8) else "nccl"
        self.exchange = exchange if self.world > 1 else "none"
        if self.exchange == "peer":
            handles = [None] * self.world
            dist.all_gather_object(handles, self.solver.comm_handle(), group=group)
            self.solver.comm_connect(self.rank, self.world, b"".join(handles))
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
