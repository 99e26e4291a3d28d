"""World-size-2 gloo test (CPU, no GPU) of the multi-GPU schedule in graph_learning_b200/sharded.py:
begin -> all-reduce -> finish -> [edge pass -> all-reduce -> node tail]*, with a host emulation of
the per-shard kernels built from the oracle's formulas over each rank's tile range.  Checks that
(a) the shard ranges partition the triangle, (b) every rank ends with identical replicated node
state, (c) the gathered weights equal the single-process oracle solve."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import logmodel_oracle as orc

T = 128


class HostShard:
    """NumPy stand-in for the CUDA Solver over tiles [tb, te): same exchange-buffer layout
    [S p | S w' | ||q-y||^2 | ||w'||^2] and the same node-tail arithmetic (node_tail.cu)."""

    def __init__(self, z_packed, n, tb, te, maxit, tol=1e-5, alpha=1.0, beta=1.0, step=0.5):
        self.n, self.maxit, self.tol, self.alpha = n, maxit, tol, alpha
        nt = -(-n // T)
        tiles = [(i, j) for i in range(nt) for j in range(i, nt)][tb:te]
        ei, ej = orc.edge_index(n)
        tid = {t: k for k, t in enumerate(tiles)}
        mine = np.array([(a // T, b // T) in tid for a, b in zip(ei, ej)])
        self.idx = np.where(mine)[0]
        self.ei, self.ej = ei[mine], ej[mine]
        self.z = z_packed[mine]
        self.w = np.zeros_like(self.z)
        self.g = step / (2 * beta + np.sqrt(2 * (n - 1)))
        self.c = 1 - 2 * self.g * beta
        self.xch = torch.zeros(2 * n + 2, dtype=torch.float64)
        self.it, self.stop = 0, False

    def _deg(self, v):
        return np.bincount(self.ei, v, self.n) + np.bincount(self.ej, v, self.n)

    def exchange(self):
        return self.xch

    def begin(self):
        x = self.xch.numpy()
        x[:] = 0
        x[self.n:2 * self.n] = self._deg(self.w)
        x[2 * self.n + 1] = np.sum(self.w ** 2)

    def _prox(self, sw):
        self.yb = self.d + self.g * sw
        self.pb = (self.yb - np.sqrt(self.yb ** 2 + 4 * self.alpha * self.g)) / 2

    def begin_finish(self):
        x = self.xch.numpy()
        self.d = x[self.n:2 * self.n].copy()
        self.gw_prev, self.gd_prev = x[2 * self.n + 1], np.sum(self.d ** 2)
        self._prox(x[self.n:2 * self.n])
        x[:] = 0

    def edge_pass(self):
        if self.stop:
            return
        g, c = self.g, self.c
        y = c * self.w - g * (self.d[self.ei] + self.d[self.ej])
        p = np.maximum(y - 2 * g * self.z, 0)
        q = c * p - g * (self.pb[self.ei] + self.pb[self.ej])
        self.w = (self.w - y) + q
        x = self.xch.numpy()
        x[:self.n] += self._deg(p)
        x[self.n:2 * self.n] += self._deg(self.w)
        x[2 * self.n] += np.sum((q - y) ** 2)
        x[2 * self.n + 1] += np.sum(self.w ** 2)

    def node_tail(self):
        if self.stop:
            return
        x = self.xch.numpy()
        sp, sw = x[:self.n].copy(), x[self.n:2 * self.n].copy()
        qb = self.pb + self.g * sp
        rb = qb - self.yb
        d_new = (self.d - self.yb) + qb
        conv = False
        if self.it >= 1:
            rp = np.sqrt(x[2 * self.n]) / np.sqrt(self.gw_prev)
            rd = np.sqrt(np.sum(rb ** 2)) / np.sqrt(self.gd_prev)
            conv = rp < self.tol and rd < self.tol
        self.gw_prev, self.gd_prev = x[2 * self.n + 1], np.sum(d_new ** 2)
        self.d = d_new
        self._prox(sw)
        x[:] = 0
        self.it += 1
        self.stop = conv or self.it >= self.maxit

    def done(self):
        return self.stop


def _worker(rank, world, port, n, maxit, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from graph_learning_b200.sharded import shard_tiles, sharded_schedule
    rng = np.random.default_rng(3)
    pts = rng.uniform(size=(n, 2))
    z = orc.to_packed(((pts[:, None] - pts[None]) ** 2).sum(-1)) * 30.0
    tb, te = shard_tiles(n, rank, world)
    sh = HostShard(z, n, tb, te, maxit)
    sharded_schedule(sh, maxit, poll_every=7)
    full = torch.zeros(z.size, dtype=torch.float64)
    full[torch.as_tensor(sh.idx)] = torch.as_tensor(sh.w)
    dist.all_reduce(full)
    dsum = torch.as_tensor(sh.d.copy())
    dall = [torch.zeros_like(dsum) for _ in range(world)]
    dist.all_gather(dall, dsum)
    if rank == 0:
        q.put((full.numpy(), sh.it, [t.numpy() for t in dall], (tb, te)))
    dist.destroy_process_group()


@pytest.mark.parametrize("n,maxit", [(300, 40), (260, 1000)])
def test_sharded_schedule_gloo_world2(n, maxit):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, maxit, q)) for r in range(2)]
    for p in procs:
        p.start()
    w, it, dall, (tb, te) = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(3)
    pts = rng.uniform(size=(n, 2))
    z = orc.to_packed(((pts[:, None] - pts[None]) ** 2).sum(-1)) * 30.0
    wo, info = orc.solve_one_pass(z, maxit=maxit, threshold=False)
    assert it == info.n_iter
    assert np.linalg.norm(w - wo) / np.linalg.norm(wo) < 1e-12
    assert np.array_equal(dall[0], dall[1])      # replicated node state is bit-identical
    assert tb == 0 and 0 < te
