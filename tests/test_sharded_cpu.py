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


# ---- kNN-support variant: rows sharded, (a, b) slices all-gathered, 4 scalars all-reduced -------------
class HostKnnShard:
    """NumPy stand-in for the CUDA ``gl_knn_*`` solver of one rank: directed slots of the own rows,
    node state of the own rows only, the interleaved (gamma d, gamma pb) array as exchanged."""

    def __init__(self, ei, ej, z, n, rb, re, per, world, dmax, maxit, tol=1e-5, alpha=1.0, beta=1.0, step=0.5):
        self.n, self.rb, self.re, self.maxit, self.tol, self.alpha = n, rb, min(re, n), maxit, tol, alpha
        src = np.concatenate([ei, ej])
        dst = np.concatenate([ej, ei])
        eid = np.concatenate([np.arange(ei.size)] * 2)
        own = (src >= rb) & (src < re)
        order = np.argsort(src[own], kind="stable")
        self.src, self.dst, self.eid = src[own][order], dst[own][order], eid[own][order]
        self.z, self.w = z[self.eid], np.zeros(self.eid.size)
        self.first_end = ei[self.eid] == self.src        # the copy whose source is the edge's first end
        self.g = step / (2 * beta + np.sqrt(2 * dmax))
        self.c = 1 - 2 * self.g * beta
        self.abv = torch.zeros((per * world, 2), dtype=torch.float64)
        self.scal = torch.zeros(4, dtype=torch.float64)
        self.d = np.zeros(n)
        self.yb = np.zeros(n)
        self.pb = np.zeros(n)
        self.it, self.stop = 0, False
        self.gw_prev = self.gd_prev = 0.0

    def ab(self):
        return self.abv

    def scalars(self):
        return self.scal

    def _rows(self, v):   # sums over the own rows' slots, by source
        return np.bincount(self.src, v, self.n)

    def _node(self, sp, sw, init):
        r = slice(self.rb, self.re)
        if init:
            dn, rbv = sw[r].copy(), np.zeros(self.re - self.rb)
        else:
            qb = self.pb[r] + self.g * sp[r]
            rbv = qb - self.yb[r]
            dn = (self.d[r] - self.yb[r]) + qb
        self.d[r] = dn
        self.yb[r] = dn + self.g * sw[r]
        self.pb[r] = (self.yb[r] - np.sqrt(self.yb[r] ** 2 + 4 * self.alpha * self.g)) / 2
        a = self.abv.numpy()
        a[r, 0], a[r, 1] = self.g * dn, self.g * self.pb[r]
        return np.sum(rbv ** 2), np.sum(dn ** 2)

    def begin(self):
        sw = self._rows(self.w)
        rb2, d2 = self._node(None, sw, True)
        self.scal[:] = torch.tensor([0.0, 0.5 * np.sum(self.w ** 2), rb2, d2])

    def edge_step(self):
        if self.stop:
            return
        a = self.abv.numpy()
        y = self.c * self.w - (a[self.src, 0] + a[self.dst, 0])
        p = np.maximum(y - 2 * self.g * self.z, 0)
        q = self.c * p - (a[self.src, 1] + a[self.dst, 1])
        self.w = (self.w - y) + q
        self._sp, self._sw = self._rows(p), self._rows(self.w)
        self._gr, self._gw = 0.5 * np.sum((q - y) ** 2), 0.5 * np.sum(self.w ** 2)

    def node_step(self):
        if self.stop:
            return
        rb2, d2 = self._node(self._sp, self._sw, False)
        self.scal[:] = torch.tensor([self._gr, self._gw, rb2, d2])

    def decide(self, init):
        g_r, g_w, r2, d2 = self.scal.tolist()
        if not init and not self.stop:
            conv = False
            if self.it >= 1:
                conv = np.sqrt(g_r) / np.sqrt(self.gw_prev) < self.tol and np.sqrt(r2) / np.sqrt(self.gd_prev) < self.tol
            self.it += 1
            self.stop = conv or self.it >= self.maxit
        elif init:
            self.stop = self.maxit <= 0
        self.gw_prev, self.gd_prev = g_w, d2

    def done(self):
        return self.stop


def _knn_problem(n, k):
    rng = np.random.default_rng(n)
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    return ei, ej, zsq[ei, ej] * (n ** (2 / 3))


def _knn_worker(rank, world, port, n, k, maxit, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from graph_learning_b200.sharded import knn_sharded_schedule, shard_rows
    ei, ej, z = _knn_problem(n, k)
    dmax = int(np.max(np.bincount(ei, minlength=n) + np.bincount(ej, minlength=n)))
    rb, re, per = shard_rows(n, rank, world)
    sh = HostKnnShard(ei, ej, z, n, rb, re, per, world, dmax, maxit)
    knn_sharded_schedule(sh, maxit, rank, world, poll_every=5)
    full = torch.zeros(z.size, dtype=torch.float64)
    keep = sh.first_end
    full[torch.as_tensor(sh.eid[keep])] = torch.as_tensor(sh.w[keep])
    dist.all_reduce(full)
    if rank == 0:
        q.put((full.numpy(), sh.it, (rb, re, per)))
    dist.destroy_process_group()


@pytest.mark.parametrize("n,k,maxit", [(300, 5, 60), (200, 4, 1000)])
def test_knn_sharded_schedule_gloo_world2(n, k, maxit):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_knn_worker, args=(r, 2, port, n, k, maxit, q)) for r in range(2)]
    for p in procs:
        p.start()
    w, it, (rb, re, per) = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ei, ej, z = _knn_problem(n, k)
    wo, info = orc.solve_knn(ei, ej, z, n, maxit=maxit, threshold=False)
    assert it == info.n_iter
    assert np.linalg.norm(w - wo) / np.linalg.norm(wo) < 1e-12
    assert rb == 0 and re == per and per % 128 == 0 and 2 * per >= n


def test_shard_rows_cover_all_nodes():
    from graph_learning_b200.sharded import shard_rows
    for n, world in [(1_000_000, 8), (300, 2), (129, 4), (50, 8)]:
        prev = 0
        for r in range(world):
            b, e, per = shard_rows(n, r, world)
            assert b == prev and e - b == per
            prev = e
        assert prev >= n
