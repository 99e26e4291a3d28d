"""The kNN iteration kernels EXECUTED ON THE CPU from their own source text.

`knn_blocked_kernel` was written after the round's GPU minutes were spent.  To run it at least once
before its first device launch, this test cuts the kernel sources (`knn_fused_kernel`,
`knn_blocked_kernel` and their argument structs) out of graph_learning_b200/csrc/knn_kernels.cuh,
compiles them with g++ against tests/cuda_emu/shim.h (one OS thread per CUDA thread, barriers for
__syncthreads and warp shuffles, CTAs one after another) and drives them from NumPy with the layouts of
oracle.knn_blocked_layout.  `knn_fused_kernel` -- validated on the B200 -- goes through the same
harness first, which checks the shim; then the blocked kernel must produce the same iterates as
oracle.solve_knn.  This checks the kernel's logic (indexing, barriers, shared-memory accumulation,
residual bookkeeping, double buffering), not its memory-model behaviour or speed on a GPU.
"""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "graph_learning_b200", "csrc")
EMU = os.path.join(ROOT, "tests", "cuda_emu")


def _extract():
    src = open(os.path.join(CSRC, "knn_kernels.cuh"), encoding="utf-8").read()
    flags = open(os.path.join(CSRC, "solver.cuh"), encoding="utf-8").read()
    enum = re.search(r"enum \{\s*FLAG_DONE.*?\};", flags, re.S).group(0)
    consts = re.search(r"constexpr int KNN_THREADS.*?template <> struct Vec2<double>[^\n]*\n", src, re.S).group(0)
    for name in ("knn_expand_kernel", "knn_window_key_kernel"):      # two build kernels from the top of the file
        consts += re.search(r"static __global__ void " + name + r"\(.*?\n}\n", src, re.S).group(0)
    a = src.index("template <typename F> struct KnnFusedArgs {")
    b = src.index("// Node phase of the kNN path")
    return "#pragma once\n" + enum + "\nnamespace gl {\n" + consts + src[a:b] + "\n}  // namespace gl\n"


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    build = tmp_path_factory.mktemp("cuda_emu")
    (build / "extracted.inc").write_text(_extract())
    so = build / "libknn_emu.so"
    cmd = ["g++", "-std=c++20", "-O1", "-shared", "-fPIC", "-pthread", "-I", str(build), "-I", EMU,
           os.path.join(EMU, "harness.cpp"), "-o", str(so)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-4000:]
    lib = ctypes.CDLL(str(so))

    class Checked:
        """Every harness call is followed by a look at the 'threads refused' flag: 256 OS threads per
        emulated CTA; if the host refuses them the harness stops cleanly and the test is skipped instead
        of comparing half-run kernels."""

        def __getattr__(self, name):
            fn = getattr(lib, name)

            def call(*args):
                out = fn(*args)
                if lib.emu_threads_refused():
                    pytest.skip("the OS refused the emulation's threads")
                return out
            return call

    return Checked()


def _case(n, k, seed):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(size=(n, 3))
    zsq = ((pts[:, None] - pts[None]) ** 2).sum(-1)
    ei, ej = orc.knn_support(zsq, k)
    return pts, ei, ej, zsq[ei, ej] * (n ** (2 / 3))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _run(lib, kind, ei, ej, z, n, lay, iters, dtype, grid, w0=None, alpha=1.0, beta=1.0, step=0.5, tol=0.0):
    """Device-side state exactly as gl_knn_create / gl_knn_begin leave it, then `iters` kernel launches."""
    ft = np.dtype(dtype)
    npad = -(-n // 128) * 128
    nslices = npad // 32
    block = lay["block"]
    w_edge = np.zeros(len(ei)) if w0 is None else np.asarray(w0, dtype=np.float64)
    # SELL-32 arrays of the slots (row space), column-major inside a slice, padding dst = -1
    cnt = np.diff(lay["slot_ptr"])
    cnt = np.concatenate([cnt, np.zeros(npad - n, dtype=cnt.dtype)])
    width = cnt.reshape(-1, 32).max(axis=1)
    slice_off = np.concatenate([[0], np.cumsum(32 * width)]).astype(np.int64)
    total = int(slice_off[-1])
    dst = np.full(max(total, 1), -1, dtype=np.int32)
    zs = np.zeros(max(total, 1), dtype=ft)
    ws = np.zeros(max(total, 1), dtype=ft)
    src = np.repeat(np.arange(n), np.diff(lay["slot_ptr"]))
    t = np.arange(len(src)) - lay["slot_ptr"][src]
    pos = slice_off[src // 32] + 32 * t + (src % 32)
    dst[pos] = lay["slot_dst"]
    zs[pos] = z[lay["slot_eid"]]
    ws[pos] = w_edge[lay["slot_eid"]]
    # intra lists
    ids = np.stack([lay["intra_a"], lay["intra_b"]], axis=1).astype(np.uint8).copy()
    iz = z[lay["intra_eid"]].astype(ft)
    iw = w_edge[lay["intra_eid"]].astype(ft)
    nblocks = -(-npad // 256) if kind == 1 else 0
    iptr = np.zeros(max(nblocks, 1) + 1, dtype=np.int32)
    if kind == 1:
        assert block == 256
        iptr[:len(lay["intra_ptr"])] = lay["intra_ptr"]
        iptr[len(lay["intra_ptr"]):] = lay["intra_ptr"][-1]
    else:
        assert len(lay["intra_eid"]) == 0
    # node state after gl_knn_begin (smooth_learning.py:144, :148, :151)
    dmax = int(np.max(np.bincount(ei, minlength=n) + np.bincount(ej, minlength=n)))
    g = step / (2.0 * beta + np.sqrt(2.0 * dmax))
    c = 1.0 - 2.0 * g * beta
    lab = lay["lab"]
    d0 = np.zeros(npad)
    d0[:n] = np.bincount(lab[ei], w_edge, n) + np.bincount(lab[ej], w_edge, n)
    yb = d0 + g * d0
    pb = np.zeros(npad)
    pb[:n] = (yb[:n] - np.sqrt(yb[:n] ** 2 + 4.0 * alpha * g)) / 2.0
    yb[n:] = 0.0
    ab0 = np.zeros((npad, 2), dtype=ft)
    ab0[:n, 0] = g * d0[:n]
    ab0[:n, 1] = g * pb[:n]
    ab1 = np.zeros_like(ab0)
    d = d0.copy()
    part = np.zeros(4 * max(grid, 1))
    gprev = np.array([np.sum(w_edge ** 2), np.sum(d0 ** 2)])
    resid = np.full(2, np.inf)
    flags = np.zeros(8, dtype=np.int32)
    lib.emu_knn_run(ctypes.c_int(kind), ctypes.c_int(int(ft == np.float64)), ctypes.c_int(grid), _p(dst), _p(zs), _p(ws),
                    _p(slice_off), ctypes.c_int(nslices), ctypes.c_int(n), ctypes.c_int(iters), _p(ab0), _p(ab1), _p(d),
                    _p(yb), _p(pb), _p(part), _p(gprev), _p(resid), _p(flags), ctypes.c_double(c),
                    ctypes.c_double(2.0 * g), ctypes.c_double(g), ctypes.c_double(4.0 * alpha * g), ctypes.c_double(tol),
                    _p(iptr), _p(ids), _p(iz), _p(iw), ctypes.c_int(nblocks), ctypes.c_int(iters))
    w = np.empty(len(ei))
    w[lay["intra_eid"]] = iw
    first = lay["slot_first"]
    w[lay["slot_eid"][first]] = ws[pos[first]]
    # both copies of an inter edge stay bit-identical
    assert np.array_equal(ws[pos[first]][np.argsort(lay["slot_eid"][first])],
                          ws[pos[~first]][np.argsort(lay["slot_eid"][~first])])
    return w, flags, resid


def _rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("grid", [1, 3])
def test_shim_runs_the_validated_fused_kernel(emu, grid):
    """knn_fused_kernel is green on the B200; through the shim it must give the oracle's iterates too."""
    n, iters = 300, 6
    pts, ei, ej, z = _case(n, 6, 7)
    order = orc.knn_row_order(ei, ej, n, None, 0)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block=1)        # block of one row: no intra edges
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=iters, tol=0.0, threshold=False)
    for kind in (0, 2):                                            # plain and evict-first variants
        w, flags, resid = _run(emu, kind, ei, ej, z, n, lay, iters, np.float64, grid)
        assert flags[1] == iters and flags[0] == 1                 # FLAG_IT, FLAG_DONE (maxit reached)
        assert _rel(w, wo) < 1e-12
        assert abs(resid[0] - io.rel_primal) < 1e-9 * io.rel_primal
        assert abs(resid[1] - io.rel_dual) < 1e-9 * io.rel_dual


@pytest.mark.parametrize("n,k,window,with_key,grid", [(600, 7, 256, True, 2), (300, 6, 0, False, 1), (257, 5, -1, True, 3),
                                                      (700, 8, 256, True, 5)])
def test_blocked_kernel_source_on_cpu(emu, n, k, window, with_key, grid):
    assert emu.emu_knn_block_rows() == 256
    iters = 5
    pts, ei, ej, z = _case(n, k, n)
    order = orc.knn_row_order(ei, ej, n, orc.morton_keys(pts.T) if with_key else None, window)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block=256)
    assert len(lay["intra_eid"]) > 0 and (len(lay["slot_eid"]) > 0 or n <= 256)
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=iters, tol=0.0, threshold=False)
    w, flags, resid = _run(emu, 1, ei, ej, z, n, lay, iters, np.float64, grid)
    assert flags[1] == iters and flags[0] == 1
    assert _rel(w, wo) < 1e-12
    assert abs(resid[0] - io.rel_primal) < 1e-9 * io.rel_primal
    assert abs(resid[1] - io.rel_dual) < 1e-9 * io.rel_dual
    # float32 instantiation, and a warm start
    w32, _, _ = _run(emu, 1, ei, ej, z, n, lay, iters, np.float32, grid)
    assert _rel(w32, wo) < 1e-5
    init = np.random.default_rng(3).uniform(0, 0.2, size=len(ei))
    wo, io = orc.solve_knn(ei, ej, z, n, edge_init=init, maxit=3, tol=0.0, threshold=False)
    w, flags, resid = _run(emu, 1, ei, ej, z, n, lay, 3, np.float64, grid, w0=init)
    assert _rel(w, wo) < 1e-12 and abs(resid[0] - io.rel_primal) < 1e-9 * io.rel_primal


def test_blocked_kernel_stop_rule(emu):
    """tol > 0: the last CTA's decision stops the solve at the oracle's iteration; later launches are no-ops."""
    n = 300
    pts, ei, ej, z = _case(n, 6, 21)
    order = orc.knn_row_order(ei, ej, n, orc.morton_keys(pts.T), 256)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block=256)
    tol = 2e-2
    wo, io = orc.solve_knn(ei, ej, z, n, maxit=60, tol=tol, threshold=False)
    assert io.converged and 2 < io.n_iter < 50
    w, flags, resid = _run(emu, 1, ei, ej, z, n, lay, io.n_iter + 4, np.float64, 2, tol=tol)
    assert flags[1] == io.n_iter and flags[2] == 1 and flags[0] == 1
    assert _rel(w, wo) < 1e-12


@pytest.mark.parametrize("grid", [1, 4])
def test_blocked_build_kernels_on_cpu(emu, grid):
    """The device-side construction of the intra lists (key -> [radix sort] -> count -> [scan] -> fill),
    the intra part of the initial degrees and the write-back, against oracle.knn_blocked_layout."""
    n = 700
    pts, ei, ej, z = _case(n, 8, 5)
    order = orc.knn_row_order(ei, ej, n, orc.morton_keys(pts.T), 256)
    lay = orc.knn_blocked_layout(ei, ej, n, order, block=256)
    npad = -(-n // 128) * 128
    nblocks = -(-npad // 256)
    e_i, e_j = ei.astype(np.int32), ej.astype(np.int32)
    lab = lay["lab"].astype(np.int32)
    nnz = len(ei)
    key = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=np.int32)
    count = np.zeros(1, dtype=np.uint64)
    emu.emu_intra_key(_p(e_i), _p(e_j), ctypes.c_longlong(nnz), _p(lab), ctypes.c_int(nblocks), _p(key), _p(val),
                      _p(count), ctypes.c_int(grid))
    n_intra = int(count[0])
    assert n_intra == len(lay["intra_eid"])
    assert np.array_equal(val, np.arange(nnz)) and np.all((key == nblocks) == ~np.isin(np.arange(nnz), lay["intra_eid"]))
    o = np.argsort(key, kind="stable")                       # cub::DeviceRadixSort::SortPairs (stable)
    key_s, val_s = key[o].copy(), val[o].copy()
    cnt = np.zeros(nblocks + 1, dtype=np.int32)
    emu.emu_intra_count(_p(key_s), ctypes.c_longlong(n_intra), _p(cnt), ctypes.c_int(grid))
    ptr = np.concatenate([[0], np.cumsum(cnt)[:-1]]).astype(np.int32)   # cub::DeviceScan::ExclusiveSum
    want = np.full(nblocks + 1, lay["intra_ptr"][-1], dtype=np.int32)
    want[:len(lay["intra_ptr"])] = lay["intra_ptr"]
    assert np.array_equal(ptr, want)
    rng = np.random.default_rng(0)
    w0 = rng.uniform(size=nnz)
    ids = np.zeros((n_intra, 2), dtype=np.uint8)
    z_i, w_i = np.empty(n_intra), np.empty(n_intra)
    eid_i = np.empty(n_intra, dtype=np.int32)
    emu.emu_intra_fill(_p(e_i), _p(e_j), _p(z), _p(w0), _p(lab), _p(val_s), ctypes.c_longlong(n_intra), _p(ids), _p(z_i),
                       _p(w_i), _p(eid_i), ctypes.c_int(grid))
    assert np.array_equal(eid_i, lay["intra_eid"])
    assert np.array_equal(ids[:, 0], lay["intra_a"]) and np.array_equal(ids[:, 1], lay["intra_b"])
    assert np.array_equal(z_i, z[eid_i]) and np.array_equal(w_i, w0[eid_i])
    # intra part of S w0 and of ||w0||^2
    sw = np.zeros(npad)
    gsum = np.zeros(2)
    emu.emu_intra_init(_p(ptr), _p(ids), _p(w_i), ctypes.c_int(nblocks), _p(sw), _p(gsum), ctypes.c_int(grid))
    ref = np.bincount(lab[ei[eid_i]], w_i, npad) + np.bincount(lab[ej[eid_i]], w_i, npad)
    assert np.allclose(sw, ref, rtol=1e-13, atol=0) and gsum[0] == 0.0
    assert abs(gsum[1] - np.sum(w_i ** 2)) < 1e-12 * np.sum(w_i ** 2)
    # write-back with the threshold
    out = np.full(nnz, -1.0)
    emu.emu_intra_writeback(_p(eid_i), _p(w_i), ctypes.c_longlong(n_intra), ctypes.c_double(0.5), _p(out),
                            ctypes.c_int(grid))
    assert np.array_equal(out[eid_i], np.where(w_i < 0.5, 0.0, w_i))
    assert np.all(out[~np.isin(np.arange(nnz), eid_i)] == -1.0)


@pytest.mark.parametrize("window,with_key", [(0, True), (64, False), (256, True), (1000, True)])
def test_row_order_and_expand_kernels_on_cpu(emu, window, with_key):
    """knn_window_key_kernel + [stable radix sort] gives oracle.knn_row_order; knn_expand_kernel keeps every
    edge's two slots by default and drops the intra-block edges in the blocked layout."""
    n = 700
    pts, ei, ej, _ = _case(n, 8, 12)
    key = orc.morton_keys(pts.T) if with_key else None
    deg = (np.bincount(ei, minlength=n) + np.bincount(ej, minlength=n)).astype(np.int32)
    arr = (np.arange(n) if key is None else np.argsort(key, kind="stable")).astype(np.int32)   # first radix sort
    wkey = np.empty(n, dtype=np.uint64)
    emu.emu_window_key(_p(arr), _p(deg), ctypes.c_int(n), ctypes.c_int(window), _p(wkey), ctypes.c_int(2))
    order = arr[np.argsort(wkey, kind="stable")]                                                # second radix sort
    assert np.array_equal(order, orc.knn_row_order(ei, ej, n, key, window))
    lab = np.empty(n, dtype=np.int32)
    lab[order] = np.arange(n, dtype=np.int32)
    e_i, e_j = ei.astype(np.int32), ej.astype(np.int32)
    npad = -(-n // 128) * 128
    for block_rows in (0, 256):
        k2 = np.empty(2 * len(ei), dtype=np.int32)
        v2 = np.empty(2 * len(ei), dtype=np.int32)
        count = np.zeros(1, dtype=np.uint64)
        emu.emu_expand(_p(e_i), _p(e_j), ctypes.c_longlong(len(ei)), _p(k2), _p(v2), ctypes.c_int(0), ctypes.c_int(npad),
                       ctypes.c_int(n), _p(count), _p(lab), ctypes.c_int(block_rows), ctypes.c_int(3))
        a, b = lab[ei], lab[ej]
        intra = (a // 256 == b // 256) if block_rows else np.zeros(len(ei), dtype=bool)
        assert int(count[0]) == 2 * int((~intra).sum())
        assert np.array_equal(v2[0::2], np.arange(len(ei))) and np.array_equal(v2[1::2], np.arange(len(ei)))
        assert np.array_equal(k2[0::2], np.where(intra, n, a)) and np.array_equal(k2[1::2], np.where(intra, n, b))
