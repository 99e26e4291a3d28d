"""Learning graphs from smooth signals -- B200-native drop-in for the reference module
``graph_learn/smooth_learning.py`` (LTS4/graph-learning).

Same public names and call semantics as the reference:

* :func:`sum_squareform`               (reference :13-44)
* :func:`gsp_learn_graph_log_degrees`  (reference :47-174)
* :func:`get_theta`                    (reference :177-235)
* :class:`LogModel`                    (reference :238-300)

All arithmetic runs in hand-written sm_100a kernels behind the C-ABI of ``include/gl_b200.h``
(``libgl_b200.so``); torch is used only to own device memory and streams.  Inputs may be NumPy
arrays (results come back as NumPy, like the reference) or CUDA torch tensors (results stay on
the device).  Extra keyword-only arguments (``dtype``, ``return_info``) default to the reference
behaviour.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from itertools import combinations_with_replacement
from typing import Optional

import numpy as np
import torch
from sklearn.base import BaseEstimator

from . import _native as nat

__all__ = [
    "sum_squareform",
    "gsp_learn_graph_log_degrees",
    "get_theta",
    "LogModel",
    "Solver",
    "pdist_sq",
    "pdist_sq_tiles",
    "SolveInfo",
    "release_solver_cache",
]


# --------------------------------------------------------------------------------------------
# device plumbing
# --------------------------------------------------------------------------------------------
def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError(
            "graph_learning_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback"
        )
    return torch.device("cuda", torch.cuda.current_device())


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _torch_dtype(dtype) -> torch.dtype:
    if dtype in (None, "float64", "f64", np.float64, torch.float64, float):
        return torch.float64
    if dtype in ("float32", "f32", np.float32, torch.float32):
        return torch.float32
    raise ValueError(f"dtype must be float32 or float64, got {dtype!r}")


def _gl_dtype(t: torch.dtype) -> int:
    return nat.GL_F32 if t == torch.float32 else nat.GL_F64


def _to_device(a, dtype: torch.dtype) -> torch.Tensor:
    """NumPy / torch -> contiguous CUDA tensor of ``dtype`` (copy only when needed)."""
    if isinstance(a, torch.Tensor):
        return a.to(device=_device(), dtype=dtype, non_blocking=True).contiguous()
    arr = np.ascontiguousarray(a)
    t = torch.from_numpy(arr)
    if arr.dtype != (np.float32 if dtype == torch.float32 else np.float64):
        t = t.to(dtype)
    return t.to(_device(), non_blocking=False).contiguous()


def _ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _n_from_edges(n_edges: int) -> int:
    # smooth_learning.py:133
    n = int(np.round((1 + np.sqrt(1 + 8 * n_edges)) / 2))
    if n * (n - 1) // 2 != n_edges:
        raise ValueError(f"{n_edges} is not a triangular number of edges")
    return n


class SolveInfo:
    """Iteration count and last relative residuals of a solve (the reference returns neither)."""

    def __init__(self, n_iter: int, rel_primal: float, rel_dual: float, converged: bool):
        self.n_iter, self.rel_primal, self.rel_dual, self.converged = n_iter, rel_primal, rel_dual, converged

    def __repr__(self):
        return (f"SolveInfo(n_iter={self.n_iter}, rel_primal={self.rel_primal:.3e}, "
                f"rel_dual={self.rel_dual:.3e}, converged={self.converged})")


# --------------------------------------------------------------------------------------------
# solver workspace (handle API)
# --------------------------------------------------------------------------------------------
class Solver:
    """Resident state of the primal-dual iteration for K graphs on n nodes (``gl_solver``).

    ``tile_range`` restricts the workspace to a shard of the triangle's 128x128 tiles; the
    caller then sums :meth:`exchange` over shards between :meth:`edge_pass` and
    :meth:`node_tail` (see ``graph_learning_b200/sharded.py``).
    """

    def __init__(self, n: int, n_graphs: int = 1, dtype=torch.float64, tile_range=None):
        self.lib = nat.lib()
        self.n, self.K, self.dtype = int(n), int(n_graphs), _torch_dtype(dtype)
        self.device = _device()
        total = self.lib.gl_num_tiles(self.n)
        self.tile_begin, self.tile_end = (0, total) if tile_range is None else map(int, tile_range)
        h = ctypes.c_void_p()
        nat.check(self.lib.gl_solver_create(ctypes.byref(h), self.n, self.K, _gl_dtype(self.dtype),
                                            self.tile_begin, self.tile_end))
        self._h = h
        self.n_edges = self.n * (self.n - 1) // 2

    def close(self):
        if getattr(self, "_h", None):
            self.lib.gl_solver_destroy(self._h)
            self._h = None

    __del__ = close

    # -- optional phase tracing of the persistent kernel (profiling aid, K == 1) -----------------------
    def trace_enable(self) -> None:
        nat.check(self.lib.gl_solver_trace_enable(self._h))

    def trace_read(self, n_iters: int) -> np.ndarray:
        """(n_iters, 8) int64 %globaltimer stamps (ns) of CTA 0, see ``gl_solver_trace_read``."""
        n_iters = min(int(n_iters), 4096)
        out = np.zeros((n_iters, 8), dtype=np.int64)
        nat.check(self.lib.gl_solver_trace_read(self._h, out.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), n_iters))
        return out

    def load(self, z_packed: torch.Tensor, z_scale=None, w_packed: Optional[torch.Tensor] = None):
        assert z_packed.is_cuda and z_packed.dtype == self.dtype and z_packed.is_contiguous()
        assert z_packed.numel() == self.K * self.n_edges
        sc = None
        if z_scale is not None:
            sc = (ctypes.c_double * self.K)(*[float(v) for v in np.broadcast_to(np.asarray(z_scale, dtype=np.float64).ravel(), (self.K,))])
        nat.check(self.lib.gl_solver_load(self._h, _ptr(z_packed), sc, _ptr(w_packed), _stream()))

    def load_rows(self, z_rows: torch.Tensor, z_scale=None, w_rows: Optional[torch.Tensor] = None):
        """:meth:`load` for a tile shard from the shard's slice of the packed vector only
        (``gl_shard_rows`` gives the slice; single graph)."""
        assert z_rows.is_cuda and z_rows.dtype == self.dtype and z_rows.is_contiguous()
        sc = None
        if z_scale is not None:
            sc = (ctypes.c_double * 1)(float(z_scale))
        nat.check(self.lib.gl_solver_load_rows(self._h, _ptr(z_rows), sc, _ptr(w_rows), _stream()))

    def shard_rows(self):
        """(row_begin, row_end, edge_begin, edge_end) of this workspace's tile range."""
        r0, r1 = ctypes.c_int(), ctypes.c_int()
        e0, e1 = ctypes.c_int64(), ctypes.c_int64()
        nat.check(self.lib.gl_shard_rows(self.n, self.tile_begin, self.tile_end, ctypes.byref(r0), ctypes.byref(r1),
                                         ctypes.byref(e0), ctypes.byref(e1)))
        return r0.value, r1.value, e0.value, e1.value

    def z_view(self, k: int = 0) -> torch.Tensor:
        """Zero-copy view of graph k's resident tiled distances (this shard's tiles)."""
        return _wrap_device_memory(self.z_tiles_ptr(k), (self.tile_end - self.tile_begin) * nat.TILE * nat.TILE,
                                   self.dtype, self.device, owner=self)

    def w_view(self, k: int = 0) -> torch.Tensor:
        return _wrap_device_memory(self.w_tiles_ptr(k), (self.tile_end - self.tile_begin) * nat.TILE * nat.TILE,
                                   self.dtype, self.device, owner=self)

    def load_w_packed(self, w_packed: Optional[torch.Tensor]):
        """Initial weights from a full packed (K, E) array into the resident tiles (zeros if None)."""
        if w_packed is None:
            for k in range(self.K):
                self.w_view(k).zero_()
            return
        assert w_packed.is_cuda and w_packed.dtype == self.dtype and w_packed.is_contiguous()
        nat.check(self.lib.gl_pack_to_tiles(_ptr(w_packed), ctypes.c_void_p(self.w_tiles_ptr(0)), self.n, self.K,
                                            _gl_dtype(self.dtype), self.tile_begin, self.tile_end, _stream()))

    def z_tiles_ptr(self, k: int = 0) -> int:
        return self.lib.gl_solver_z_tiles(self._h, k)

    def w_tiles_ptr(self, k: int = 0) -> int:
        return self.lib.gl_solver_w_tiles(self._h, k)

    def scale_z(self, z_scale):
        sc = (ctypes.c_double * self.K)(*[float(v) for v in np.broadcast_to(np.asarray(z_scale, dtype=np.float64).ravel(), (self.K,))])
        nat.check(self.lib.gl_solver_scale_z(self._h, sc, _stream()))

    def begin(self, alpha=1.0, beta=1.0, step_size=0.5, maxit=1000, tol=1e-5, finish=True):
        nat.check(self.lib.gl_solver_begin(self._h, float(alpha), float(beta), float(step_size), int(maxit),
                                           float(tol), _stream()))
        if finish:
            self.begin_finish()

    def begin_finish(self):
        nat.check(self.lib.gl_solver_begin_finish(self._h, _stream()))

    def edge_pass(self):
        nat.check(self.lib.gl_solver_edge_pass(self._h, _stream()))

    def node_tail(self):
        nat.check(self.lib.gl_solver_node_tail(self._h, _stream()))

    def exchange(self) -> torch.Tensor:
        """float64 view of the exchange buffer (degree partials + residual Grams)."""
        n = ctypes.c_int64()
        p = self.lib.gl_solver_exchange_ptr(self._h, ctypes.byref(n))
        return _wrap_device_memory(p, n.value, torch.float64, self.device, owner=self)

    def run(self, poll_every: int = 0):
        nat.check(self.lib.gl_solver_run(self._h, int(poll_every), _stream()))

    # -- peer (NVLink) exchange ---------------------------------------------------------------------
    def comm_handle(self) -> bytes:
        """64-byte CUDA IPC handle of this rank's comm block (to be all-gathered by the caller)."""
        buf = ctypes.create_string_buffer(64)
        nat.check(self.lib.gl_solver_comm_handle(self._h, buf))
        return buf.raw

    def comm_connect(self, rank: int, world: int, handles: bytes):
        assert len(handles) == 64 * world
        nat.check(self.lib.gl_solver_comm_connect(self._h, int(rank), int(world), handles))

    def comm_publish(self):
        nat.check(self.lib.gl_solver_comm_publish(self._h, _stream()))

    def comm_ptr(self) -> int:
        return self.lib.gl_solver_comm_ptr(self._h, None)

    def comm_connect_ptrs(self, rank: int, world: int, ptrs):
        arr = (ctypes.c_void_p * world)(*[ctypes.c_void_p(int(p)) for p in ptrs])
        nat.check(self.lib.gl_solver_comm_connect_ptrs(self._h, int(rank), int(world), arr))

    def status(self) -> SolveInfo:
        it, conv = ctypes.c_int(), ctypes.c_int()
        rp, rd = ctypes.c_double(), ctypes.c_double()
        nat.check(self.lib.gl_solver_status(self._h, ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd),
                                            ctypes.byref(conv), _stream()))
        return SolveInfo(it.value, rp.value, rd.value, bool(conv.value))

    def store(self, edge_tol: float = 1e-3, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Packed (K, E) weights with the edge_tol threshold (smooth_learning.py:169).  For a
        sharded workspace only the shard's edges are written (``out`` should start zeroed)."""
        if out is None:
            out = torch.zeros((self.K, self.n_edges), dtype=self.dtype, device=self.device)
        nat.check(self.lib.gl_solver_store(self._h, _ptr(out), float(edge_tol), _stream()))
        return out

    def square(self, edge_tol: float = 1e-3) -> torch.Tensor:
        """Dense symmetric (n, n) weights of graph 0 straight from the tiled state."""
        assert self.tile_begin == 0 and self.tile_end == self.lib.gl_num_tiles(self.n)
        out = torch.empty((self.n, self.n), dtype=self.dtype, device=self.device)
        nat.check(self.lib.gl_tiles_to_square(ctypes.c_void_p(self.w_tiles_ptr(0)), _ptr(out), self.n,
                                              _gl_dtype(self.dtype), float(edge_tol), _stream()))
        return out

    def node_state(self) -> torch.Tensor:
        d = torch.empty((self.K, self.n), dtype=torch.float64, device=self.device)
        nat.check(self.lib.gl_solver_node_state(self._h, _ptr(d), _stream()))
        return d


def _wrap_device_memory(ptr: int, numel: int, dtype: torch.dtype, device: torch.device, owner=None):
    """Zero-copy torch view of library-owned device memory (via __cuda_array_interface__)."""

    class _Holder:
        pass

    h = _Holder()
    typestr = {torch.float64: "<f8", torch.float32: "<f4", torch.int32: "<i4"}[dtype]
    h.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (int(ptr), False), "version": 2}
    h._owner = owner
    return torch.as_tensor(h, device=device)


# --------------------------------------------------------------------------------------------
# the edge <-> degree operator pair (reference sum_squareform, :13-44)
# --------------------------------------------------------------------------------------------
class _SumOperator:
    """Matrix-free S (n x E) or S^T (E x n).  Supports ``@`` on (E,), (E,K), (n,), (n,K) NumPy
    arrays or CUDA tensors, ``.shape`` and ``.T`` -- what the reference's callers use
    (smooth_learning.py:136,144-154; temporal/graph_factor.py:101; dictionary/log_model.py:28)."""

    def __init__(self, n: int, transposed: bool):
        self.n, self.transposed = int(n), transposed
        e = self.n * (self.n - 1) // 2
        self.shape = (e, self.n) if transposed else (self.n, e)

    @property
    def T(self):
        return _SumOperator(self.n, not self.transposed)

    def __matmul__(self, v):
        is_np = not isinstance(v, torch.Tensor)
        dt = torch.float32 if (getattr(v, "dtype", None) in (np.float32, torch.float32)) else torch.float64
        x = _to_device(v, dt)
        if x.shape[0] != self.shape[1]:
            raise ValueError(f"dimension mismatch: operator {self.shape} @ operand {tuple(x.shape)}")
        vec = x.ndim == 1
        xk = (x[None, :] if vec else x.t()).contiguous()  # (K, len)
        k = xk.shape[0]
        lib = nat.lib()
        out = torch.empty((k, self.shape[0]), dtype=dt, device=x.device)
        fn = lib.gl_sum_op_adj if self.transposed else lib.gl_sum_op
        nat.check(fn(_ptr(xk), _ptr(out), self.n, k, _gl_dtype(dt), _stream()))
        res = out[0] if vec else out.t().contiguous()
        return res.cpu().numpy() if is_np else res


def sum_squareform(n: int):
    """Operator S with ``S @ w`` = weighted degrees of the packed edge vector w, and its
    transpose (``St @ d`` = d_i + d_j per edge).  Reference :13-44 materialises both as CSR
    matrices with O(n^2) Python list work; here they are matrix-free CUDA kernels."""
    return _SumOperator(n, False), _SumOperator(n, True)


# --------------------------------------------------------------------------------------------
# pairwise squared distances (reference :278) and theta (reference :177-235)
# --------------------------------------------------------------------------------------------
# tensor-core Gram form (two-slice fp16 since round 2): used by "auto" for float32 when there are enough
# samples to pay for the split pre-pass; the exact direct-difference kernel stays the float64 path and
# the parity anchor.
GRAM_MIN_SAMPLES = 256
_gram_ws = {}


def _gram_workspace(m: int, n: int, terms: int, device) -> torch.Tensor:
    """Scratch for the split operand planes; one cached buffer per device, grown on demand."""
    need = int(nat.lib().gl_gram_workspace_bytes(m, n, terms))
    if need < 0:
        raise ValueError(f"bad Gram problem size (m={m}, n={n}, terms={terms})")
    key = (device.type, device.index)
    buf = _gram_ws.get(key)
    if buf is None or buf.numel() < need:
        _gram_ws.pop(key, None)
        buf = torch.empty((need,), dtype=torch.uint8, device=device)
        _gram_ws[key] = buf
    return buf


def release_gram_workspace() -> None:
    """Drop the cached Gram workspace (6 GB at n=20,000, m=50,000)."""
    _gram_ws.clear()


def _resolve_method(method: str, dt: torch.dtype, m: int) -> str:
    if method not in ("auto", "direct", "gram", "gram2", "gram16"):
        raise ValueError(f"method must be auto, direct, gram, gram16 or gram2, got {method!r}")
    if method == "auto":
        return "gram16" if (dt == torch.float32 and m >= GRAM_MIN_SAMPLES) else "direct"
    if method != "direct" and dt != torch.float32:
        raise ValueError("the tensor-core Gram form is float32 only; use method='direct' for float64")
    return method


def _pdist_into(xd, rs, cs, out, layout, tile_begin, tile_end, method, chunk=0, variant=0) -> None:
    """Launch the chosen distance kernel writing ``out`` (packed (E,) or a tile range)."""
    m, n = xd.shape
    if method == "direct":
        nat.check(nat.lib().gl_pdist_sq(_ptr(xd), m, n, xd.stride(0), _ptr(rs), _ptr(cs), _ptr(out),
                                        layout, tile_begin, tile_end, _gl_dtype(xd.dtype), _stream()))
        return
    terms = 2 if method in ("gram2", "gram16") else 3
    if method == "gram16" and variant not in (7, 8):
        variant = 7      # fp16 two-slice form (normalised planes, four products); 8 = the same on CTA pairs
    ws = _gram_workspace(m, n, terms, xd.device)
    nat.check(nat.lib().gl_pdist_sq_gram(_ptr(xd), m, n, xd.stride(0), _ptr(rs), _ptr(cs), _ptr(out),
                                         layout, tile_begin, tile_end, terms, int(chunk), int(variant),
                                         _ptr(ws), ws.numel(), _stream()))


def pdist_sq(x, row_scale=None, col_shift=None, dtype=None, method="auto", chunk=0, variant=0) -> torch.Tensor:
    """Packed squared distances between the COLUMNS of x (m, n): ``pdist(x.T) ** 2``.

    With ``row_scale`` (m,) / ``col_shift`` (n,) the signals are first mapped to
    ``row_scale[:, None] * (x - col_shift)`` -- the GLMM M-step form (glmm.py:61-68).
    ``method``: "direct" = exact differences on the FMA pipe (float32/float64); "gram16" = tcgen05
    Gram GEMM on two fp16 slices of per-node normalised signals (float32 accuracy: weights within
    3e-6 of the float64 solve; four products per k-step; the float32 default); "gram" = the same on
    three bf16 slices (six products, 1.35x slower, same accuracy); "gram2" = 2-term bf16 splits (faster,
    ~16 bits); "auto" = gram for float32 with >= 256 samples, direct otherwise.
    Returns a CUDA tensor of shape (E,).
    """
    dt = _torch_dtype(dtype if dtype is not None else (x.dtype if getattr(x, "dtype", None) in (np.float32, torch.float32) else None))
    xd = _to_device(x, dt)
    if xd.ndim != 2:
        raise ValueError("signals must be a (n_samples, n_nodes) matrix")
    m, n = xd.shape
    rs = None if row_scale is None else _to_device(row_scale, dt)
    cs = None if col_shift is None else _to_device(col_shift, dt)
    out = torch.empty((n * (n - 1) // 2,), dtype=dt, device=xd.device)
    _pdist_into(xd, rs, cs, out, nat.LAYOUT_PACKED, 0, 0, _resolve_method(method, dt, m), chunk, variant)
    return out


def pdist_sq_tiles(x, tile_begin: int, tile_end: int, out=None, dtype=None, method="auto", chunk=0,
                   variant=0) -> torch.Tensor:
    """Same quantity written in the solver's resident tiled layout for tiles
    [tile_begin, tile_end) -- what a GPU shard builds for itself."""
    dt = _torch_dtype(dtype if dtype is not None else (x.dtype if getattr(x, "dtype", None) in (np.float32, torch.float32) else None))
    xd = _to_device(x, dt)
    m, n = xd.shape
    if out is None:
        out = torch.empty(((tile_end - tile_begin) * nat.TILE * nat.TILE,), dtype=dt, device=xd.device)
    _pdist_into(xd, None, None, out, nat.LAYOUT_TILED, tile_begin, tile_end, _resolve_method(method, dt, m),
                chunk, variant)
    return out


def _theta_dense(mat: torch.Tensor, avg_degree: int) -> float:
    th = ctypes.c_double()
    rows, cols = mat.shape
    nat.check(nat.lib().gl_theta_dense(_ptr(mat), rows, cols, mat.stride(0), int(avg_degree),
                                       _gl_dtype(mat.dtype), ctypes.byref(th), _stream()))
    return th.value


def _theta_packed(z_packed: torch.Tensor, n: int, avg_degree: int) -> float:
    th = ctypes.c_double()
    nat.check(nat.lib().gl_theta(_ptr(z_packed), nat.LAYOUT_PACKED, n, int(avg_degree),
                                 _gl_dtype(z_packed.dtype), ctypes.byref(th), _stream()))
    return th.value


def _theta_tiles(z_tiles: torch.Tensor, n: int, avg_degree: int, tile_range=None, group=None) -> float:
    """theta from the resident tiled distances.  With a tile shard per rank (``tile_range``, a process
    ``group``) every rank extracts the avg_degree + 1 smallest entries of each row among its own tiles
    (``gl_theta_partial``); the candidate lists are all-gathered (n x world x (k + 1) values) and merged
    by the same row statistic -- identical on every rank, equal to the single-GPU value."""
    lib = nat.lib()
    if tile_range is None:
        th = ctypes.c_double()
        nat.check(lib.gl_theta(_ptr(z_tiles), nat.LAYOUT_TILED, n, int(avg_degree), _gl_dtype(z_tiles.dtype),
                               ctypes.byref(th), _stream()))
        return th.value
    import torch.distributed as dist
    k1 = int(avg_degree) + 1
    cand = torch.empty((n, k1), dtype=z_tiles.dtype, device=z_tiles.device)
    nat.check(lib.gl_theta_partial(_ptr(z_tiles), n, int(tile_range[0]), int(tile_range[1]), int(avg_degree),
                                   _gl_dtype(z_tiles.dtype), _ptr(cand), _stream()))
    world = dist.get_world_size(group)
    allc = torch.empty((world, n, k1), dtype=cand.dtype, device=cand.device)
    dist.all_gather_into_tensor(allc, cand, group=group)
    merged = allc.permute(1, 0, 2).reshape(n, world * k1).contiguous()
    return _theta_dense(merged, avg_degree)


def _upload_signals_distributed(x, dt: torch.dtype, group) -> torch.Tensor:
    """Host signals (m, n), identical on every rank -> the full matrix on every rank's device, with each
    rank moving only ITS block of rows over PCIe and the blocks all-gathered over NVLink (8 ranks: 1/8 of
    the 4 GB of config 3 per host link instead of all of it on each, then 900 GB/s links for the rest)."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
    if xt.ndim != 2:
        raise ValueError("signals must be a (n_samples, n_nodes) matrix")
    m, n = xt.shape
    if world == 1 or m < world:
        return _to_device(x, dt)
    per = -(-m // world)
    full = torch.empty((per * world, n), dtype=dt, device=_device())
    lo, hi = min(rank * per, m), min((rank + 1) * per, m)
    mine = full[rank * per:(rank + 1) * per]
    if hi > lo:
        mine[:hi - lo].copy_(xt[lo:hi], non_blocking=True)      # dtype conversion (if any) happens on the device side of the copy
    if hi - lo < per:
        mine[hi - lo:].zero_()
    dist.all_gather_into_tensor(full.view(-1), mine.reshape(-1), group=group)
    return full[:m]


def get_theta(sq_pdists, avg_degree, blocks=None):
    """Theta giving the requested average degree (Kalofolias & Perraudin 2019).

    ``sq_pdists`` is the SQUARE (n, n) matrix of squared distances, as in the reference
    (:177-235); rows are ranked with their zero diagonal included (:202).  With ``blocks`` the
    result is the packed (E,) vector of per-block-pair thetas (:220-235).
    """
    dt = torch.float32 if getattr(sq_pdists, "dtype", None) in (np.float32, torch.float32) else torch.float64
    is_np = not isinstance(sq_pdists, torch.Tensor)
    sq = _to_device(sq_pdists, dt)
    if blocks is None:
        return _theta_dense(sq, avg_degree)
    blocks = np.asarray(blocks.cpu() if isinstance(blocks, torch.Tensor) else blocks)
    n = blocks.size
    thetas = torch.zeros((n, n), dtype=dt, device=sq.device)
    for pair in combinations_with_replacement(np.unique(blocks).tolist(), 2):
        idx1 = torch.as_tensor(np.where(blocks == pair[0])[0], device=sq.device)
        idx2 = torch.as_tensor(np.where(blocks == pair[1])[0], device=sq.device)
        block_degree = avg_degree[pair] if isinstance(avg_degree, dict) else avg_degree
        sub = sq[idx1[:, None], idx2[None, :]].contiguous()
        theta = _theta_dense(sub, block_degree)
        thetas[idx1[:, None], idx2[None, :]] = theta
        thetas[idx2[:, None], idx1[None, :]] = theta
    iu = torch.triu_indices(n, n, 1, device=sq.device)
    packed = thetas[iu[0], iu[1]]
    return packed.cpu().numpy() if is_np else packed


# --------------------------------------------------------------------------------------------
# the solver (reference :47-174)
# --------------------------------------------------------------------------------------------
_solver_cache: dict = {}


def _cached_solver(n: int, n_graphs: int, dt: torch.dtype) -> "Solver":
    """The workspace of the last (n, K, dtype, device) solve is kept for the next call: the reference
    rebuilds its operator on every call (:133), callers in a loop (GLMM / k-Graphs outer iterations, a
    benchmark) should not pay two n^2-sized device allocations and frees per solve.  One workspace at a
    time; :func:`release_solver_cache` frees it."""
    key = (int(n), int(n_graphs), dt, torch.cuda.current_device())
    ws = _solver_cache.get(key)
    if ws is None or getattr(ws, "_h", None) is None:
        release_solver_cache()
        ws = Solver(n, n_graphs, dt)
        _solver_cache[key] = ws
    return ws


def release_solver_cache() -> None:
    """Free the cached single-GPU solver workspace (w, z and the node vectors of the last problem size)."""
    for key in list(_solver_cache):
        _solver_cache.pop(key).close()


def gsp_learn_graph_log_degrees(
    distances,
    alpha: float,
    beta: float,
    edge_init=None,
    maxit: int = 1000,
    tol: float = 1e-5,
    step_size: float = 0.5,
    edge_tol: float = 1e-3,
    *,
    dtype=None,
    return_info: bool = False,
    out=None,
    z_scale=None,
    distributed=False,
    result_on: str = "all",
):
    r"""Weighted adjacency from squared pairwise distances under the smoothness prior
    (Kalofolias 2016): minimise :math:`\sum_{ij} w_{ij} z_{ij} - \alpha \sum_i \log(\sum_j w_{ij})
    + \beta \|w\|^2` over :math:`w \ge 0` with the forward-backward-forward primal-dual method.

    Call-compatible with the reference (:47-56): ``distances`` is (E,) -> returns the dense
    symmetric (n, n) matrix, or (K, E) -> returns (K, E) (K graphs sharing ONE stopping decision
    with spectral-norm residuals, :128-129, :158-159); more dimensions raise ``ValueError``
    (:130-131).  ``edge_init`` uses the reference's internal layout: (E,) or (E, K) (:140-141).
    Entries below ``edge_tol`` are zeroed (:169).

    ``dtype`` (extra): arithmetic type of the edge arrays, default float64 (the reference's);
    float32 halves the memory traffic.  ``return_info`` (extra): also return a :class:`SolveInfo`.
    ``out`` (extra): preallocated result (e.g. a pinned CPU tensor) to copy the result into.
    A CPU torch tensor as input is treated like a NumPy array (host in, host out) and is copied
    asynchronously when pinned.
    ``z_scale`` (extra): scalar or K values multiplied into the distances while they are loaded (the
    ``theta`` of :289, ``glmm.py:79``, ``kgraphs.py:102``) -- saves the caller a full pass over Z.
    ``distributed`` (extra): ``True`` (default process group) or a ``torch.distributed`` group: one
    process per GPU, every rank calls with the SAME arguments; a single graph's triangle tiles are
    sharded over the ranks (each rank uploads only its slice of ``distances``), the result arrives on
    every rank (``result_on="all"``) or on rank 0 only (``"root"``; the other ranks get ``None``).  A
    (K, E) batch is not sharded: every rank solves it (replicas).
    """
    is_np = not isinstance(distances, torch.Tensor)
    is_host_tensor = isinstance(distances, torch.Tensor) and not distances.is_cuda
    n_dim = len(distances.shape)
    if n_dim > 2:
        raise ValueError("Distances must be vector or matrix")
    if n_dim == 0:
        raise ValueError("Distances must be vector or matrix")
    if dtype is None and getattr(distances, "dtype", None) in (np.float32, torch.float32):
        dtype = torch.float32
    dt = _torch_dtype(dtype)
    if distributed is not False and distributed is not None and n_dim == 1:
        from .sharded import solve_distributed
        return solve_distributed(distances, alpha, beta, edge_init, maxit, tol, step_size, edge_tol, dt,
                                 None if distributed is True else distributed, result_on, out, return_info,
                                 z_scale)
    z = _to_device(distances, dt)
    zk = z[None, :] if n_dim == 1 else z
    n_graphs, n_edges = zk.shape
    n = _n_from_edges(n_edges)

    w0 = None
    if edge_init is not None:
        w0 = _to_device(edge_init, dt)
        w0 = (w0[None, :] if n_dim == 1 else w0.t()).contiguous()  # internal (E,K) -> (K,E)
        if tuple(w0.shape) != (n_graphs, n_edges):
            raise ValueError(f"edge_init has shape {tuple(edge_init.shape)}, expected "
                             f"{(n_edges,) if n_dim == 1 else (n_edges, n_graphs)}")

    solver = _cached_solver(n, n_graphs, dt)
    nvtx = torch.cuda.nvtx.range
    try:
        with nvtx("gl.load"):
            solver.load(zk.contiguous(), z_scale, w0)
        with nvtx("gl.solve"):
            solver.begin(alpha, beta, step_size, maxit, tol)
            solver.run()
            info = solver.status()
        with nvtx("gl.store"):
            result = solver.square(edge_tol) if n_dim == 1 else solver.store(edge_tol)
    except BaseException:
        release_solver_cache()
        raise
    res = result
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


# --------------------------------------------------------------------------------------------
# the estimator (reference :238-300)
# --------------------------------------------------------------------------------------------
class LogModel(BaseEstimator):
    """Extract the graph on which data is smooth (Kalofolias 2016 model, Kalofolias & Perraudin
    2019 parametrisation).  Drop-in for the reference ``LogModel`` (:238-300).

    Parameters:
        avg_degree (int, optional): Desired average number of neighbours per node.
        edge_init (array, optional): Prior on graph weights, packed (E,).
        maxit (int): Max optimisation iterations. Defaults to 1000.
        tol (float): Optimisation tolerance. Defaults to 1e-5.
        step_size (float): Optimisation step size. Defaults to 0.5.
        edge_tol (float): Tolerance to keep positive edges. Defaults to 1e-3.
        dtype (extra, keyword-only): "float64" (reference arithmetic, default) or "float32".
        distributed (extra, keyword-only): ``False`` (one GPU), ``True`` (the default
            ``torch.distributed`` process group) or a process group: one process per GPU, every rank
            calls ``fit`` with the same signals; each rank builds the distance tiles of its own shard of
            the triangle, theta comes from per-rank row candidates merged after one
            all-gather, the solve exchanges its degree partials over NVLink peer memory.
        weights_on (extra, keyword-only): "all" (every rank gets ``weights_``) or "root" (rank 0 only;
            ``None`` elsewhere) for distributed fits.

    Attributes:
        theta_ (float): Sparsity parameter rescaling the pairwise distances.
        weights_ (ndarray): Dense symmetric (n, n) edge weights of the learned graph.
        n_iter_ (int), rel_primal_ (float), rel_dual_ (float), converged_ (bool): extras.
    """

    def __init__(
        self,
        avg_degree: int = None,
        *,
        edge_init=None,
        maxit: int = 1000,
        tol: float = 1e-5,
        step_size: float = 0.5,
        edge_tol: float = 1e-3,
        dtype="float64",
        distributed=False,
        weights_on: str = "all",
    ) -> None:
        self.avg_degree = avg_degree
        self.edge_init = edge_init
        self.maxit = maxit
        self.tol = tol
        self.step_size = step_size
        self.edge_tol = edge_tol
        self.dtype = dtype
        self.distributed = distributed
        self.weights_on = weights_on

    def _initialize(self, x):
        """Squared pairwise distances (packed, on the device) and ``theta_`` (:277-284).  Subclasses
        override this (``WindowLogModel``); :meth:`fit` itself takes the resident route below when it
        is not overridden."""
        sq_pdists = pdist_sq(x, dtype=self.dtype)
        if self.avg_degree is None:
            self.theta_ = 1
        else:
            n = x.shape[1]
            self.theta_ = _theta_packed(sq_pdists, n, self.avg_degree)
        return sq_pdists

    def _group(self):
        d = getattr(self, "distributed", False)
        if d is False or d is None:
            return False
        return None if d is True else d

    def _fit_resident(self, x, out=None):
        """X -> Z tiles written straight into the solver's resident buffer (this rank's shard) ->
        theta -> in-place scaling -> solve -> (n, n).  No packed Z, no eager elementwise pass, no
        re-tiling (:277-300)."""
        dt = _torch_dtype(self.dtype)
        group = self._group()
        if group is not False and not (isinstance(x, torch.Tensor) and x.is_cuda):
            xd = _upload_signals_distributed(x, dt, group)
        else:
            xd = _to_device(x, dt)
        if xd.ndim != 2:
            raise ValueError("signals must be a (n_samples, n_nodes) matrix")
        m, n = xd.shape
        sharded = None
        if group is False:
            solver = _cached_solver(n, 1, dt)
            tile_range = None
        else:
            from .sharded import sharded_workspace
            sharded = sharded_workspace(n, dt, group)
            solver = sharded.solver
            tile_range = sharded.tile_range if sharded.world > 1 else None
        nvtx = torch.cuda.nvtx.range      # phase ranges for nsys / ncu --nvtx
        try:
            method = _resolve_method("auto", dt, m)
            with nvtx("gl.pdist_sq_tiles"):
                _pdist_into(xd, None, None, solver.z_view(0), nat.LAYOUT_TILED, solver.tile_begin, solver.tile_end, method)
            if self.avg_degree is None:
                self.theta_ = 1
            else:
                with nvtx("gl.theta"):
                    self.theta_ = _theta_tiles(solver.z_view(0), n, self.avg_degree, tile_range,
                                               None if group is False else group)
                    solver.scale_z(self.theta_)
            w0 = None
            if self.edge_init is not None:
                w0 = _to_device(self.edge_init, dt).reshape(1, -1).contiguous()
                if w0.shape[1] != n * (n - 1) // 2:
                    raise ValueError(f"edge_init has shape {tuple(np.shape(self.edge_init))}, expected {(n * (n - 1) // 2,)}")
            solver.load_w_packed(w0)
            if sharded is None:
                with nvtx("gl.solve"):
                    solver.begin(1, 1, self.step_size, self.maxit, self.tol)
                    solver.run()
                    info = solver.status()
                with nvtx("gl.tiles_to_square"):
                    result = solver.square(self.edge_tol)
            else:
                with nvtx("gl.solve_sharded"):
                    sharded.begin(1, 1, self.step_size, self.maxit, self.tol)
                    sharded.run()
                    info = sharded.status()
                with nvtx("gl.gather_square"):
                    result = sharded.gather_square(self.edge_tol, self.weights_on)
        except BaseException:
            if sharded is None:
                release_solver_cache()
            raise
        return result, info

    def fit(self, x, out=None):
        """x: (n_samples, n_nodes) signals, nodes are COLUMNS (:286-300).  ``out`` (extra): a
        preallocated (n, n) host array / pinned tensor that receives ``weights_``."""
        is_np = not isinstance(x, torch.Tensor)
        if type(self)._initialize is LogModel._initialize:
            weights, info = self._fit_resident(x)
        else:
            sq_pdists = self._initialize(x)
            if not isinstance(sq_pdists, torch.Tensor):
                sq_pdists = _to_device(sq_pdists, _torch_dtype(self.dtype))
            theta = self.theta_
            z_scale = None
            if isinstance(theta, (np.ndarray, torch.Tensor)) and np.size(theta) == (sq_pdists.shape[0] if sq_pdists.ndim == 2 else 1):
                z_scale = np.asarray(theta.cpu() if isinstance(theta, torch.Tensor) else theta, dtype=np.float64).ravel()
            elif isinstance(theta, (np.ndarray, torch.Tensor)):
                sq_pdists = sq_pdists * _to_device(theta, sq_pdists.dtype)   # per-edge theta (block form, :220-235)
            else:
                z_scale = float(theta)
            group = self._group()
            weights, info = gsp_learn_graph_log_degrees(
                sq_pdists,
                1,
                1,
                edge_init=self.edge_init,
                maxit=self.maxit,
                tol=self.tol,
                step_size=self.step_size,
                edge_tol=self.edge_tol,
                dtype=self.dtype,
                return_info=True,
                z_scale=z_scale,
                distributed=False if group is False else (True if group is None else group),
                result_on=self.weights_on,
            )
        if weights is None:
            self.weights_ = None
        elif out is not None:
            out_t = torch.from_numpy(out) if isinstance(out, np.ndarray) else out
            out_t.copy_(weights, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            self.weights_ = out
        else:
            self.weights_ = weights.cpu().numpy() if (is_np and isinstance(weights, torch.Tensor)) else weights
        self.n_iter_, self.rel_primal_, self.rel_dual_ = info.n_iter, info.rel_primal, info.rel_dual
        self.converged_ = info.converged
        return self
