"""Assignment step of the batched callers on the device.

``QuadForm`` wraps ``gl_quadform_*`` (csrc/quadform.cu): K square matrices -- the learned Laplacians --
are compacted once into non-zero entry lists, then ``apply`` evaluates
``Q[s, k] = (x_s - shift_k)^T (A_k + diag_add I) (x_s - shift_k)`` for every sample, optionally the
dots ``D[s, k] = U_k . (x_s - shift_k)`` and the fused ``argmin_k``.  It replaces

* ``np.einsum("ni,kij,nj->nk", x, laplacians, x)`` + ``argmin`` of the k-Graphs assignment step
  (reference ``kgraphs.py:87-88``, ``:112-113``), and
* the eigenvector projection + weighted square sum of the GLMM E-step (reference
  ``glmm.py:253-274``) through ``sum_{j>=1} (lam_j + r)(v_j . y)^2 = y^T (L + r I) y - (lam_0 + r)(v_0 . y)^2``.

There is no host fallback: without the CUDA library or a device the calls raise.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from .. import _native as nat
from ..smooth_learning import _gl_dtype, _ptr, _stream, _to_device, _torch_dtype


class QuadForm:
    """Per-sample quadratic forms against K (n, n) matrices.  Use as a context manager or call
    :meth:`close`; the handle owns device memory proportional to the non-zero count."""

    def __init__(self, matrices, *, diag_add: float = 0.0, dtype=None):
        dt = _torch_dtype(dtype)
        a = _to_device(matrices, dt)
        if a.ndim == 2:
            a = a[None]
        if a.ndim != 3 or a.shape[1] != a.shape[2]:
            raise ValueError(f"matrices must have shape (K, n, n), got {tuple(a.shape)}")
        self.dtype = dt
        self.n_graphs, self.n_nodes = int(a.shape[0]), int(a.shape[1])
        self._h = ctypes.c_void_p()
        nat.check(nat.lib().gl_quadform_create(ctypes.byref(self._h), _ptr(a), self.n_graphs, self.n_nodes,
                                               _gl_dtype(dt), float(diag_add), _stream()))
        self.nnz = int(nat.lib().gl_quadform_nnz(self._h))

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            nat.lib().gl_quadform_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def apply(self, x, shift=None, vectors=None, *, labels: bool = False):
        """``x`` (m, n) -> ``Q`` (m, K) device tensor; with ``vectors`` (K, n) also ``D`` (m, K); with
        ``labels=True`` also the int32 argmin over k.  Returns ``Q`` or a tuple in that order."""
        if not self._h.value:
            raise RuntimeError("QuadForm is closed")
        xd = _to_device(x, self.dtype)
        if xd.ndim != 2 or xd.shape[1] != self.n_nodes:
            raise ValueError(f"x must have shape (n_samples, {self.n_nodes}), got {tuple(xd.shape)}")
        m = int(xd.shape[0])
        sd: Optional[torch.Tensor] = None
        if shift is not None:
            sd = _to_device(shift, self.dtype)
            if tuple(sd.shape) != (self.n_graphs, self.n_nodes):
                raise ValueError(f"shift must have shape ({self.n_graphs}, {self.n_nodes}), got {tuple(sd.shape)}")
        ud: Optional[torch.Tensor] = None
        if vectors is not None:
            ud = _to_device(vectors, self.dtype)
            if tuple(ud.shape) != (self.n_graphs, self.n_nodes):
                raise ValueError(f"vectors must have shape ({self.n_graphs}, {self.n_nodes}), got {tuple(ud.shape)}")
        q = torch.empty((m, self.n_graphs), dtype=self.dtype, device=xd.device)
        d = torch.empty_like(q) if ud is not None else None
        lab = torch.empty((m,), dtype=torch.int32, device=xd.device) if labels else None
        nat.check(nat.lib().gl_quadform_apply(self._h, _ptr(xd), m, int(xd.stride(0)) if m else self.n_nodes,
                                              _ptr(sd), _ptr(ud), _ptr(q), _ptr(d), _ptr(lab), _stream()))
        out = (q,) + ((d,) if d is not None else ()) + ((lab,) if lab is not None else ())
        return out[0] if len(out) == 1 else out


def smoothness(x, laplacians, means=None, *, dtype=None, labels: bool = False):
    """x_n^T L_k x_n (or the mean-centred form) for every sample and cluster on the device."""
    with QuadForm(laplacians, dtype=dtype) as qf:
        return qf.apply(x, shift=means, labels=labels)
