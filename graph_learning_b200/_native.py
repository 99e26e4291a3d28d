"""ctypes binding of libgl_b200.so (the C-ABI declared in include/gl_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import of
:func:`lib`, and every compute call fails with the CUDA error when no device is present.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, c_char_p, c_double, c_int, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgl_b200.so")
CSRC = os.path.join(_HERE, "csrc")

GL_F32, GL_F64 = 0, 1
GL_OK, GL_E_SHAPE, GL_E_ARG, GL_E_CUDA, GL_E_ALLOC, GL_E_STATE = 0, 1, 2, 3, 4, 5
LAYOUT_PACKED, LAYOUT_TILED = 0, 1
TILE = 128

_lib = None

_pd = POINTER(c_double)
_pi = POINTER(c_int)

# name -> (restype, argtypes); must list every symbol of include/gl_b200.h
SIGNATURES = {
    "gl_version": (c_int, []),
    "gl_last_error": (c_char_p, []),
    "gl_launch_count": (c_int64, []),
    "gl_num_edges": (c_int64, [c_int]),
    "gl_nodes_from_edges": (c_int, [c_int64]),
    "gl_num_tiles": (c_int, [c_int]),
    "gl_tiled_elems": (c_int64, [c_int]),
    "gl_shard_tiles": (c_int, [c_int, c_int, c_int, _pi, _pi]),
    "gl_shard_rows": (c_int, [c_int, c_int, c_int, _pi, _pi, POINTER(c_int64), POINTER(c_int64)]),
    "gl_pack_to_tiles": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "gl_tiles_to_packed": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "gl_tiles_to_square": (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_void_p]),
    "gl_sum_op": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "gl_sum_op_adj": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "gl_laplacian_from_packed": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "gl_laplacian_adjoint": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "gl_pdist_sq": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int,
                            c_int, c_int, c_void_p]),
    "gl_gram_workspace_bytes": (c_int64, [c_int64, c_int, c_int]),
    "gl_pdist_sq_gram": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                 c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "gl_theta": (c_int, [c_void_p, c_int, c_int, c_int, c_int, _pd, c_void_p]),
    "gl_theta_noalloc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, _pd, c_void_p]),
    "gl_theta_partial": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "gl_theta_dense": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, _pd, c_void_p]),
    "gl_solver_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int]),
    "gl_solver_destroy": (c_int, [c_void_p]),
    "gl_solver_z_tiles": (c_void_p, [c_void_p, c_int]),
    "gl_solver_w_tiles": (c_void_p, [c_void_p, c_int]),
    "gl_solver_load": (c_int, [c_void_p, c_void_p, _pd, c_void_p, c_void_p]),
    "gl_solver_load_rows": (c_int, [c_void_p, c_void_p, _pd, c_void_p, c_void_p]),
    "gl_solver_scale_z": (c_int, [c_void_p, _pd, c_void_p]),
    "gl_solver_begin": (c_int, [c_void_p, c_double, c_double, c_double, c_int, c_double, c_void_p]),
    "gl_solver_begin_finish": (c_int, [c_void_p, c_void_p]),
    "gl_solver_edge_pass": (c_int, [c_void_p, c_void_p]),
    "gl_solver_node_tail": (c_int, [c_void_p, c_void_p]),
    "gl_solver_comm_handle": (c_int, [c_void_p, c_void_p]),
    "gl_solver_comm_connect": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "gl_solver_comm_first_nodes": (c_int, [c_void_p, _pi]),
    "gl_solver_comm_ptr": (c_void_p, [c_void_p, POINTER(c_int64)]),
    "gl_solver_comm_connect_ptrs": (c_int, [c_void_p, c_int, c_int, POINTER(c_void_p)]),
    "gl_solver_comm_publish": (c_int, [c_void_p, c_void_p]),
    "gl_solver_exchange_ptr": (c_void_p, [c_void_p, POINTER(c_int64)]),
    "gl_solver_run": (c_int, [c_void_p, c_int, c_void_p]),
    "gl_solver_status": (c_int, [c_void_p, _pi, _pd, _pd, _pi, c_void_p]),
    "gl_solver_trace_enable": (c_int, [c_void_p]),
    "gl_solver_trace_read": (c_int, [c_void_p, POINTER(c_int64), c_int]),
    "gl_solver_store": (c_int, [c_void_p, c_void_p, c_double, c_void_p]),
    "gl_solver_node_state": (c_int, [c_void_p, c_void_p, c_void_p]),
    "gl_logmodel_solve": (c_int, [c_void_p, _pd, c_int, c_int, c_int, c_double, c_double, c_void_p, c_int,
                                  c_double, c_double, c_double, c_void_p, _pi, _pd, _pd, c_void_p]),
    "gl_logmodel_solve_host": (c_int, [c_void_p, _pd, c_int, c_int, c_int, c_double, c_double, c_void_p, c_int,
                                       c_double, c_double, c_double, c_void_p, _pi, _pd, _pd]),
    "gl_knn_max_degree": (c_int, [c_void_p, c_void_p, c_int64, c_int, _pi, c_void_p]),
    "gl_knn_create": (c_int, [POINTER(c_void_p), c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                              c_int, c_int64, c_void_p]),
    "gl_knn_create_ordered": (c_int, [POINTER(c_void_p), c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int,
                                      c_int, c_int, c_int64, c_void_p, c_int, c_int, c_void_p]),
    "gl_knn_destroy": (c_int, [c_void_p]),
    "gl_knn_local_slots": (c_int64, [c_void_p]),
    "gl_knn_padded_slots": (c_int64, [c_void_p]),
    "gl_knn_set_fused": (c_int, [c_void_p, c_int]),
    "gl_knn_comm_handle": (c_int, [c_void_p, c_void_p]),
    "gl_knn_comm_connect": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "gl_knn_mark_far_ends": (c_int, [c_void_p, c_void_p, c_void_p]),
    "gl_knn_comm_need": (c_int, [c_void_p, c_void_p, c_void_p]),
    "gl_knn_begin": (c_int, [c_void_p, c_double, c_double, c_double, c_int, c_double, c_int, c_void_p]),
    "gl_knn_edge_step": (c_int, [c_void_p, c_void_p]),
    "gl_knn_node_step": (c_int, [c_void_p, c_void_p]),
    "gl_knn_decide": (c_int, [c_void_p, c_int, c_void_p]),
    "gl_knn_ab_ptr": (c_void_p, [c_void_p, POINTER(c_int64)]),
    "gl_knn_scalars_ptr": (c_void_p, [c_void_p]),
    "gl_knn_status": (c_int, [c_void_p, _pi, _pd, _pd, _pi, c_void_p]),
    "gl_knn_flags_ptr": (c_void_p, [c_void_p]),
    "gl_knn_store": (c_int, [c_void_p, c_void_p, c_double, c_void_p]),
    "gl_knn_build": (c_int, [c_void_p, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "gl_quadform_create": (c_int, [POINTER(c_void_p), c_void_p, c_int, c_int, c_int, c_double, c_void_p]),
    "gl_quadform_destroy": (c_int, [c_void_p]),
    "gl_quadform_nnz": (c_int64, [c_void_p]),
    "gl_quadform_apply": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p]),
    "gl_knn_logmodel_solve": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_double, c_double,
                                      c_int, c_double, c_double, c_double, c_void_p, _pi, _pd, _pd, c_void_p]),
    "gl_knn_logmodel_solve_ordered": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_double, c_double,
                                              c_int, c_double, c_double, c_double, c_void_p, _pi, _pd, _pd, c_void_p,
                                              c_int, c_int, c_void_p]),
}


def build(verbose: bool = False) -> str:
    """Compile libgl_b200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libgl_b200.so failed (see output above)")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C graph_learning_b200/csrc`.  There is no CPU fallback for this path."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class GLError(RuntimeError):
    pass


def check(rc: int) -> None:
    """Map C-ABI error codes onto the exceptions the reference raises."""
    if rc == GL_OK:
        return
    msg = lib().gl_last_error().decode("utf-8", "replace")
    if rc == GL_E_SHAPE:
        raise ValueError(msg)
    raise GLError(f"gl_b200 error {rc}: {msg}")


def launch_count() -> int:
    return int(lib().gl_launch_count())
