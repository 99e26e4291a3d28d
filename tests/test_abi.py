"""The C-ABI library loads without a GPU and exports every symbol include/gl_b200.h declares;
host-only entry points (sizes, shard planning) are checked against closed forms."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def nat():
    from graph_learning_b200 import _native
    if not os.path.exists(_native.LIB_PATH):
        _native.build()
    return _native


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gl_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gl_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(nat):
    lib = nat.lib()
    names = header_symbols()
    assert len(names) >= 30
    for name in names:
        assert hasattr(lib, name), f"{name} declared in gl_b200.h but not exported"
        assert name in nat.SIGNATURES, f"{name} has no ctypes signature in _native.py"
    assert sorted(nat.SIGNATURES) == names


def test_sizes(nat):
    lib = nat.lib()
    assert lib.gl_version() >= 100
    for n in (2, 100, 128, 129, 2000, 20000):
        e = n * (n - 1) // 2
        nt = -(-n // 128)
        assert lib.gl_num_edges(n) == e
        assert lib.gl_nodes_from_edges(e) == n
        assert lib.gl_num_tiles(n) == nt * (nt + 1) // 2
        assert lib.gl_tiled_elems(n) == nt * (nt + 1) // 2 * 128 * 128
    assert lib.gl_nodes_from_edges(7) == -1
    # Gram workspace (host-only arithmetic): 3 bf16 planes of (n_pad, m_pad) dominate
    need = lib.gl_gram_workspace_bytes(50000, 20000, 3)
    assert 3 * 20096 * 50048 * 2 < need < 3 * 20096 * 50048 * 2 + (64 << 20)
    assert lib.gl_gram_workspace_bytes(50000, 20000, 2) < need
    assert lib.gl_gram_workspace_bytes(10, 1, 3) == -1 and lib.gl_gram_workspace_bytes(10, 5, 4) == -1


@pytest.mark.parametrize("n,world", [(100, 2), (2000, 2), (20000, 8), (20000, 3), (300, 8)])
def test_shards_partition_the_triangle(nat, n, world):
    lib = nat.lib()
    total = lib.gl_num_tiles(n)
    nt = -(-n // 128)
    weights = np.array([1 if i == j else 2 for i in range(nt) for j in range(i, nt)])
    prev_end, loads = 0, []
    for r in range(world):
        b, e = ctypes.c_int(), ctypes.c_int()
        assert lib.gl_shard_tiles(n, r, world, ctypes.byref(b), ctypes.byref(e)) == 0
        assert b.value == prev_end and e.value >= b.value
        prev_end = e.value
        loads.append(weights[b.value:e.value].sum())
    assert prev_end == total
    if total >= 8 * world:
        assert max(loads) - min(loads) <= 0.02 * weights.sum() / world + 4


def test_errors_without_device(nat):
    """Argument errors are reported before any CUDA call; codes map to the reference's exceptions."""
    lib = nat.lib()
    h = ctypes.c_void_p()
    assert lib.gl_solver_create(ctypes.byref(h), 1, 1, 0, 0, 0) == nat.GL_E_SHAPE
    assert b"at least 2 nodes" in lib.gl_last_error()
    with pytest.raises(ValueError):
        nat.check(nat.GL_E_SHAPE)
    assert lib.gl_solver_create(ctypes.byref(h), 100, 1, 7, 0, 1) == nat.GL_E_ARG
    with pytest.raises(nat.GLError):
        nat.check(nat.GL_E_ARG)
    assert lib.gl_shard_tiles(100, 3, 2, None, None) == nat.GL_E_ARG
    # a batch cannot be tile-sharded (shared spectral stop rule): refused before any allocation
    assert lib.gl_solver_create(ctypes.byref(h), 300, 4, 0, 0, 2) == nat.GL_E_ARG
    assert b"cannot be tile-sharded" in lib.gl_last_error()


def test_theta_noalloc_limit_without_device(nat):
    """The allocation-free theta entry point refuses sizes beyond its static scratch before touching the device;
    the limit in the message is the header's GL_THETA_NOALLOC_MAX_NODES."""
    lib = nat.lib()
    src = open(os.path.join(ROOT, "include", "gl_b200.h")).read()
    limit = int(re.search(r"#define\s+GL_THETA_NOALLOC_MAX_NODES\s+(\d+)", src).group(1))
    th = ctypes.c_double(-1.0)
    dummy = ctypes.c_void_p(256)       # never dereferenced: the checks come first
    rc = lib.gl_theta_noalloc(dummy, nat.LAYOUT_PACKED, limit + 1, 3, nat.GL_F32, ctypes.byref(th), None)
    assert rc == nat.GL_E_SHAPE and f"at most {limit} nodes".encode() in lib.gl_last_error()
    assert lib.gl_theta_noalloc(dummy, 99, 100, 3, nat.GL_F32, ctypes.byref(th), None) == nat.GL_E_ARG
    assert lib.gl_theta_noalloc(dummy, nat.LAYOUT_PACKED, 1, 3, nat.GL_F32, ctypes.byref(th), None) == nat.GL_E_SHAPE
    assert lib.gl_theta_noalloc(dummy, nat.LAYOUT_PACKED, 100, 100, nat.GL_F32, ctypes.byref(th), None) == nat.GL_E_ARG
    assert lib.gl_theta_noalloc(None, nat.LAYOUT_PACKED, 100, 3, nat.GL_F32, ctypes.byref(th), None) == nat.GL_E_ARG
    assert lib.gl_theta_noalloc(dummy, nat.LAYOUT_PACKED, 100, 3, 9, ctypes.byref(th), None) == nat.GL_E_ARG
    assert th.value == -1.0
