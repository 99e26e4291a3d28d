"""gl_theta_noalloc: the allocation-free variant of gl_theta (scratch in a file-scope static device array,
include/gl_b200.h) against the existing entry point, through the C-ABI.  No counterpart in the reference;
the statistic itself is get_theta, graph_learn/smooth_learning.py:201-219.

Collected last on purpose (file name): the suite runs with -x, so a failure here must not hide other tests.
The two entry points launch the same kernels on the same data with fixed-order reductions, so the results
agree to the last bit; the assertion uses the 1e-12 relative tolerance of the other theta tests.
"""
import ctypes

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

RTOL = 1e-12


@pytest.fixture(scope="module")
def nat():
    from graph_learning_b200 import _native as m
    return m


def _both(nat, z, layout, n, k, gdt):
    lib = nat.lib()
    a, b = ctypes.c_double(), ctypes.c_double()
    nat.check(lib.gl_theta(z.data_ptr(), layout, n, k, gdt, ctypes.byref(a), None))
    nat.check(lib.gl_theta_noalloc(z.data_ptr(), layout, n, k, gdt, ctypes.byref(b), None))
    return a.value, b.value


def _tiles(nat, zp, n, gdt):
    lib = nat.lib()
    t = torch.zeros(lib.gl_tiled_elems(n), dtype=zp.dtype, device="cuda")
    nat.check(lib.gl_pack_to_tiles(zp.data_ptr(), t.data_ptr(), n, 1, gdt, 0, lib.gl_num_tiles(n), None))
    return t


def _random_packed(n, seed, dtype):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((24, n))
    return torch.as_tensor(orc.pdist_sq(x), device="cuda").to(dtype)


def test_noalloc_matches_gl_theta_two_calls_in_a_row(nat, golden):
    """Two calls back to back on different problems, the second smaller than the first so the scratch
    still holds the first call's rows beyond its own: both must equal gl_theta (and the golden value)."""
    n1, k1 = 700, 8                                   # 6 x 6 tiles, float32, tiled layout
    z1 = _tiles(nat, _random_packed(n1, 0, torch.float32), n1, nat.GL_F32)
    ref1, got1 = _both(nat, z1, nat.LAYOUT_TILED, n1, k1, nat.GL_F32)
    assert np.isfinite(ref1) and abs(got1 - ref1) <= RTOL * abs(ref1)

    g = golden("theta_n40")                           # float64, packed layout, reference's own values
    zp = torch.as_tensor(orc.to_packed(g["zsq"]), device="cuda")
    for k, want in zip(g["ks"][1:], g["thetas"][1:]):
        ref2, got2 = _both(nat, zp, nat.LAYOUT_PACKED, 40, int(k), nat.GL_F64)
        assert abs(got2 - ref2) <= RTOL * abs(ref2)
        assert abs(got2 - want) <= RTOL * abs(want)

    # and the first problem again: same answer as before, bit for bit
    th = ctypes.c_double()
    nat.check(nat.lib().gl_theta_noalloc(z1.data_ptr(), nat.LAYOUT_TILED, n1, k1, nat.GL_F32, ctypes.byref(th), None))
    assert th.value == got1


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_noalloc_packed_and_tiled_on_a_side_stream(nat, dtype):
    tdt = getattr(torch, dtype)
    gdt = nat.GL_F32 if dtype == "float32" else nat.GL_F64
    n, k = 333, 5
    zp = _random_packed(n, 3, tdt)
    zt = _tiles(nat, zp, n, gdt)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    lib = nat.lib()
    vals = []
    for z, layout in ((zp, nat.LAYOUT_PACKED), (zt, nat.LAYOUT_TILED)):
        ref, got = _both(nat, z, layout, n, k, gdt)
        assert abs(got - ref) <= RTOL * abs(ref)
        s = ctypes.c_double()
        nat.check(lib.gl_theta_noalloc(z.data_ptr(), layout, n, k, gdt, ctypes.byref(s), ctypes.c_void_p(st.cuda_stream)))
        assert s.value == got
        vals.append(got)
    assert abs(vals[0] - vals[1]) <= RTOL * abs(vals[0])
    want = orc.theta_for_degree(orc.to_square(zp.double().cpu().numpy()), k)
    assert abs(vals[0] - want) <= 1e-12 * abs(want)


def test_noalloc_refuses_sizes_beyond_the_static_scratch(nat):
    lib = nat.lib()
    z = torch.zeros(8, dtype=torch.float32, device="cuda")     # never read: the size check comes first
    th = ctypes.c_double(-1.0)
    with pytest.raises(ValueError, match="at most 65536 nodes"):
        nat.check(lib.gl_theta_noalloc(z.data_ptr(), nat.LAYOUT_TILED, 65537, 8, nat.GL_F32, ctypes.byref(th), None))
    assert th.value == -1.0
    with pytest.raises(nat.GLError):
        nat.check(lib.gl_theta_noalloc(z.data_ptr(), nat.LAYOUT_PACKED, 40, 40, nat.GL_F32, ctypes.byref(th), None))
    with pytest.raises(nat.GLError):
        nat.check(lib.gl_theta_noalloc(None, nat.LAYOUT_PACKED, 40, 3, nat.GL_F32, ctypes.byref(th), None))
    torch.cuda.synchronize()
