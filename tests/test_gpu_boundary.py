"""The drop-in boundary, exercised the way a reference maintainer would bind it (INTEGRATION.md
route B): ``integration/_gl_b200.py`` is a NumPy + ctypes stub with no torch and nothing from the
Python host layer; it calls ``gl_logmodel_solve_host`` (host buffers in, host buffers out) and
replaces ``gsp_learn_graph_log_degrees`` (reference graph_learn/smooth_learning.py:47-56).  Checked
against the goldens of the unmodified reference.  ``gl_logmodel_solve`` (device buffers) is driven
through raw ctypes as well.
"""
import ctypes
import importlib.util
import os

import numpy as np
import pytest

from oracle import logmodel_oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def stub():
    spec = importlib.util.spec_from_file_location("_gl_b200_stub", os.path.join(ROOT, "integration", "_gl_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("its", [1, 5, 50])
def test_stub_iterates_1d(golden, stub, its):
    g = golden("solver_n100")
    w = stub.gsp_learn_graph_log_degrees(g["z"] * g["theta"], 1, 1, maxit=its, edge_tol=-np.inf)
    assert w.shape == (100, 100) and stub.last_info["n_iter"] == its
    assert rel(w, g[f"w_raw_it{its}"]) < 1e-12


def test_stub_converged_1d_fp64_and_fp32(golden, stub):
    g = golden("solver_n100")
    z = g["z"] * g["theta"]
    _, oinfo = orc.solve_two_pass(z)
    w = stub.gsp_learn_graph_log_degrees(z, 1, 1)
    assert stub.last_info["n_iter"] == oinfo.n_iter
    assert rel(w, g["w_final"]) < 1e-9
    assert np.array_equal(w > 0, g["w_final"] > 0)
    assert abs(stub.last_info["rel_primal"] - oinfo.rel_primal) < 1e-6 * oinfo.rel_primal
    w32 = stub.gsp_learn_graph_log_degrees(z, 1, 1, dtype=np.float32)
    assert w32.dtype == np.float32 and abs(stub.last_info["n_iter"] - oinfo.n_iter) <= 2
    assert rel(w32, g["w_final"]) < 1e-4
    ref = g["w_final"]
    band = np.abs(ref - 1e-3) < 1e-5
    assert np.array_equal((w32 > 0)[~band], (ref > 0)[~band])


def test_stub_parameters_and_warm_start_1d(golden, stub):
    g = golden("solver_n100")
    z = g["z"] * g["theta"]
    w = stub.gsp_learn_graph_log_degrees(z, 0.7, 1.9, maxit=400, tol=1e-6, step_size=0.8, edge_tol=1e-4)
    assert rel(w, g["w_alt"]) < 1e-9
    w = stub.gsp_learn_graph_log_degrees(z, 1, 1, edge_init=g["edge_init"].copy(), maxit=30)
    assert rel(w, g["w_warm"]) < 1e-11


def test_stub_batched_2d(golden, stub):
    """(K, E) distances -> (K, E) weights with the shared spectral stop; edge_init arrives in the
    reference's internal (E, K) layout and is transposed by the stub."""
    g = golden("batched_n30")
    z = g["z"] * g["theta"]
    for its in (1, 7, 40):
        w = stub.gsp_learn_graph_log_degrees(z, 1, 1, maxit=its, edge_tol=-np.inf)
        assert w.shape == z.shape
        assert rel(w, g[f"w_raw_it{its}"]) < 1e-12
    w = stub.gsp_learn_graph_log_degrees(z, 1, 1)
    _, oinfo = orc.solve_two_pass(z)
    assert stub.last_info["n_iter"] == oinfo.n_iter
    assert rel(w, g["w"]) < 1e-9
    assert g["edge_init"].shape == (z.shape[1], z.shape[0])
    w = stub.gsp_learn_graph_log_degrees(z, 1, 1, edge_init=g["edge_init"].copy(), maxit=25)
    assert rel(w, g["w_warm"]) < 1e-11


def test_stub_errors(stub):
    with pytest.raises(ValueError, match="vector or matrix"):
        stub.gsp_learn_graph_log_degrees(np.zeros((2, 2, 3)), 1, 1)
    # a single "edge" list of length 1 means n = 2 (fine); an empty one is a shape error (code 1)
    with pytest.raises(ValueError):
        stub.gsp_learn_graph_log_degrees(np.zeros((1, 0)), 1, 1)


def test_host_entry_error_codes(stub):
    lib = stub._lib
    z = np.ones(3)
    w = np.empty(3)
    it, rp, rd = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
    args = (1.0, 1.0, None, 5, 1e-5, 0.5, 1e-3, w.ctypes.data, ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd))
    assert lib.gl_logmodel_solve_host(z.ctypes.data, None, 1, 1, 1, *args) == 1      # GL_E_SHAPE: n < 2
    assert b"2 nodes" in lib.gl_last_error()
    assert lib.gl_logmodel_solve_host(z.ctypes.data, None, 3, 0, 1, *args) == 1      # GL_E_SHAPE: K < 1
    assert lib.gl_logmodel_solve_host(z.ctypes.data, None, 3, 1, 7, *args) == 2      # GL_E_ARG: dtype
    assert lib.gl_logmodel_solve_host(z.ctypes.data, None, 3, 1, 1, *args) == 0
    assert it.value == 5


def test_host_entry_z_scale(golden, stub):
    """theta passed as z_scale_host (K doubles multiplied into Z on load: smooth_learning.py:289)."""
    g = golden("batched_n30")
    z = np.ascontiguousarray(g["z"])
    k, e = z.shape
    sc = (ctypes.c_double * k)(*g["theta"].ravel())
    w = np.empty_like(z)
    it, rp, rd = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
    rc = stub._lib.gl_logmodel_solve_host(z.ctypes.data, sc, 30, k, 1, 1.0, 1.0, None, 1000, 1e-5, 0.5, 1e-3,
                                          w.ctypes.data, ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd))
    assert rc == 0
    assert rel(w, g["w"]) < 1e-9


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_device_entry(golden, dtype):
    """gl_logmodel_solve on device arrays (torch only provides the memory)."""
    from graph_learning_b200 import _native as nat
    lib = nat.lib()
    g = golden("solver_n100")
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    z = torch.as_tensor((g["z"] * g["theta"]).astype(dtype), device="cuda")
    w0 = torch.as_tensor(g["edge_init"].astype(dtype), device="cuda")
    w = torch.empty_like(z)
    it, rp, rd = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
    gdt = nat.GL_F64 if dtype == np.float64 else nat.GL_F32
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    nat.check(lib.gl_logmodel_solve(z.data_ptr(), None, 100, 1, gdt, 1.0, 1.0, None, 1000, 1e-5, 0.5, 1e-3,
                                    w.data_ptr(), ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd), st))
    ref = orc.to_packed(g["w_final"])
    assert w.dtype == tdt
    assert rel(w.cpu().numpy(), ref) < (1e-9 if dtype == np.float64 else 1e-4)
    nat.check(lib.gl_logmodel_solve(z.data_ptr(), None, 100, 1, gdt, 1.0, 1.0, w0.data_ptr(), 30, 1e-5, 0.5, 1e-3,
                                    w.data_ptr(), ctypes.byref(it), ctypes.byref(rp), ctypes.byref(rd), st))
    assert it.value == 30
    assert rel(w.cpu().numpy(), orc.to_packed(g["w_warm"])) < (1e-11 if dtype == np.float64 else 1e-4)
