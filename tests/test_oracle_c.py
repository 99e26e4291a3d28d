"""Pin the C oracle to the NumPy oracle (itself pinned to the reference) and to the goldens."""
import numpy as np

from oracle import c_oracle
from oracle import logmodel_oracle as orc


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / np.linalg.norm(b)


def test_c_solver_matches_golden(golden):
    g = golden("solver_n100")
    z = g["z"] * g["theta"]
    for its in (1, 2, 5, 50):
        w, it, _, _ = c_oracle.solve(z, maxit=its, threshold=False)
        assert it == its
        assert rel(orc.to_square(w), g[f"w_raw_it{its}"]) < 1e-13
    w, it, rp, rd = c_oracle.solve(z)
    _, info = orc.solve_two_pass(z)
    assert it == info.n_iter
    assert rel(orc.to_square(w), g["w_final"]) < 1e-12
    assert abs(rp - info.rel_primal) < 1e-9 * rp and abs(rd - info.rel_dual) < 1e-9 * rd
    w, _, _, _ = c_oracle.solve(z, 0.7, 1.9, maxit=400, tol=1e-6, step_size=0.8, edge_tol=1e-4)
    assert rel(orc.to_square(w), g["w_alt"]) < 1e-12
    w, _, _, _ = c_oracle.solve(z, edge_init=g["edge_init"], maxit=30)
    assert rel(orc.to_square(w), g["w_warm"]) < 1e-12


def test_c_pdist(golden):
    g = golden("logmodel_n60")
    assert rel(c_oracle.pdist_sq(g["x"]), g["pdist_sq"]) < 1e-14
