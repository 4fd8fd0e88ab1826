"""Committed golden vectors (tests/golden/vectors.npz, written by tests/golden/make_golden.py from the CPU oracle with
seeded inputs): the oracle must keep reproducing them (CPU), and the library is compared with the file directly (GPU)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "vectors.npz"))


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_oracle_reproduces_golden_vectors(orc, mgb):
    now = make_golden.generate()
    assert sorted(now) == sorted(GOLD.files)
    for k in GOLD.files:
        # (the solver runs sum in OpenMP reductions: rounding differs with the number of threads)
        assert rel(now[k], GOLD[k]) <= (1e-9 if "solver" in k else 1e-13), k


@pytest.mark.gpu
@pytest.mark.parametrize("p", range(1, 9))
def test_library_against_golden_fe_q_and_dg(mgb, orc, p):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    nsub, nref = make_golden.fe_q_case(p)
    s = mgb.MultigridSolver(p, nsub, nref)
    lex = s.lex_map()
    x = np.zeros(s.n_dofs())
    x[lex] = make_golden.seeded(s.n_dofs(), p)
    for prec, dt, tol in ((mgb.F64, torch.float64, 1e-12), (mgb.F32, torch.float32, 1e-5)):
        dst = torch.zeros(s.n_dofs(), dtype=dt, device="cuda")
        (s.matrix_dp() if prec == mgb.F64 else s.matrix()).vmult(dst, torch.tensor(x, dtype=dt, device="cuda"))
        assert rel(dst.double().cpu().numpy()[lex], GOLD["fe_q_p%d" % p]) <= tol, ("fe_q", p, prec)
    for prec, dt, tol in ((mgb.F64, torch.float64, 1e-12), (mgb.F32, torch.float32, 1e-5)):
        g = mgb.LaplaceOperatorDG(p, (1, 1, 1), 1, trafo=orc.dg_trafo(), precision=prec)
        xd = make_golden.seeded(g.m(), 100 + p)
        dst = torch.zeros(g.m(), dtype=dt, device="cuda")
        g.vmult(dst, torch.tensor(xd, dtype=dt, device="cuda"))
        assert rel(dst.double().cpu().numpy(), GOLD["dg_p%d" % p]) <= tol, ("dg", p, prec)


@pytest.mark.gpu
def test_library_against_golden_adaptive(mgb):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from adaptive_cases import CASES, make
    for i, name in enumerate(sorted(CASES)):
        op = make(mgb, name)
        order = make_golden.sorted_order(op.dof_points())
        x = np.zeros(op.m())
        x[order] = make_golden.seeded(op.m(), 200 + i)
        dst = torch.zeros(op.m(), dtype=torch.float64, device="cuda")
        op.vmult(dst, torch.tensor(x, device="cuda"))
        assert rel(dst.cpu().numpy()[order], GOLD["adaptive_" + name]) <= 1e-12, name


@pytest.mark.gpu
@pytest.mark.parametrize("p", range(1, 9))
@pytest.mark.parametrize("name", ["hermite", "gauss"])
def test_library_against_golden_dg_bases(mgb, orc, p, name):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    kind = {"hermite": mgb.DG_BASIS_HERMITE, "gauss": mgb.DG_BASIS_GAUSS}[name]
    g = mgb.LaplaceOperatorDG(p, (2, 2, 2), 0, trafo=orc.dg_trafo(), basis=kind)
    x = make_golden.seeded(g.m(), 100 + p)
    y = g.initialize_dof_vector()
    g.vmult(y, torch.tensor(x, device="cuda"))
    assert rel(y.cpu().numpy(), GOLD["dg_%s_p%d" % (name, p)]) <= 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("prefix", ["dg_solver", "dg_hermite_solver"])
def test_library_against_golden_dg_solver(mgb, prefix):
    """MultigridSolverDG against the stored oracle run: the float cycle against the double one (2e-4), the same
    number of CG iterations, the residual history, the L2 error to 3 digits"""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    basis = mgb.DG_BASIS_HERMITE if "hermite" in prefix else mgb.DG_BASIS_GAUSS_LOBATTO
    g = mgb.MultigridSolverDG(*make_golden.DG_SOLVER_CASE, basis=basis)
    src = make_golden.seeded(g.n_dofs(), 300)
    dst = torch.zeros(g.n_dofs(), dtype=torch.float64, device="cuda")
    g.vmult(dst, torch.tensor(src, device="cuda"))
    assert rel(dst.cpu().numpy(), GOLD[prefix + "_cycle"]) <= 2e-4
    g.solve_cg(1e-8)
    hist, want = g.residual_history(), GOLD[prefix + "_history"]
    assert hist.size == want.size
    assert np.all(np.abs(hist - want) <= 0.1 * want + 1e-6 * want[0])
    assert abs(g.compute_l2_error() - GOLD[prefix + "_l2_error"][0]) <= 1e-3 * GOLD[prefix + "_l2_error"][0]
    assert rel(g.get_solution(), GOLD[prefix + "_solution"]) <= 1e-6
