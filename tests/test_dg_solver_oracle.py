"""The oracle's DG multigrid solver (oracle/orc.py MultigridSolverDG, restating common/multigrid_solver_dg.h:44-735).
The reference ships no golden numbers for poisson_dg, so the restatement is pinned by the properties the algorithm
must have: the cycle is a symmetric positive preconditioner, CG converges at a mesh-independent rate, and the
discretisation error falls with order p+1."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def solver(orc):
    return orc.MultigridSolverDG(2, (1, 1, 1), 2)


def test_cycle_is_symmetric_positive(solver):
    rng = np.random.default_rng(5)
    a, b = rng.standard_normal(solver.n_dofs()), rng.standard_normal(solver.n_dofs())
    Ma, Mb = solver.vmult(a), solver.vmult(b)
    assert abs(Ma @ b - a @ Mb) <= 1e-4 * abs(Ma @ b)   # the FE_Q part runs in float
    assert a @ Ma > 0 and b @ Mb > 0


def test_cg_rate_and_estimate(solver):
    est, rate, hist = solver.solve_cg(1e-3)
    assert solver.last_step == len(hist) - 1 and hist[-1] <= 1e-3 * hist[0] < hist[-2]
    assert rate < 0.12
    assert abs(est - np.log(1e-3) / np.log(rate)) < 1e-12
    x = solver.solution
    assert np.linalg.norm(solver.rhs - solver.op.vmult(x)) <= 1.0001 * hist[-1]


def test_l2_error_order(orc):
    errs = []
    for nref in (2, 3):
        s = orc.MultigridSolverDG(2, (1, 1, 1), nref, preconditioner=0)
        s.solve_cg(1e-8)
        errs.append(s.compute_l2_error())
    assert 5.5 < errs[0] / errs[1] < 12    # order p + 1 = 3 -> factor 8 (pre-asymptotic on 4^3 cells)


def test_smoother_settings(orc):
    s = orc.MultigridSolverDG(3, (1, 1, 1), 2, degree_pre=3, degree_post=3, preconditioner=0)
    L = s.fe.n_levels - 1
    assert s.fe.chebyshev_info(L)["degree"] == 2 and s.fe.chebyshev_info(L)["cg_iterations"] == 15
    assert s.fe.chebyshev_info(1)["degree"] == 3
    assert s.smooth.degree == 3 and s.smooth.cg_iterations == 15
    assert 1.0 < s.smooth.max_eig < 4.0
