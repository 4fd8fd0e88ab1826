"""GPU parity of the DG multigrid solver (csrc/dg_solver.cuh through the C ABI, mgb_dg_solver_*) against the
oracle's composition (oracle/orc.py MultigridSolverDG): same CG iteration count, residual history and reduction
rate, the L2 error to 3 digits, and one preconditioner cycle applied to the same vector."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("no GPU")
    return t


CASES = [  # degree, n_subdivisions, n_refinements, lower, upper, degree_pre, preconditioner
    (2, (1, 1, 1), 2, (-1.,) * 3, (1.,) * 3, 3, 1),
    (3, (1, 1, 1), 2, (-1.,) * 3, (1.,) * 3, 3, 0),
    (4, (2, 1, 1), 1, (-1.,) * 3, (3., 1., 1.), 3, 1),     # doubling mesh, remainder 1 (poisson_dg/program.cc:357-376)
    (1, (3, 3, 3), 1, (-0.9,) * 3, (1.,) * 3, 2, 1),       # subdivided_hyper_cube(3, -0.9, 1) (:390)
    (2, (1, 1, 1), 0, (-1.,) * 3, (1.,) * 3, 3, 1),        # one cell: the coarse solver alone
]


@pytest.mark.parametrize("case", CASES)
def test_solve_cg_matches_oracle(mgb, orc, torch, case):
    p, nsub, nref, lower, upper, pre, prec = case
    g = mgb.MultigridSolverDG(p, nsub, nref, lower, upper, pre, pre, prec)
    o = orc.MultigridSolverDG(p, nsub, nref, lower, upper, pre, pre, prec)
    assert g.n_dofs() == o.n_dofs()
    rhs = g.get_vector(1)
    assert np.abs(rhs - o.rhs).max() <= 1e-12 * np.abs(o.rhs).max()
    gi = g.smoother_info()
    assert gi["degree"] == o.smooth.degree and gi["cg_iterations"] == o.smooth.cg_iterations
    assert abs(gi["max_eig"] - o.smooth.max_eig) <= 2e-3 * o.smooth.max_eig
    L = o.fe.n_levels - 1
    for level in {0, L}:
        a, b = g.fe_smoother_info(level), o.fe.chebyshev_info(level)
        assert a["degree"] == b["degree"], (level, a, b)
        assert abs(a["max_eig"] - b["max_eig"]) <= 2e-3 * b["max_eig"]
    # one cell: the float cycle is a direct solve to 1e-7, the double oracle to 1e-16 -- compare the first tolerance only
    for tol in ((1e-3,) if nref == 0 else (1e-3, 1e-8)):
        est_o, rate_o, hist_o = o.solve_cg(tol)
        est_g, rate_g = g.solve_cg(tol)
        hist_g = g.residual_history()
        assert g.last_step == o.last_step, (hist_g, hist_o)
        assert np.abs(hist_g - hist_o).max() <= 2e-2 * np.abs(hist_o).max()
        assert np.all(np.abs(hist_g - hist_o) <= 0.1 * hist_o + 1e-6 * hist_o[0])   # float cycle: floor 1e-6
        if rate_o > 1e-4:   # (a direct coarse solve reaches the rounding floor of its precision in one step)
            assert abs(rate_g - rate_o) <= 3e-2 * rate_o
            assert abs(est_g - est_o) <= 2e-2 * est_o
        eg, eo = g.compute_l2_error(), o.compute_l2_error()
        assert abs(eg - eo) <= 1e-3 * eo
    xs, xo = g.get_solution(), o.solution
    assert np.abs(xs - xo).max() <= 1e-6 * np.abs(xo).max()


def test_cycle_matches_oracle(mgb, orc, torch):
    p, nsub, nref = 3, (1, 2, 1), 2
    g = mgb.MultigridSolverDG(p, nsub, nref, (-1.,) * 3, (1., 3., 1.))
    o = orc.MultigridSolverDG(p, nsub, nref, (-1.,) * 3, (1., 3., 1.))
    rng = np.random.default_rng(3)
    src = rng.standard_normal(o.n_dofs())
    dst = torch.zeros(g.n_dofs(), dtype=torch.float64, device="cuda")
    g.vmult(dst, torch.tensor(src, device="cuda"))
    ref = o.vmult(src)
    assert np.abs(dst.cpu().numpy() - ref).max() <= 2e-4 * np.abs(ref).max()   # float cycle against a double one
    # symmetric positive preconditioner
    a, b = torch.tensor(rng.standard_normal(o.n_dofs()), device="cuda"), torch.tensor(src, device="cuda")
    Ma, Mb = torch.zeros_like(a), torch.zeros_like(a)
    g.vmult(Ma, a)
    g.vmult(Mb, b)
    assert abs(float(Ma @ b) - float(a @ Mb)) <= 1e-4 * abs(float(Ma @ b))
    assert float(a @ Ma) > 0


def test_host_call_and_errors(mgb, torch):
    g = mgb.MultigridSolverDG(2, (1, 1, 1), 2)
    out = np.zeros(g.n_dofs())
    est, rate = g.solve_cg_host(1e-6, out)
    assert 0 < rate < 0.15 and g.last_step >= 3
    assert np.array_equal(out, g.get_solution())
    assert g.launch_count() > 0
    with pytest.raises(RuntimeError):
        g.solve_cg(2.0)
    with pytest.raises(RuntimeError):
        mgb.MultigridSolverDG(2, (1, 1, 1), 1, degree_pre=3, degree_post=2)
    with pytest.raises(RuntimeError):
        mgb.MultigridSolverDG(9, (1, 1, 1), 1)


def test_l2_error_at_scale(mgb, torch):
    """degree 4 on 16^3 cells (0.5 M DoFs): the discretisation error of the sine solution, halving h -> 2^-5"""
    errs = []
    for nref in (3, 4):
        g = mgb.MultigridSolverDG(4, (1, 1, 1), nref)
        g.solve_cg(1e-9)
        errs.append(g.compute_l2_error())
        assert g.last_step <= 12
    assert 24 < errs[0] / errs[1] < 40


@pytest.mark.parametrize("factor", [0., -0.37])
def test_vmult_with_residual_update(mgb, torch, factor):
    """multigrid_solver_dg.h:445-527: the merged call equals its parts -- cycle on residual + factor * update, the two vector updates, the two sums"""
    g = mgb.MultigridSolverDG(2, (1, 1, 1), 2)
    rng = np.random.default_rng(11)
    res0, upd0 = rng.standard_normal(g.n_dofs()), rng.standard_normal(g.n_dofs())
    res, upd = torch.tensor(res0, device="cuda"), torch.tensor(upd0, device="cuda")
    a, b = g.vmult_with_residual_update(res, upd, factor)
    new_res = res0 + factor * upd0
    mg = torch.zeros_like(res)
    g.vmult(mg, torch.tensor(new_res, device="cuda"))
    mg = mg.cpu().numpy()
    assert np.abs(res.cpu().numpy() - new_res).max() <= 1e-14 * np.abs(new_res).max()
    assert np.abs(upd.cpu().numpy() - mg).max() <= 1e-6 * np.abs(mg).max()
    assert abs(a - mg @ new_res) <= 1e-6 * abs(mg @ new_res)
    expect_b = mg @ (factor * upd0) if factor != 0. else mg @ new_res
    assert abs(b - expect_b) <= 1e-6 * abs(mg @ new_res)
