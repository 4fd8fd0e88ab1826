"""GPU parity of the pure DG h-multigrid solver (csrc/dg_plain_solver.cuh through the C ABI, mgb_dg_plain_solver_*)
against the oracle's composition (oracle/orc.py MultigridSolverDGPlain): the level transfer on the same vectors, the
smoother set-up of every level, one preconditioner cycle on the same vector, and solve_cg -- iteration count,
residual history, reduction rate, L2 error and solution."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("no GPU")
    return t


CASES = [  # degree, n_subdivisions, n_refinements, lower, upper, degree_pre, basis
    (2, (1, 1, 1), 2, (-1.,) * 3, (1.,) * 3, 3, 1),
    (3, (1, 1, 1), 2, (-1.,) * 3, (1.,) * 3, 3, 0),
    (4, (2, 1, 1), 1, (-1.,) * 3, (3., 1., 1.), 3, 2),     # doubling mesh, remainder 1 (poisson_dg_plain/program.cc:357-376)
    (1, (3, 3, 3), 1, (-0.9,) * 3, (1.,) * 3, 2, 1),       # subdivided_hyper_cube(3, -0.9, 1) (:390)
    (2, (2, 2, 1), 2, (-1.,) * 3, (3., 3., 1.), 4, 0),     # remainder 2, another smoother degree
]


@pytest.mark.parametrize("case", CASES)
def test_transfer_matches_oracle(mgb, orc, torch, case):
    p, nsub, nref, lower, upper, pre, basis = case
    g = mgb.MultigridSolverDGPlain(p, nsub, nref, lower, upper, pre, pre, basis=basis)
    o = orc.MultigridSolverDGPlain(p, nsub, nref, lower, upper, pre, pre, basis=basis)
    assert g.n_levels() == o.n_levels and g.n_dofs() == o.n_dofs()
    rng = np.random.default_rng(p)
    for level in range(1, o.n_levels):
        nc, nf = o.ops[level - 1].n_dofs(), o.ops[level].n_dofs()
        assert (g.level_n_dofs(level - 1), g.level_n_dofs(level)) == (nc, nf)
        u, v = rng.standard_normal(nc).astype(np.float32), rng.standard_normal(nf).astype(np.float32)
        du, dv = torch.tensor(u, device="cuda"), torch.tensor(v, device="cuda")
        out_f, out_c = torch.zeros(nf, dtype=torch.float32, device="cuda"), torch.zeros(nc, dtype=torch.float32, device="cuda")
        g.prolongate(level, out_f, du)
        ref = o.prolongate(level, u)
        assert np.abs(out_f.cpu().numpy() - ref).max() <= 2e-6 * np.abs(ref).max()
        g.prolongate(level, out_f, du, add=True)
        assert np.abs(out_f.cpu().numpy() - 2 * ref).max() <= 4e-6 * np.abs(ref).max()
        g.restrict(level, out_c, dv)
        ref = o.restrict(level, v)
        assert np.abs(out_c.cpu().numpy() - ref).max() <= 4e-6 * np.abs(ref).max()
        g.restrict(level, out_c, dv, add=True)
        assert np.abs(out_c.cpu().numpy() - 2 * ref).max() <= 8e-6 * np.abs(ref).max()


@pytest.mark.parametrize("case", CASES)
def test_solve_cg_matches_oracle(mgb, orc, torch, case):
    p, nsub, nref, lower, upper, pre, basis = case
    g = mgb.MultigridSolverDGPlain(p, nsub, nref, lower, upper, pre, pre, basis=basis)
    o = orc.MultigridSolverDGPlain(p, nsub, nref, lower, upper, pre, pre, basis=basis)
    rhs = g.get_vector(1)
    assert np.abs(rhs - o.rhs).max() <= 1e-12 * np.abs(o.rhs).max()
    for level in range(o.n_levels):
        a, b = g.smoother_info(level), o.smooth[level]
        assert abs(a["max_eig"] - b.max_eig) <= 2e-3 * b.max_eig, (level, a)
        if level > 0:
            assert a["degree"] == b.degree and a["cg_iterations"] == b.cg_iterations == 15
        else:
            # coarse solver: the float Lanczos process does not see the 1e-10 residual the double one stops at, and the
            # degree follows from the smallest eigenvalue through a floor()
            assert abs(a["min_eig"] - b.min_eig) <= 2e-2 * b.min_eig, (a, b.min_eig)
            assert abs(a["degree"] - b.degree) <= 1
    for tol in (1e-3, 1e-8):
        est_o, rate_o, hist_o = o.solve_cg(tol)
        est_g, rate_g = g.solve_cg(tol)
        hist_g = g.residual_history()
        assert g.last_step == o.last_step, (hist_g, hist_o)
        assert np.abs(hist_g - hist_o).max() <= 2e-2 * np.abs(hist_o).max()
        assert np.all(np.abs(hist_g - hist_o) <= 0.1 * hist_o + 1e-6 * hist_o[0])   # float cycle: floor 1e-6
        assert abs(rate_g - rate_o) <= 3e-2 * rate_o
        assert abs(est_g - est_o) <= 2e-2 * est_o
        eg, eo = g.compute_l2_error(), o.compute_l2_error()
        assert abs(eg - eo) <= 1e-3 * eo
    xs, xo = g.get_solution(), o.solution
    assert np.abs(xs - xo).max() <= 1e-6 * np.abs(xo).max()


def test_cycle_matches_oracle(mgb, orc, torch):
    p, nsub, nref = 3, (1, 2, 1), 2
    g = mgb.MultigridSolverDGPlain(p, nsub, nref, (-1.,) * 3, (1., 3., 1.))
    o = orc.MultigridSolverDGPlain(p, nsub, nref, (-1.,) * 3, (1., 3., 1.), basis=orc.DG_BASIS_HERMITE)
    rng = np.random.default_rng(3)
    src = rng.standard_normal(o.n_dofs())
    dst = torch.zeros(g.n_dofs(), dtype=torch.float64, device="cuda")
    g.vmult(dst, torch.tensor(src, device="cuda"))
    ref = o.vmult(src)
    assert np.abs(dst.cpu().numpy() - ref).max() <= 2e-4 * np.abs(ref).max()   # float cycle against a double one
    a, b = torch.tensor(rng.standard_normal(o.n_dofs()), device="cuda"), torch.tensor(src, device="cuda")
    Ma, Mb = torch.zeros_like(a), torch.zeros_like(a)
    g.vmult(Ma, a)
    g.vmult(Mb, b)
    assert abs(float(Ma @ b) - float(a @ Mb)) <= 1e-4 * abs(float(Ma @ b))
    assert float(a @ Ma) > 0


def test_host_call_and_errors(mgb, torch):
    g = mgb.MultigridSolverDGPlain(2, (1, 1, 1), 2)
    out = np.zeros(g.n_dofs())
    est, rate = g.solve_cg_host(1e-6, out)
    assert 0 < rate < 0.4 and g.last_step >= 3
    assert np.array_equal(out, g.get_solution())
    assert g.launch_count() > 0
    g.do_matvec()
    g.do_matvec_smoother()
    g.sync()
    with pytest.raises(RuntimeError):
        g.solve_cg(2.0)
    with pytest.raises(RuntimeError):
        mgb.MultigridSolverDGPlain(2, (1, 1, 1), 1, degree_pre=3, degree_post=2)
    with pytest.raises(RuntimeError):
        mgb.MultigridSolverDGPlain(9, (1, 1, 1), 1)
    with pytest.raises(RuntimeError):
        g.smoother_info(7)


def test_l2_error_at_scale(mgb, torch):
    """degree 4 on 16^3 / 32^3 cells: the discretisation error of the sine solution, halving h -> 2^-5, at an
    iteration count that does not grow with the mesh"""
    errs, its = [], []
    for nref in (3, 4):
        g = mgb.MultigridSolverDGPlain(4, (1, 1, 1), nref)
        g.solve_cg(1e-9)
        errs.append(g.compute_l2_error())
        its.append(g.last_step)
    assert 24 < errs[0] / errs[1] < 40
    assert its[1] <= its[0] + 2 and max(its) <= 25


@pytest.mark.parametrize("factor", [0., -0.37])
def test_vmult_with_residual_update(mgb, torch, factor):
    """multigrid_solver_dg_plain.h:340-427: the merged call equals its parts -- cycle on residual + factor * update, the two vector updates, the two sums"""
    g = mgb.MultigridSolverDGPlain(2, (1, 1, 1), 2)
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
