"""GPU parity of the DG operator, its preconditioners, the DG <-> FE_Q transfer and the DG multigrid solver on the other two
bases of the reference (template parameter `type` of LaplaceOperatorCompactCombine, laplace_operator_dg.h:346-350):
FE_DGQHermite (type 0, what poisson_dg/program.cc:174 uses) and FE_DGQArbitraryNodes(QGauss) (type 2), against the oracle
with the tables of oracle/orc.py dg_basis()."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = {1: 1e-12, 0: 1e-5}
KINDS = [1, 2]  # DG_BASIS_HERMITE, DG_BASIS_GAUSS


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("no GPU")
    return t


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


MESHES = [  # p, n_subdivisions, n_refinements, sheared
    (1, (2, 2, 2), 1, True), (2, (1, 2, 1), 1, True), (3, (2, 1, 3), 1, True), (4, (1, 1, 1), 2, True),
    (4, (3, 2, 1), 0, False), (5, (2, 1, 1), 1, True), (6, (1, 2, 2), 0, True), (7, (2, 2, 1), 0, True), (8, (1, 1, 2), 1, True),
]


def make(mgb, orc, p, nsub, nref, sheared, precision, kind):
    trafo = orc.dg_trafo() if sheared else np.eye(3)
    g = mgb.LaplaceOperatorDG(p, nsub, nref, trafo=trafo, precision=precision, basis=kind)
    o = orc.DGOperator(p, [s << nref for s in nsub], trafo=trafo, basis=kind)
    assert g.m() == o.n_dofs()
    return g, o


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", MESHES)
def test_vmult_family(mgb, orc, torch, mesh, precision, kind):
    p, nsub, nref, sheared = mesh
    g, o = make(mgb, orc, p, nsub, nref, sheared, precision, kind)
    rng = np.random.default_rng(42)
    x, b = rng.random(g.m()) - 0.5, rng.random(g.m())
    dx = torch.tensor(x, dtype=g.dtype, device="cuda")
    db = torch.tensor(b, dtype=g.dtype, device="cuda")
    y = g.initialize_dof_vector()
    g.vmult(y, dx)
    ref = o.vmult(x)
    assert rel(y.double().cpu().numpy(), ref) <= TOL[precision], ("vmult", rel(y.double().cpu().numpy(), ref))
    g.vmult_residual(db, dx, y)
    assert rel(y.double().cpu().numpy(), b - ref) <= TOL[precision]
    r_, q_, p_, x_ = (rng.random(g.m()) for _ in range(4))
    dr, dq, dp, dxx = (torch.tensor(v, dtype=g.dtype, device="cuda") for v in (r_, q_, p_, x_))
    sums = g.vmult_with_cg_update(0.3, 0.7, dr, dq, dp, dxx)
    p_new = 0.7 * p_ + q_
    q_new = o.vmult(p_new)
    assert np.allclose(sums, [q_new @ p_new, r_ @ r_, q_new @ r_, q_new @ q_new], rtol=1e-11 if precision else 2e-5)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", [(2, (1, 2, 3), 0, True), (4, (2, 1, 2), 0, True), (3, (1, 1, 1), 1, False), (5, (1, 2, 1), 0, True)])
def test_preconditioners(mgb, orc, torch, mesh, precision, kind):
    """point Jacobi (exact diagonal in this basis) and JacobiTransformed, alone and inside the fused Chebyshev update"""
    p, nsub, nref, sheared = mesh
    g, o = make(mgb, orc, p, nsub, nref, sheared, precision, kind)
    rng = np.random.default_rng(11)
    n = g.m()
    r = rng.random(n) - 0.5
    dr = torch.tensor(r, dtype=g.dtype, device="cuda")
    z = g.initialize_dof_vector()
    g.set_preconditioner(g.POINT_JACOBI)
    g.preconditioner_vmult(z, dr)
    assert rel(z.double().cpu().numpy(), r / o.diagonal()) <= (1e-12 if precision else 1e-6)
    J = orc.JacobiTransformed(o)
    g.set_preconditioner(g.JACOBI_TRANSFORMED)
    g.preconditioner_vmult(z, dr)
    tol = 1e-11 if precision else 2e-5
    assert rel(z.double().cpu().numpy(), J.vmult(r)) <= tol
    b, x0, x1 = rng.random(n), rng.random(n), rng.random(n)
    f1, f2 = 0.6, 0.2
    db = torch.tensor(b, dtype=g.dtype, device="cuda")
    sol = torch.tensor(x1, dtype=g.dtype, device="cuda")
    old = torch.tensor(x0, dtype=g.dtype, device="cuda")
    g.vmult_with_chebyshev_update(db, 2, f1, f2, sol, old)
    ref = (1 + f1) * x1 - f1 * x0 + f2 * J.vmult(b - o.vmult(x1))
    assert rel(old.double().cpu().numpy(), ref) <= 2 * tol


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", [(4, (1, 1, 1), 2), (3, (2, 1, 3), 1), (2, (1, 2, 1), 2)])
def test_dg_fe_q_transfer(mgb, orc, torch, mesh, precision, kind):
    """prolongate_add_cg_to_dg / vmult_residual_and_restrict_to_cg with the change of basis of LocalBasisTransformer<dim,1,...>
    (laplace_operator_dg.h:106-141): DG coefficients from the FE_Q values in the Gauss-Lobatto points, restriction transposed"""
    p, nsub, nref = mesh
    lower, upper = (-0.9,) * 3, (1.0,) * 3
    s = mgb.MultigridSolver(p, nsub, nref, lower=lower, upper=upper)
    g = mgb.LaplaceOperatorDG(p, nsub, nref, lower=lower, upper=upper, trafo=np.eye(3), precision=precision, basis=kind)
    o = orc.DGOperator(p, [k << nref for k in nsub], lower=lower, upper=upper, trafo=np.eye(3), basis=kind)
    Binv = np.linalg.inv(o.tables["B"])
    L = s.n_levels - 1
    nd = s.dof_grid(L)
    nc = g.n_cells()
    n = p + 1
    lexmap = s.lex_map(L)
    cz, cy, cx, k, j, i = np.meshgrid(np.arange(nc[2]), np.arange(nc[1]), np.arange(nc[0]), np.arange(n), np.arange(n),
                                      np.arange(n), indexing="ij")
    lex = (((cz * p + k) * nd[1] + cy * p + j) * nd[0] + cx * p + i).ravel()
    mask = np.ones((nd[2], nd[1], nd[0]))
    mask[0], mask[-1], mask[:, 0], mask[:, -1], mask[:, :, 0], mask[:, :, -1] = (0,) * 6
    mask = mask.ravel()

    def change(v, M):
        return np.einsum("kz,jy,ix,czyx->ckji", M, M, M, v.reshape(-1, n, n, n)).ravel()
    rng = np.random.default_rng(21)
    cg_lex = rng.random(s.n_dofs(L))
    dg0 = rng.random(g.m())
    cg = np.zeros_like(cg_lex)
    cg[lexmap] = cg_lex
    d_dg = torch.tensor(dg0, dtype=g.dtype, device="cuda")
    g.prolongate_add_cg_to_dg(s, L, d_dg, torch.tensor(cg, dtype=g.dtype, device="cuda"))
    assert rel(d_dg.double().cpu().numpy(), dg0 + change((cg_lex * mask)[lex], Binv)) <= (1e-12 if precision else 2e-6)
    rhs, lhs = rng.random(g.m()), rng.random(g.m())
    out = torch.full((s.n_dofs(L),), 7.0, dtype=g.dtype, device="cuda")
    g.vmult_residual_and_restrict_to_cg(s, L, torch.tensor(rhs, dtype=g.dtype, device="cuda"),
                                        torch.tensor(lhs, dtype=g.dtype, device="cuda"), out)
    ref_lex = np.zeros(s.n_dofs(L))
    np.add.at(ref_lex, lex, change(rhs - o.vmult(lhs), Binv.T))
    ref_lex *= mask
    assert rel(out.double().cpu().numpy()[lexmap], ref_lex) <= (1e-12 if precision else 2e-5)


CASES = [  # degree, n_subdivisions, n_refinements, degree_pre, preconditioner
    (2, (1, 1, 1), 2, 3, 1), (3, (1, 1, 1), 2, 3, 0), (4, (2, 1, 1), 1, 3, 1),
]


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("case", CASES)
def test_solve_cg_matches_oracle(mgb, orc, torch, case, kind):
    """poisson_dg on its own basis (FE_DGQHermite): right-hand side, CG steps, residual history, L2 error and solution
    coefficients against the oracle's composition in the same basis"""
    p, nsub, nref, pre, prec = case
    lower, upper = (-1.,) * 3, tuple(float(2 * k - 1) for k in nsub)
    g = mgb.MultigridSolverDG(p, nsub, nref, lower, upper, pre, pre, prec, basis=kind)
    o = orc.MultigridSolverDG(p, nsub, nref, lower, upper, pre, pre, prec, basis=kind)
    rhs = g.get_vector(1)
    assert np.abs(rhs - o.rhs).max() <= 1e-12 * np.abs(o.rhs).max()
    gi = g.smoother_info()
    assert gi["degree"] == o.smooth.degree
    assert abs(gi["max_eig"] - o.smooth.max_eig) <= 2e-3 * o.smooth.max_eig
    for tol in (1e-3, 1e-8):
        est_o, rate_o, hist_o = o.solve_cg(tol)
        est_g, rate_g = g.solve_cg(tol)
        hist_g = g.residual_history()
        assert g.last_step == o.last_step, (hist_g, hist_o)
        assert np.abs(hist_g - hist_o).max() <= 2e-2 * np.abs(hist_o).max()
        eg, eo = g.compute_l2_error(), o.compute_l2_error()
        assert abs(eg - eo) <= 1e-3 * eo
    xs, xo = g.get_solution(), o.solution
    assert np.abs(xs - xo).max() <= 1e-6 * np.abs(xo).max()
    # the same problem on FE_DGQ: same discrete solution, hence the same L2 error
    ref = mgb.MultigridSolverDG(p, nsub, nref, lower, upper, pre, pre, prec)
    ref.solve_cg(1e-8)
    assert abs(ref.compute_l2_error() - eg) <= 1e-4 * eg
