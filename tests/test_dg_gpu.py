"""GPU parity of the DG operator (csrc/dg_kernels.cuh through the C ABI) against the face-based CPU oracle
(oracle/dg_oracle.cpp): matvec relative error <= 1e-12 in double and <= 1e-5 in float, as north_star asks."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = {1: 1e-12, 0: 1e-5}


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
    (4, (3, 2, 1), 0, True), (4, (1, 1, 1), 0, True), (4, (2, 2, 2), 1, False), (5, (2, 1, 1), 1, True),
    (6, (1, 2, 2), 0, True), (7, (2, 2, 1), 0, True), (8, (1, 1, 2), 1, True),
]


def make(mgb, orc, p, nsub, nref, sheared, precision):
    trafo = orc.dg_trafo() if sheared else np.eye(3)
    g = mgb.LaplaceOperatorDG(p, nsub, nref, trafo=trafo, precision=precision)
    o = orc.DGOperator(p, [s << nref for s in nsub], trafo=trafo)
    assert g.m() == o.n_dofs() and g.n_cells() == tuple(s << nref for s in nsub)
    return g, o


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", MESHES)
def test_vmult_family(mgb, orc, torch, mesh, precision):
    p, nsub, nref, sheared = mesh
    g, o = make(mgb, orc, p, nsub, nref, sheared, precision)
    rng = np.random.default_rng(42)
    x, b = rng.random(g.m()), rng.random(g.m())
    dx = torch.tensor(x, dtype=g.dtype, device="cuda")
    db = torch.tensor(b, dtype=g.dtype, device="cuda")
    y = g.initialize_dof_vector()
    g.vmult(y, dx)
    ref = o.vmult(x)
    assert rel(y.double().cpu().numpy(), ref) <= TOL[precision], ("vmult", rel(y.double().cpu().numpy(), ref))
    g.vmult_residual(db, dx, y)
    assert rel(y.double().cpu().numpy(), b - ref) <= TOL[precision]
    assert np.allclose(g.get_penalty(), o.penalty(), rtol=1e-14)
    # vmult_with_cg_update (laplace_operator_dg.h:862-905)
    r_, q_, p_, x_ = (rng.random(g.m()) for _ in range(4))
    dr, dq, dp, dxx = (torch.tensor(v, dtype=g.dtype, device="cuda") for v in (r_, q_, p_, x_))
    sums = g.vmult_with_cg_update(0.3, 0.7, dr, dq, dp, dxx)
    x_new, p_new = x_ + 0.3 * p_, 0.7 * p_ + q_
    q_new = o.vmult(p_new)
    sums_ref = [q_new @ p_new, r_ @ r_, q_new @ r_, q_new @ q_new]
    assert np.allclose(sums, sums_ref, rtol=1e-11 if precision else 2e-5)
    assert rel(dxx.double().cpu().numpy(), x_new) <= TOL[precision] and rel(dq.double().cpu().numpy(), q_new) <= TOL[precision]


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", [(2, (2, 2, 2), 0, True), (4, (2, 2, 2), 0, True), (3, (3, 2, 2), 0, False), (4, (1, 1, 1), 1, True)])
def test_diagonal_and_chebyshev(mgb, orc, torch, mesh, precision):
    p, nsub, nref, sheared = mesh
    g, o = make(mgb, orc, p, nsub, nref, sheared, precision)
    inv = g.get_matrix_diagonal_inverse().double().cpu().numpy()
    diag = o.diagonal()  # unit-vector probing of the oracle
    assert rel(inv, 1.0 / diag) <= (1e-12 if precision else 1e-6)
    rng = np.random.default_rng(7)
    n = g.m()
    b, x0, x1 = rng.random(n), rng.random(n), rng.random(n)
    f1, f2 = 0.6, 0.2  # matvec_dg_cheby/program.cc:119-121
    db = torch.tensor(b, dtype=g.dtype, device="cuda")
    sol = torch.tensor(x1, dtype=g.dtype, device="cuda")
    old = torch.tensor(x0, dtype=g.dtype, device="cuda")
    g.vmult_with_chebyshev_update(db, 2, f1, f2, sol, old)
    ref = (1 + f1) * x1 - f1 * x0 + f2 / diag * (b - o.vmult(x1))
    assert rel(old.double().cpu().numpy(), ref) <= 2 * TOL[precision]
    old = torch.tensor(x0, dtype=g.dtype, device="cuda")
    g.vmult_with_chebyshev_update(db, 1, f1, f2, sol, old)
    ref = (1 + f1) * x1 + f2 / diag * (b - o.vmult(x1))
    assert rel(old.double().cpu().numpy(), ref) <= 2 * TOL[precision]
    g.vmult_with_chebyshev_update(db, 0, f1, f2, sol, old)
    assert rel(sol.double().cpu().numpy(), f2 / diag * b) <= 2 * TOL[precision]


def test_properties_at_scale(mgb, torch):
    """131 M DoF (the matvec_dg benchmark size): symmetry and linearity of the kernel, float against double"""
    g = mgb.LaplaceOperatorDG(4, (2, 2, 1), 6, precision=1)   # s = 20: 128 x 128 x 64 cells
    n = g.m()
    assert n == 131072000
    gen = torch.Generator(device="cuda").manual_seed(1)
    x = torch.rand(n, dtype=torch.float64, device="cuda", generator=gen)
    y = torch.rand(n, dtype=torch.float64, device="cuda", generator=gen)
    Ax, Ay = torch.empty_like(x), torch.empty_like(x)
    g.vmult(Ax, x)
    g.vmult(Ay, y)
    xAy, yAx = torch.dot(x, Ay).item(), torch.dot(y, Ax).item()
    assert abs(xAy - yAx) <= 1e-11 * abs(xAy)
    assert torch.dot(x, Ax).item() > 0
    gf = mgb.LaplaceOperatorDG(4, (2, 2, 1), 6, precision=0)
    Axf = torch.empty(n, dtype=torch.float32, device="cuda")
    gf.vmult(Axf, x.float())
    assert (Axf.double() - Ax).abs().max().item() <= 1e-5 * Ax.abs().max().item()


def test_vmult_host(mgb, orc, torch):
    g, o = make(mgb, orc, 4, (2, 1, 1), 0, True, 1)
    x = np.random.default_rng(3).random(g.m())
    y = np.zeros_like(x)
    g.vmult_host(y, x)
    assert rel(y, o.vmult(x)) <= 1e-12


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", [(2, (1, 2, 3), 0, True), (4, (3, 2, 2), 0, True), (3, (1, 1, 1), 2, False), (4, (2, 1, 2), 1, True),
                                  (6, (2, 2, 1), 0, True)])
def test_jacobi_transformed(mgb, orc, torch, mesh, precision):
    """multigrid::JacobiTransformed (laplace_operator_dg.h:2028-2256): V D^-1 V^T per cell, alone and inside the fused
    Chebyshev update, against the numpy/scipy restatement (LAPACK eigenvectors, diagonals probed from the oracle)"""
    p, nsub, nref, sheared = mesh
    g, o = make(mgb, orc, p, nsub, nref, sheared, precision)
    J = orc.JacobiTransformed(o)
    g.set_preconditioner(g.JACOBI_TRANSFORMED)
    rng = np.random.default_rng(11)
    n = g.m()
    r = rng.random(n) - 0.5
    dr = torch.tensor(r, dtype=g.dtype, device="cuda")
    z = g.initialize_dof_vector()
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
    old = torch.tensor(x0, dtype=g.dtype, device="cuda")
    g.vmult_with_chebyshev_update(db, 1, f1, f2, sol, old)
    ref = (1 + f1) * x1 + f2 * J.vmult(b - o.vmult(x1))
    assert rel(old.double().cpu().numpy(), ref) <= 2 * tol
    g.vmult_with_chebyshev_update(db, 0, f1, f2, sol, old)
    assert rel(sol.double().cpu().numpy(), f2 * J.vmult(b)) <= 2 * tol
    # back to point Jacobi
    g.set_preconditioner(g.POINT_JACOBI)
    g.preconditioner_vmult(z, dr)
    assert rel(z.double().cpu().numpy(), r / o.diagonal()) <= (1e-12 if precision else 1e-6)


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", [(4, (1, 1, 1), 2), (3, (2, 1, 3), 1), (2, (1, 2, 1), 2), (5, (1, 1, 2), 1)])
def test_dg_fe_q_transfer(mgb, orc, torch, mesh, precision):
    """prolongate_add_cg_to_dg / vmult_residual_and_restrict_to_cg (laplace_operator_dg.h:851-860,1798-1819,1863-1894):
    the same-degree transfer is the FE_Q cell-to-DoF map; restriction is the transpose, Dirichlet DoFs excluded"""
    p, nsub, nref = mesh
    lower, upper = (-0.9,) * 3, (1.0,) * 3
    s = mgb.MultigridSolver(p, nsub, nref, lower=lower, upper=upper)
    g = mgb.LaplaceOperatorDG(p, nsub, nref, lower=lower, upper=upper, trafo=np.eye(3), precision=precision)
    o = orc.DGOperator(p, [k << nref for k in nsub], lower=lower, upper=upper, trafo=np.eye(3))
    L = s.n_levels - 1
    nd = s.dof_grid(L)
    nc = g.n_cells()
    n = p + 1
    lexmap = s.lex_map(L)
    # lexicographic FE_Q index of every DG DoF
    cz, cy, cx, k, j, i = np.meshgrid(np.arange(nc[2]), np.arange(nc[1]), np.arange(nc[0]), np.arange(n), np.arange(n),
                                      np.arange(n), indexing="ij")
    lex = (((cz * p + k) * nd[1] + cy * p + j) * nd[0] + cx * p + i).ravel()
    mask = np.ones((nd[2], nd[1], nd[0]))
    mask[0], mask[-1], mask[:, 0], mask[:, -1], mask[:, :, 0], mask[:, :, -1] = (0,) * 6
    mask = mask.ravel()
    rng = np.random.default_rng(21)
    cg_lex = rng.random(s.n_dofs(L))
    dg0 = rng.random(g.m())
    cg = np.zeros_like(cg_lex)
    cg[lexmap] = cg_lex
    tol = 1e-13 if precision else 1e-6
    d_dg = torch.tensor(dg0, dtype=g.dtype, device="cuda")
    g.prolongate_add_cg_to_dg(s, L, d_dg, torch.tensor(cg, dtype=g.dtype, device="cuda"))
    assert rel(d_dg.double().cpu().numpy(), dg0 + (cg_lex * mask)[lex]) <= tol
    # restriction of the residual
    rhs, lhs = rng.random(g.m()), rng.random(g.m())
    out = torch.full((s.n_dofs(L),), 7.0, dtype=g.dtype, device="cuda")  # must be overwritten (restricted_residual = 0 first)
    g.vmult_residual_and_restrict_to_cg(s, L, torch.tensor(rhs, dtype=g.dtype, device="cuda"),
                                        torch.tensor(lhs, dtype=g.dtype, device="cuda"), out)
    res = rhs - o.vmult(lhs)
    ref_lex = np.zeros(s.n_dofs(L))
    np.add.at(ref_lex, lex, res)
    ref_lex *= mask
    got = out.double().cpu().numpy()[lexmap]
    assert rel(got, ref_lex) <= (1e-12 if precision else 2e-5)
    # transpose property: <P c, d> = <c, P^T d> on unconstrained entries
    c2 = rng.random(s.n_dofs(L)) * mask
    assert abs(((c2)[lex]) @ res - c2 @ ref_lex) <= 1e-10 * abs(c2 @ ref_lex)


CG_MESHES = [  # p, n_subdivisions, n_refinements, sheared, basis
    (3, (2, 1, 1), 1, True, 2),      # solver_dg/program.cc: degree 3, FE_DGQArbitraryNodes(QGauss), sheared mesh
    (4, (1, 1, 1), 1, True, 0), (2, (2, 2, 1), 1, False, 1), (5, (1, 2, 1), 0, True, 2), (7, (1, 1, 1), 1, True, 2),
]


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("mesh", CG_MESHES)
def test_cg_solver_plain_and_interleaved(mgb, orc, torch, mesh, precision):
    """solver_dg/program.cc:180-262: n iterations of CG with the inverse lumped mass matrix as preconditioner; the
    cell-based and the interleaved loop against a numpy CG on the oracle's operator (the program's own verification is
    the difference of the variants, :229-230)"""
    p, nsub, nref, sheared, basis = mesh
    trafo = orc.dg_trafo() if sheared else np.eye(3)
    g = mgb.LaplaceOperatorDG(p, nsub, nref, trafo=trafo, precision=precision, basis=basis)
    nc = [s << nref for s in nsub]
    o = orc.DGOperator(p, nc, trafo=trafo, basis=basis)
    t = o.tables
    col = t["wq"] @ t["S"]                                        # (1, phi_i) in 1-D
    h = [(o.upper[d] - o.lower[d]) / nc[d] for d in range(3)]
    detJ = abs(np.linalg.det(np.asarray(trafo).reshape(3, 3) @ np.diag(h)))
    D = 1. / (detJ * np.einsum("k,j,i->kji", col, col, col).ravel())
    D = np.tile(D, o.n_dofs() // D.size)
    dD = g.lumped_mass_inverse()
    assert rel(dD.double().cpu().numpy(), D) <= (1e-14 if precision else 1e-6)
    rng = np.random.default_rng(7)
    b = rng.random(o.n_dofs())
    n_it = 12
    # reference CG
    x, r = np.zeros_like(b), b.copy()
    z = D * r
    pp, rz = z.copy(), r @ z
    for _ in range(n_it):
        v = o.vmult(pp)
        alpha = rz / (pp @ v)
        x += alpha * pp
        r -= alpha * v
        z = D * r
        rz_new = r @ z
        pp = z + (rz_new / rz) * pp
        rz = rz_new
    db = torch.tensor(b, dtype=g.dtype, device="cuda")
    tol = 1e-9 if precision else 2e-3
    out = {}
    for interleaved in (False, True):
        dx = g.initialize_dof_vector()
        its, res = g.solve_cg(dx, db, dD, n_it, interleaved)
        assert its == n_it
        out[interleaved] = dx.double().cpu().numpy()
        assert rel(out[interleaved], x) <= tol, (interleaved, rel(out[interleaved], x))
        assert abs(res - np.linalg.norm(r)) <= (1e-7 if precision else 2e-2) * np.linalg.norm(r)
    assert rel(out[True], out[False]) <= tol
