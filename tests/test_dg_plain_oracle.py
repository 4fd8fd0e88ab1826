"""The oracle's pure DG h-multigrid solver (oracle/orc.py MultigridSolverDGPlain, restating
common/multigrid_solver_dg_plain.h:54-590).  The reference ships no golden numbers for poisson_dg_plain, so the
restatement is pinned by the properties the algorithm must have: the level transfer reproduces polynomials of the
cell's space exactly and restriction is its transpose, the cycle is a symmetric positive preconditioner, CG converges
at a mesh-independent rate, the coarse Chebyshev solver reaches its tolerance, the error falls with order p+1."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def solver(orc):
    return orc.MultigridSolverDGPlain(2, (1, 1, 1), 2, basis=orc.DG_BASIS_HERMITE)


@pytest.mark.parametrize("basis", [0, 1, 2])
@pytest.mark.parametrize("degree", [1, 3])
def test_prolongation_embeds_the_coarse_space(orc, basis, degree):
    """a polynomial of degree p per direction, written in the coarse basis and prolongated, is the same function on
    the children: compared in the Gauss points of the fine cells"""
    s = orc.MultigridSolverDGPlain(degree, (1, 2, 1), 1, upper=(1., 3., 1.), basis=basis)
    n = degree + 1
    t = s.op.tables
    S, xq = t["S"], t["xq"]
    rng = np.random.default_rng(degree)
    co = rng.standard_normal((3, n))

    def poly(x, y, z):   # on the coarse mesh of cells [0,1] x [0,2] x [0,1] in cell units
        return np.polyval(co[0], x) * np.polyval(co[1], y) * np.polyval(co[2], z)
    Sinv = np.linalg.inv(S)
    ncx, ncy, ncz = s.ops[0].ncell
    coarse = np.zeros((ncz, ncy, ncx, n, n, n))
    for cz in range(ncz):
        for cy in range(ncy):
            for cx in range(ncx):
                g = poly(cx + xq[None, None, :], cy + xq[None, :, None], cz + xq[:, None, None])
                coarse[cz, cy, cx] = np.einsum("cr,bq,ap,rqp->cba", Sinv, Sinv, Sinv, g)
    fine = s.prolongate(1, coarse.ravel()).reshape(2 * ncz, 2 * ncy, 2 * ncx, n, n, n)
    for cz in range(2 * ncz):
        for cy in range(2 * ncy):
            for cx in range(2 * ncx):
                g = poly(0.5 * (cx + xq[None, None, :]), 0.5 * (cy + xq[None, :, None]), 0.5 * (cz + xq[:, None, None]))
                got = np.einsum("rc,qb,pa,cba->rqp", S, S, S, fine[cz, cy, cx])
                assert np.abs(got - g).max() <= 1e-11 * max(1., np.abs(g).max())
    u, v = rng.standard_normal(s.ops[0].n_dofs()), rng.standard_normal(s.ops[1].n_dofs())
    assert abs(s.prolongate(1, u) @ v - u @ s.restrict(1, v)) <= 1e-12 * np.linalg.norm(u) * np.linalg.norm(v)


def test_smoother_settings(solver):
    """:192-213"""
    assert [m.degree for m in solver.smooth[1:]] == [3, 2]
    assert [m.cg_iterations for m in solver.smooth[1:]] == [15, 15]
    c = solver.smooth[0]
    assert c.degree >= 8 and c.cg_iterations <= solver.ops[0].n_dofs()
    # the coarse solver is a solver: A x = b to the tolerance 1e-5 it was built for (in the norm Chebyshev bounds)
    rng = np.random.default_rng(2)
    b = rng.standard_normal(solver.ops[0].n_dofs())
    x = c.vmult(b)
    A = np.array([solver.ops[0].vmult(e) for e in np.eye(b.size)])
    exact = np.linalg.solve(A, b)
    assert np.linalg.norm(x - exact) <= 1e-4 * np.linalg.norm(exact)


def test_cycle_is_symmetric_positive(solver):
    rng = np.random.default_rng(5)
    a, b = rng.standard_normal(solver.n_dofs()), rng.standard_normal(solver.n_dofs())
    Ma, Mb = solver.vmult(a), solver.vmult(b)
    assert abs(Ma @ b - a @ Mb) <= 1e-8 * abs(Ma @ b)
    assert a @ Ma > 0 and b @ Mb > 0


def test_cg_rate_and_estimate(solver):
    est, rate, hist = solver.solve_cg(1e-3)
    assert solver.last_step == len(hist) - 1 and hist[-1] <= 1e-3 * hist[0] < hist[-2]
    assert rate < 0.3
    assert abs(est - np.log(1e-3) / np.log(rate)) < 1e-12
    assert np.linalg.norm(solver.rhs - solver.op.vmult(solver.solution)) <= 1.0001 * hist[-1]


def test_mesh_independence_and_order(orc):
    errs, its = [], []
    for nref in (2, 3):
        s = orc.MultigridSolverDGPlain(2, (1, 1, 1), nref, basis=orc.DG_BASIS_GAUSS_LOBATTO)
        s.solve_cg(1e-8)
        errs.append(s.compute_l2_error())
        its.append(s.last_step)
    assert 5.5 < errs[0] / errs[1] < 12    # order p + 1 = 3 -> factor 8 (pre-asymptotic on 4^3 cells)
    assert abs(its[0] - its[1]) <= 2 and max(its) <= 16
