"""The DG bases of the reference (template parameter `type` of LaplaceOperatorCompactCombine, laplace_operator_dg.h:346-350;
matvec_dg/program.cc:44-49; poisson_dg/program.cc:174 uses FE_DGQHermite) in the CPU oracle: the Hermite-like and the Gauss
basis span the same space as FE_DGQ, so every operator in those bases is the FE_DGQ one under the change of basis."""
import numpy as np
import pytest

KINDS = [1, 2]  # DG_BASIS_HERMITE, DG_BASIS_GAUSS


def test_hermite_like_basis_known_values(orc):
    """Polynomials::HermiteLikeInterpolation: the hard-coded low degrees of deal.II and the defining properties"""
    t = orc.dg_basis(1, orc.DG_BASIS_HERMITE)
    assert np.allclose(t["f0"], [1, 0]) and np.allclose(t["f1"], [0, 1]) and np.allclose(t["e0"], [-1, 1])
    t = orc.dg_basis(2, orc.DG_BASIS_HERMITE)   # (x-1)^2, -2 x (x-1), x^2
    assert np.allclose(t["e0"], [-2, 2, 0]) and np.allclose(t["e1"], [0, -2, 2])
    r = orc._hermite_like_roots(3)               # -3.5 (x - 2/7)(x-1)^2, 5.5 x (x-1)^2, mirrored
    assert np.isclose(r[0][0][0], 2. / 7., rtol=1e-14) and np.isclose(r[0][1], -3.5) and np.isclose(r[1][1], 5.5)
    assert np.isclose(r[2][1], -5.5) and np.isclose(r[3][1], 3.5) and np.isclose(sorted(r[3][0])[2], 5. / 7.)
    for p in range(1, 9):
        t = orc.dg_basis(p, orc.DG_BASIS_HERMITE)
        n = p + 1
        assert np.allclose(t["f0"], np.eye(n)[0], atol=1e-14) and np.allclose(t["f1"], np.eye(n)[p], atol=1e-14)
        assert np.allclose(t["e0"][2:], 0, atol=1e-12) and np.allclose(t["e1"][:-2], 0, atol=1e-12)
        assert np.isclose(t["e0"][0], -t["e0"][1]) and np.isclose(t["e1"][p], -t["e1"][p - 1])
        assert np.allclose(t["S"], t["S"][::-1, ::-1]) and np.allclose(t["G"], -t["G"][::-1, ::-1])
        if p >= 3:   # entries (1,0) and (p,p-1) of the mass matrix vanish
            M = t["S"].T @ np.diag(t["wq"]) @ t["S"]
            assert abs(M[0, 1]) < 1e-15 and abs(M[p, p - 1]) < 1e-15
        assert np.linalg.cond(t["B"]) < 10
        # partition of unity (value 1 and slope 0 at both ends, 1 in the interior nodes: p + 1 conditions on degree p)
        assert np.allclose(t["S"].sum(axis=1), 1, atol=1e-14) and np.allclose(t["G"].sum(axis=1), 0, atol=1e-12)


def test_gauss_basis_is_collocated(orc):
    for p in range(1, 9):
        t = orc.dg_basis(p, orc.DG_BASIS_GAUSS)
        assert np.allclose(t["S"], np.eye(p + 1), atol=1e-15)
        assert np.allclose(t["G"], orc.basis(p)["D"], atol=1e-11)
        assert np.isclose(t["f0"].sum(), 1) and np.isclose(t["e0"].sum(), 0, atol=1e-10)


def change(v, M, n):
    return np.einsum("kz,jy,ix,czyx->ckji", M, M, M, v.reshape(-1, n, n, n)).ravel()


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("p,ncell", [(1, (2, 2, 2)), (2, (3, 2, 2)), (3, (2, 2, 2)), (4, (2, 1, 2)), (5, (1, 2, 1)), (8, (1, 1, 2))])
def test_operator_is_the_fe_dgq_operator_in_another_basis(orc, kind, p, ncell):
    """A_kind = T^T A_FE_DGQ T with T = B x B x B, B[j, i] = phi_i(Gauss-Lobatto point j)"""
    a = orc.DGOperator(p, ncell, basis=kind)
    ref = orc.DGOperator(p, ncell)
    B = a.tables["B"]
    rng = np.random.default_rng(5)
    x = rng.random(a.n_dofs()) - 0.5
    y = a.vmult(x)
    y_ref = change(ref.vmult(change(x, B, p + 1)), B.T, p + 1)
    assert np.abs(y - y_ref).max() <= 1e-12 * np.abs(y_ref).max()
    z = rng.random(a.n_dofs()) - 0.5
    assert abs(z @ y - x @ a.vmult(z)) <= 1e-12 * abs(z @ y)


@pytest.mark.parametrize("kind", KINDS)
def test_jacobi_transformed_is_basis_independent(orc, kind):
    """V D^-1 V^T with generalised eigenvectors: P_kind^-1 = T^-1 P^-1 T^-T"""
    p, ncell = 3, (2, 2, 1)
    a, ref = orc.DGOperator(p, ncell, basis=kind), orc.DGOperator(p, ncell)
    Ja, Jr = orc.JacobiTransformed(a), orc.JacobiTransformed(ref)
    assert np.allclose(Ja.eigenvalues, Jr.eigenvalues, rtol=1e-10)
    Binv = np.linalg.inv(a.tables["B"])
    r = np.random.default_rng(3).random(a.n_dofs())
    got = Ja.vmult(r)
    want = change(Jr.vmult(change(r, Binv.T, p + 1)), Binv, p + 1)
    assert np.abs(got - want).max() <= 1e-10 * np.abs(want).max()


@pytest.mark.parametrize("kind", KINDS)
def test_solver_iterations_and_error_do_not_depend_on_the_basis(orc, kind):
    """MultigridSolverDG on FE_DGQHermite (poisson_dg) / Gauss: preconditioned CG produces the same iterates in every
    basis (operator and cycle transform consistently); only the Euclidean norm of the residual in the stopping test
    depends on the basis, so the step count may differ by one and the results agree to the tolerance"""
    args = (2, 1, 2)
    s, ref = orc.MultigridSolverDG(*args, basis=kind), orc.MultigridSolverDG(*args)
    s.solve_cg(1e-8)
    ref.solve_cg(1e-8)
    assert abs(s.last_step - ref.last_step) <= 1
    assert np.isclose(s.compute_l2_error(), ref.compute_l2_error(), rtol=1e-5)
    assert np.allclose(change(s.solution, s.op.tables["B"], 3), ref.solution, rtol=0, atol=1e-6 * np.abs(ref.solution).max())
    # one cycle as an operator: M_kind^-1 = T^-1 M^-1 T^-T up to the eigenvalue estimate of the Chebyshev smoother
    # (15 CG steps from a start vector that is not the same function in every basis)
    Binv = np.linalg.inv(s.op.tables["B"])
    r = np.random.default_rng(9).random(s.n_dofs()) - 0.5
    want = change(ref.vmult(change(r, Binv.T, 3)), Binv, 3)
    assert np.abs(s.vmult(r) - want).max() <= 2e-3 * np.abs(want).max()


@pytest.fixture(scope="module")
def basis_dump(tmp_path_factory):
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path_factory.mktemp("dgb") / "dg_basis_dump")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-I", os.path.join(root, "multigrid_b200", "csrc"),
                           os.path.join(root, "tests", "cpp", "dg_basis_dump.cc"), "-o", exe])

    def run(p, kind):
        out = subprocess.check_output([exe, str(p), str(kind)], text=True)
        return {l.split()[0]: np.array(l.split()[1:], dtype=np.float64) for l in out.splitlines()}
    return run


@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("p", range(1, 9))
def test_product_tables_match_the_oracle(orc, basis_dump, p, kind):
    """host code of the library (csrc/basis.h: long double, Gegenbauer recurrence + bisection for the Jacobi roots) against
    the oracle's numpy / scipy construction: the kernels and the oracle work with the same basis functions"""
    t, d = orc.dg_basis(p, kind), basis_dump(p, kind)
    n = p + 1
    for k in ("S", "G", "B"):
        assert np.abs(d[k].reshape(n, n) - t[k]).max() <= 2e-12 * max(1., np.abs(t[k]).max()), k
    for k in ("f0", "f1", "e0", "e1", "xq", "wq"):
        assert np.abs(d[k] - t[k]).max() <= 2e-12 * max(1., np.abs(t[k]).max()), k
    assert np.abs(d["Binv"].reshape(n, n) @ t["B"] - np.eye(n)).max() <= 1e-12
    assert np.abs(d["D"].reshape(n, n) - orc.basis(p)["D"]).max() <= 1e-12 * np.abs(d["D"]).max()


def test_a_gauss_basis_solve_serves_a_hermite_caller(orc):
    """The plan of DESIGN.md section 8 for the next round, pinned on the oracle: with C = B_H^-1 B_G (Gauss coefficients ->
    Hermite coefficients, per direction) the right-hand side is a dual vector, b_G = C^T b_H, the solution a primal one,
    x_H = C x_G; the L2 errors agree and the residual norm of the Hermite stopping test is |C^-T r_G|."""
    args = (2, 1, 2)
    h = orc.MultigridSolverDG(*args, basis=orc.DG_BASIS_HERMITE)
    g = orc.MultigridSolverDG(*args, basis=orc.DG_BASIS_GAUSS)
    C = np.linalg.inv(h.op.tables["B"]) @ g.op.tables["B"]
    assert np.abs(change(h.rhs, C.T, 3) - g.rhs).max() <= 1e-12 * np.abs(g.rhs).max()
    h.solve_cg(1e-10)
    g.solve_cg(1e-10)
    assert np.abs(change(g.solution, C, 3) - h.solution).max() <= 1e-7 * np.abs(h.solution).max()
    assert np.isclose(g.compute_l2_error(), h.compute_l2_error(), rtol=1e-6)
    r_g = g.rhs - g.op.vmult(g.solution)
    r_h = h.rhs - h.op.vmult(change(g.solution, C, 3))
    assert np.isclose(np.linalg.norm(change(r_g, np.linalg.inv(C).T, 3)), np.linalg.norm(r_h), rtol=1e-6, atol=1e-12)
