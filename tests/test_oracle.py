"""CPU tests of the oracle: pinned against the reference's only golden numbers (README.md:135-159)
and against an independent numpy restatement (Kronecker-assembled stiffness matrix)."""
import json
import os

import numpy as np
import pytest

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_table.json")))


def _mesh_for(n):
    nsub, nref = n, 0
    if nsub > 1:
        while nsub % 2 == 0:
            nsub //= 2
            nref += 1
    return nsub, nref  # poisson_cube/program.cc:532-539


@pytest.mark.parametrize("row", [r for r in GOLDEN["rows"] if r[1] <= 300000])
def test_readme_table(orc, row):
    cells, dofs, reduction, fmg_l2, cg_l2, cg_its, cg_red = row
    nsub, nref = _mesh_for(round(cells ** (1 / 3)))
    s = orc.MultigridSolver(4, nsub, nref, degree_pre=3, degree_post=3, n_cycles=2)
    assert s.n_dofs() == dofs
    red, _ = s.solve(True)
    e_fmg = s.compute_l2_error()
    its, rate, _ = s.solve_cg()
    e_cg = s.compute_l2_error()
    assert its == cg_its
    # discretisation error: converged CG solution, 4 printed digits
    assert abs(e_cg - cg_l2) <= 1.5e-3 * cg_l2
    if nref > 0:
        # algebraic quantities depend (weakly) on the numbering-dependent eigenvalue estimate
        assert abs(e_fmg - fmg_l2) <= 0.03 * fmg_l2
        assert abs(red - reduction) <= 0.10 * reduction
        assert abs(rate - cg_red) <= 0.05 * cg_red
    else:
        assert abs(e_fmg - fmg_l2) <= 1.5e-3 * fmg_l2


def test_readme_config1(orc):
    """config 1: 28^3 cells, 1 442 897 DoF (README.md:150)."""
    row = [r for r in GOLDEN["rows"] if r[0] == 21952][0]
    s = orc.MultigridSolver(4, 7, 2, degree_pre=3, degree_post=3, n_cycles=2)
    assert s.n_dofs() == row[1]
    s.solve(False)
    e_fmg = s.compute_l2_error()
    its, rate, _ = s.solve_cg()
    e_cg = s.compute_l2_error()
    assert its == 8
    assert abs(e_cg - row[4]) <= 1e-3 * row[4]
    assert abs(e_fmg - row[3]) <= 5e-3 * row[3]
    assert abs(rate - row[6]) <= 0.05 * row[6]


def test_fmg_trace_first_levels(orc):
    """README.md:75-102: the coarse part of the 135 M DoF FMG trace is mesh-size independent."""
    s = orc.MultigridSolver(4, 1, 4, degree_pre=3, degree_post=3, n_cycles=2)
    _, trace = s.solve(True)
    for lvl in (1, 2, 3, 4):
        ref = GOLDEN["fmg_trace_135M"]["levels"][str(lvl)]
        assert np.allclose(trace[lvl], ref, rtol=0.03), (lvl, trace[lvl], ref)
    # level 1 is reproduced to the printed digits
    assert np.allclose(trace[1], GOLDEN["fmg_trace_135M"]["levels"]["1"], rtol=2e-4)


# ---------------------------------------------------------------- independent numpy check
def _gll(p):
    from numpy.polynomial import legendre as L
    c = np.zeros(p + 1)
    c[p] = 1
    x = np.concatenate(([-1.0], np.sort(L.legroots(L.legder(c))), [1.0])) if p > 1 else np.array([-1.0, 1.0])
    return 0.5 * (x + 1)


def _lagrange(nodes, x):
    V = np.ones((len(x), len(nodes)))
    dV = np.zeros_like(V)
    for i in range(len(nodes)):
        for l in range(len(nodes)):
            if l != i:
                V[:, i] *= (x - nodes[l]) / (nodes[i] - nodes[l])
        for m in range(len(nodes)):
            if m != i:
                t = np.ones(len(x)) / (nodes[i] - nodes[m])
                for l in range(len(nodes)):
                    if l != i and l != m:
                        t *= (x - nodes[l]) / (nodes[i] - nodes[l])
                dV[:, i] += t
    return V, dV


def _assembled_1d(p, ncell, h):
    from numpy.polynomial import legendre as L
    xg, wg = L.leggauss(p + 3)  # over-integrate: exact for mass and stiffness
    xg, wg = 0.5 * (xg + 1), 0.5 * wg
    V, dV = _lagrange(_gll(p), xg)
    Me = V.T @ (wg[:, None] * V) * h
    Ke = dV.T @ (wg[:, None] * dV) / h
    N = ncell * p + 1
    M, K = np.zeros((N, N)), np.zeros((N, N))
    for c in range(ncell):
        sl = slice(c * p, c * p + p + 1)
        M[sl, sl] += Me
        K[sl, sl] += Ke
    return M, K


@pytest.mark.parametrize("p,ncell", [(1, (2, 2, 2)), (2, (2, 1, 2)), (3, (2, 2, 1)), (4, (2, 2, 2)), (5, (1, 2, 1))])
def test_oracle_against_assembled_matrix(orc, p, ncell):
    lo, hi = (-0.9, -0.5, 0.1), (1.0, 0.7, 2.0)
    s = orc.MultigridSolver(p, ncell, 0, lower=lo, upper=hi)
    mats = [_assembled_1d(p, ncell[d], (hi[d] - lo[d]) / ncell[d]) for d in range(3)]
    (Mx, Kx), (My, Ky), (Mz, Kz) = mats
    A = np.kron(Mz, np.kron(My, Kx)) + np.kron(Mz, np.kron(Ky, Mx)) + np.kron(Kz, np.kron(My, Mx))
    nd = s.dof_grid()
    idx = np.arange(A.shape[0]).reshape(nd[2], nd[1], nd[0])
    bnd = np.zeros_like(idx, dtype=bool)
    bnd[0], bnd[-1], bnd[:, 0], bnd[:, -1], bnd[:, :, 0], bnd[:, :, -1] = (True,) * 6
    bnd = bnd.ravel()
    A[bnd, :] = 0
    A[:, bnd] = 0
    A[bnd, bnd] = 1
    x = np.random.default_rng(1).random(A.shape[0])
    y = s.vmult(x)
    assert np.abs(y - A @ x).max() <= 1e-12 * np.abs(A @ x).max()
    d = s.inverse_diagonal_dp()
    assert np.allclose(d, 1 / np.diag(A), rtol=1e-12)
    r = s.vmult_residual(np.ones_like(x), x)
    assert np.abs(r - (1 - A @ x)).max() <= 1e-12 * np.abs(A @ x).max()


def test_transfer_is_interpolation_and_adjoint(orc):
    s = orc.MultigridSolver(3, (1, 2, 1), 1)
    rng = np.random.default_rng(0)
    c, f = rng.random(s.n_dofs(0)), rng.random(s.n_dofs(1))
    Pc = s.prolongate(c, 1)
    Rf = s.restrict_and_add(f, 1)
    assert abs(Pc @ f - c_dot(c, Rf, s, 0)) <= 1e-5 * abs(Pc @ f)
    # no-boundary prolongation reproduces a trilinear function exactly
    nd0, nd1 = s.dof_grid(0), s.dof_grid(1)
    b = orc.basis(3)["gll"]

    def coords(nd, ncell):
        out = []
        for d in range(3):
            pts = np.concatenate([(cc + b[:-1]) for cc in range(ncell[d])] + [[ncell[d]]]) / ncell[d]
            out.append(pts)
        return np.meshgrid(out[2], out[1], out[0], indexing="ij")

    z0, y0, x0 = coords(nd0, (1, 2, 1))
    z1, y1, x1 = coords(nd1, (2, 4, 2))
    fun = lambda x, y, z: 1 + 2 * x - y + 0.5 * z + x * y * z
    fine = s.prolongate_no_boundary(fun(x0, y0, z0).ravel(), 1)
    assert np.allclose(fine, fun(x1, y1, z1).ravel(), atol=1e-13)


def c_dot(c, Rf, s, level):
    # constrained transfer: boundary entries of the coarse vector do not take part
    nd = s.dof_grid(level)
    m = np.ones((nd[2], nd[1], nd[0]), dtype=bool)
    m[0], m[-1], m[:, 0], m[:, -1], m[:, :, 0], m[:, :, -1] = (False,) * 6
    return (c * m.ravel()) @ Rf


def test_chebyshev_matches_textbook_recurrence(orc):
    s = orc.MultigridSolver(2, 2, 1, degree_pre=4, degree_post=4)
    info = s.chebyshev_info(1)
    n = s.n_dofs(1)
    b = np.random.default_rng(3).random(n)
    P = s.get_vector(orc.VEC_INV_DIAG, 1)
    A = lambda v: s.vmult(v, 1, double=False)
    theta, delta = info["theta"], info["delta"]
    # step() from a nonzero guess
    x0 = np.random.default_rng(4).random(n)
    x_old, x = x0.copy(), x0 + P * (b - A(x0)) / theta
    rho = delta / theta
    sigma = theta / delta
    for k in range(info["degree"] - 1):
        rho_new = 1 / (2 * sigma - rho)
        x_new = x + rho_new * rho * (x - x_old) + 2 * rho_new / delta * P * (b - A(x))
        x_old, x, rho = x, x_new, rho_new
    got = s.chebyshev_step(x0, b, 1)
    assert np.abs(got - x).max() <= 2e-5 * np.abs(x).max()


def _even_odd(A):
    """deal.II's even-odd compression of a 1-D table A[i][q] (basis function i at point q), the layout
    `apply_1d_matvec_kernel` reads (matrix_vector_kernel.h:47-101): rows i < n/2 hold the even parts
    (A[i][q] + A[i][n-1-q]) / 2, rows n-1-i the odd parts, the middle row the plain values; columns q < (n+1)/2."""
    n = A.shape[0]
    off, mid = (n + 1) // 2, n // 2
    eo = np.zeros((n, off))
    for i in range(mid):
        for q in range(off):
            eo[i, q] = 0.5 * (A[i, q] + A[i, n - 1 - q])
            eo[n - 1 - i, q] = 0.5 * (A[i, q] - A[i, n - 1 - q])
    if n % 2 == 1:
        eo[mid, :] = A[mid, :off]
    return np.ascontiguousarray(eo)


@pytest.mark.parametrize("p", range(1, 9))
def test_reference_even_odd_kernel(orc, p):
    """The one piece of the reference that can be executed here: common/matrix_vector_kernel.h (compiled from where it
    lies into oracle/_ref/libref1d.so by `make -C oracle ref`).  Its even-odd 1-D product, fed with the oracle's own
    tables S (values of the Gauss-Lobatto basis in the Gauss points) and D (collocation derivative), must reproduce the
    dense products the oracle and the CUDA kernels' host tables are built from -- evaluation (contract over rows) and
    integration (the transpose), types 0 (symmetric) and 1 (anti-symmetric)."""
    import ctypes as C
    import subprocess
    here = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")
    so = os.path.join(here, "_ref", "libref1d.so")
    if os.path.exists("/root/reference/common/matrix_vector_kernel.h"):
        subprocess.check_call(["make", "-C", here, "ref"], stdout=subprocess.DEVNULL)
    if not os.path.exists(so):
        pytest.skip("oracle/_ref/libref1d.so not built and /root/reference absent")
    lib = C.CDLL(so)
    dp = C.POINTER(C.c_double)
    lib.ref_apply_1d.argtypes = [C.c_int, C.c_int, C.c_int, dp, dp, dp]
    b = orc.basis(p)
    n = p + 1
    x = np.random.default_rng(p).random(n)
    out = np.zeros(n)

    def ref(table_iq, typ, rows):
        eo = _even_odd(table_iq)
        assert lib.ref_apply_1d(n, typ, int(rows), eo.ctypes.data_as(dp), x.ctypes.data_as(dp), out.ctypes.data_as(dp)) == 0
        return out.copy()

    S, D = b["S"], b["D"]            # S[q][i], D[q][r]
    # the symmetries the compressed layout relies on
    assert np.allclose(S, S[::-1, ::-1], atol=1e-14) and np.allclose(D, -D[::-1, ::-1], atol=1e-13)
    assert np.allclose(ref(S.T, 0, True), S @ x, atol=1e-14)       # evaluate: out[q] = sum_i S[q][i] x[i]
    assert np.allclose(ref(S.T, 0, False), S.T @ x, atol=1e-14)    # integrate: out[i] = sum_q S[q][i] x[q]
    assert np.allclose(ref(D.T, 1, True), D @ x, atol=1e-12)       # collocation derivative
    assert np.allclose(ref(D.T, 1, False), D.T @ x, atol=1e-12)    # and its transpose
    # G = D S is the gradient table the oracle's cell kernel applies (oracle/basis1d.h)
    x_q = ref(S.T, 0, True)
    x, keep = x_q, x
    assert np.allclose(ref(D.T, 1, True), b["G"] @ keep, atol=1e-11)  # evaluate then differentiate, as the reference does
