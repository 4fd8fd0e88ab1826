"""ORACLE (test infrastructure, not product code): ctypes view of oracle/_build/liborc.so.

CPU restatement of the reference's FE_Q multigrid path (see oracle/mg_oracle.cpp for the
file:line map). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liborc.so")


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("mg_oracle.cpp", "dg_oracle.cpp", "basis1d.h", "Makefile")]
    if (not force and os.path.exists(_LIB)
            and all(os.path.getmtime(_LIB) >= os.path.getmtime(s) for s in src)):
        return _LIB
    subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def omp_threads(n=None):
    """set (n given) and return the thread count the OpenMP runtime of the oracle uses"""
    if n is not None:
        lib().orc_omp_set_threads(int(n))
    return int(lib().orc_omp_max_threads())


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        dp = C.POINTER(C.c_double)
        ip = C.POINTER(C.c_int)
        L.orc_mg_create.restype = C.c_void_p
        L.orc_mg_create.argtypes = [C.c_int, ip, C.c_int, dp, dp, C.c_uint, C.c_uint, C.c_uint, C.c_int]
        L.orc_mg_create_mapped.restype = C.c_void_p
        L.orc_mg_create_mapped.argtypes = [C.c_int, ip, C.c_int, dp, dp, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, dp]
        L.orc_mg_quadrature_points.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_support_points.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_set_coefficient_values.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_reinit_smoothers.argtypes = [C.c_void_p]
        L.orc_mg_set_rhs_values.argtypes = [C.c_void_p, C.c_int, dp, dp]
        L.orc_mg_set_exact_values.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_destroy.argtypes = [C.c_void_p]
        L.orc_mg_ms_set_boundary_values.argtypes = [C.c_void_p, dp]
        L.orc_mg_ms_newton_step.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_ms_residual.argtypes = [C.c_void_p, dp, C.c_int, dp]
        L.orc_mg_ms_level_solution.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_n_levels.argtypes = [C.c_void_p]
        L.orc_mg_n_dofs.restype = C.c_longlong
        L.orc_mg_n_dofs.argtypes = [C.c_void_p, C.c_int]
        L.orc_mg_dof_grid.argtypes = [C.c_void_p, C.c_int, ip]
        L.orc_mg_solve.restype = C.c_double
        L.orc_mg_solve.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_solve_cg.restype = C.c_double
        L.orc_mg_solve_cg.argtypes = [C.c_void_p, C.POINTER(C.c_uint), dp, C.c_int]
        L.orc_mg_l2_error.restype = C.c_double
        L.orc_mg_l2_error.argtypes = [C.c_void_p, C.c_int]
        L.orc_mg_vmult.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp]
        L.orc_mg_vmult_residual.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp, dp]
        L.orc_mg_vmult_with_cg_update.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, dp, dp, dp, dp, dp]
        L.orc_mg_get_vector.argtypes = [C.c_void_p, C.c_int, C.c_int, dp]
        L.orc_mg_inverse_diagonal_dp.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_chebyshev_info.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_mg_chebyshev_apply.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp]
        L.orc_mg_smoother_reinit.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_int]
        L.orc_mg_transfer.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp]
        L.orc_mg_vcycle.argtypes = [C.c_void_p, dp, dp]
        L.orc_mg_time_matvec.restype = C.c_double
        L.orc_mg_time_matvec.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_omp_max_threads.restype = C.c_int
        L.orc_omp_set_threads.argtypes = [C.c_int]
        L.orc_mg_setup_times.argtypes = [C.c_void_p, dp]
        L.orc_basis.argtypes = [C.c_int] + [dp] * 7
        L.dgo_create.restype = C.c_void_p
        L.dgo_create.argtypes = [C.c_int, ip, dp, dp, dp]
        L.dgo_destroy.argtypes = [C.c_void_p]
        L.dgo_set_basis.argtypes = [C.c_void_p] + [dp] * 6
        L.dgo_n_dofs.restype = C.c_longlong
        L.dgo_n_dofs.argtypes = [C.c_void_p]
        L.dgo_penalty.argtypes = [C.c_void_p, dp]
        L.dgo_vmult.argtypes = [C.c_void_p, dp, dp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def basis(p):
    n = p + 1
    S, G, D = (np.zeros((n, n)) for _ in range(3))
    P1 = np.zeros((2 * p + 1, n))
    gll, xq, wq = (np.zeros(n) for _ in range(3))
    lib().orc_basis(p, _dp(S), _dp(G), _dp(D), _dp(P1), _dp(gll), _dp(xq), _dp(wq))
    return dict(S=S, G=G, D=D, P1=P1, gll=gll, xq=xq, wq=wq)


DG_BASIS_GAUSS_LOBATTO, DG_BASIS_HERMITE, DG_BASIS_GAUSS = 0, 1, 2   # reference `type` 1, 0, 2


def _hermite_like_roots(p):
    """Polynomials::HermiteLikeInterpolation of deal.II (1-D basis of FE_DGQHermite; call sites
    poisson_dg/program.cc:174, matvec_dg/program.cc:44-45), restated from its definition: every function in
    product form w prod (x - r_k).  phi_0 / phi_1 carry value / derivative at 0 (phi_0(0) = 1, phi_0'(0) + phi_1'(0) = 0,
    int phi_0 phi_1 = 0 fixes the extra root x* of phi_0), phi_p / phi_{p-1} mirror them, the interior functions have
    double roots at both ends and are 1 at one root of the Jacobi polynomial P^(4,4)_{p-3} (scipy.special.roots_jacobi)
    and 0 at the others.  deal.II itself is not available here: PARITY UNPINNED for the Hermite coefficients, the
    space and the operator are pinned by the change-of-basis tests (tests/test_dg_basis.py)."""
    if p == 1:
        return [([1.], -1.), ([0.], 1.)]
    if p == 2:
        return [([1., 1.], 1.), ([0., 1.], -2.), ([0., 0.], 1.)]
    jr = []
    if p > 3:
        import scipy.special
        jr = list(0.5 * (scipy.special.roots_jacobi(p - 3, 4, 4)[0] + 1.))
    # int_0^1 x (x - x*) common^2 = 0 with the shared factor common = prod (x - r_j) (x - 1)^2 of phi_0 and phi_1,
    # integrated with numpy's Gauss-Legendre rule on p + 2 points (exact)
    xg, wg = np.polynomial.legendre.leggauss(p + 2)
    xg, wg = 0.5 * (xg + 1.), 0.5 * wg
    c2 = xg * np.prod([(xg - rk) ** 2 for rk in jr + [1., 1.]], axis=0)
    xs = np.sum(wg * xg * c2) / np.sum(wg * c2)
    val = lambda r, w, x: w * np.prod([x - rk for rk in r])
    der = lambda r, w, x: w * sum(np.prod([x - r[k] for k in range(len(r)) if k != m]) for m in range(len(r)))
    r0, r1 = [xs] + jr + [1., 1.], [0.] + jr + [1., 1.]
    w0 = 1. / val(r0, 1., 0.)
    w1 = -der(r0, w0, 0.) / der(r1, 1., 0.)
    out = [(r0, w0), (r1, w1)]
    for i in range(2, p - 1):
        r = [0., 0.] + [jr[j] for j in range(p - 3) if j != i - 2] + [1., 1.]
        out.append((r, 1. / val(r, 1., jr[i - 2])))
    sign = -1. if p % 2 else 1.
    out.append(([1. - rk for rk in r1], w1 * sign))
    out.append(([1. - rk for rk in r0], w0 * sign))
    return out


def dg_basis(p, kind=DG_BASIS_GAUSS_LOBATTO):
    """1-D tables of a DG basis: S[q, i] = phi_i(x_q), G = phi_i'(x_q) at the Gauss points, f0/f1 = phi_i(0/1),
    e0/e1 = phi_i'(0/1), B[j, i] = phi_i(Gauss-Lobatto point j)"""
    b = basis(p)
    n = p + 1
    if kind == DG_BASIS_HERMITE:
        polys = _hermite_like_roots(p)
    else:
        x = b["gll"] if kind == DG_BASIS_GAUSS_LOBATTO else b["xq"]
        polys = [([x[l] for l in range(n) if l != i], 1. / np.prod([x[i] - x[l] for l in range(n) if l != i]))
                 for i in range(n)]

    def val(i, x):
        r, w = polys[i]
        return w * np.prod([x - rk for rk in r])

    def der(i, x):
        r, w = polys[i]
        return w * sum(np.prod([x - r[k] for k in range(len(r)) if k != m]) for m in range(len(r)))
    t = dict(xq=b["xq"], wq=b["wq"], gll=b["gll"], kind=kind)
    t["S"] = np.array([[val(i, x) for i in range(n)] for x in b["xq"]])
    t["G"] = np.array([[der(i, x) for i in range(n)] for x in b["xq"]])
    t["B"] = np.array([[val(i, x) for i in range(n)] for x in b["gll"]])
    t["f0"], t["f1"] = np.array([val(i, 0.) for i in range(n)]), np.array([val(i, 1.) for i in range(n)])
    # embedding of the cell's space into its two 1-D children (the prolongation matrices of the element, used by
    # MGTransferMatrixFree): E_o[j, i] = coefficient j on child o of phi_i restricted to it
    for o in (0, 1):
        t["E%d" % o] = np.linalg.solve(t["B"], np.array([[val(i, 0.5 * (o + x)) for i in range(n)] for x in b["gll"]]))
    t["e0"], t["e1"] = np.array([der(i, 0.) for i in range(n)]), np.array([der(i, 1.) for i in range(n)])
    for k in ("S", "G", "B", "f0", "f1", "e0", "e1", "E0", "E1"):
        t[k] = np.ascontiguousarray(np.where(np.abs(t[k]) < 1e-15, 0., t[k]))
    return t


VEC_SOLUTION, VEC_RHS, VEC_INV_DIAG, VEC_UPDATE, VEC_RESIDUAL, VEC_DEFECT, VEC_CHEB_GUESS = range(7)


class MultigridSolver:
    """multigrid_solver.h:96-782 restated on the CPU; vectors in lexicographic numbering."""

    def __init__(self, degree, n_subdivisions, n_refine, lower=(-0.9,) * 3, upper=(1.0,) * 3,
                 degree_pre=2, degree_post=2, n_cycles=2, problem=0, geometry=0, geometry_params=None):
        """geometry 0: box [lower, upper]; 1: one coarse cell of hyper_shell(0, r0, r1, 6) and its refinements,
        geometry_params = (r0, r1, opening half-angle), default (0.5, 1.0, pi/4); 2: affine, geometry_params = A (9)"""
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        self.degree = degree
        nsub = (C.c_int * 3)(*n_subdivisions)
        lo = (C.c_double * 3)(*lower)
        hi = (C.c_double * 3)(*upper)
        gp = [0.5, 1.0, np.pi / 4] + [0.0] * 9
        if geometry_params is not None:
            gp[:len(geometry_params)] = list(geometry_params)
        par = (C.c_double * 12)(*gp)
        self.h = lib().orc_mg_create_mapped(degree, nsub, n_refine, lo, hi, degree_pre, degree_post, n_cycles, problem,
                                            geometry, par)
        if not self.h:
            raise ValueError("unsupported degree")
        self.n_levels = lib().orc_mg_n_levels(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mg_destroy(self.h)
            self.h = None

    def n_dofs(self, level=-1):
        return lib().orc_mg_n_dofs(self.h, self._l(level))

    def dof_grid(self, level=-1):
        nd = (C.c_int * 3)()
        lib().orc_mg_dof_grid(self.h, self._l(level), nd)
        return tuple(nd)

    def _l(self, level):
        return level if level >= 0 else self.n_levels + level

    # ---- problem data given by samples (the Function<dim> arguments of multigrid_solver.h:100-106)
    def n_cells(self, level=-1):
        nd = self.dof_grid(level)
        return int(np.prod([(v - 1) // self.degree for v in nd]))

    def quadrature_points(self, level=-1):
        out = np.zeros((self.n_cells(level) * (self.degree + 1) ** 3, 3))
        lib().orc_mg_quadrature_points(self.h, self._l(level), _dp(out))
        return out

    def support_points(self, level=-1):
        out = np.zeros((self.n_dofs(level), 3))
        lib().orc_mg_support_points(self.h, self._l(level), _dp(out))
        return out

    def set_problem(self, coefficient=None, right_hand_side=None, boundary_values=None, exact_solution=None):
        ev = lambda f, x: np.ascontiguousarray(f(x[:, 0], x[:, 1], x[:, 2]), dtype=np.float64)
        if coefficient is not None:
            for l in range(self.n_levels):
                lib().orc_mg_set_coefficient_values(self.h, l, _dp(ev(coefficient, self.quadrature_points(l))))
            lib().orc_mg_reinit_smoothers(self.h)
        for l in range(self.n_levels):
            if right_hand_side is not None:
                f = ev(right_hand_side, self.quadrature_points(l))
                g = ev(boundary_values, self.support_points(l)) if boundary_values is not None else None
                lib().orc_mg_set_rhs_values(self.h, l, _dp(f), _dp(g) if g is not None else None)
            if exact_solution is not None:
                lib().orc_mg_set_exact_values(self.h, l, _dp(ev(exact_solution, self.quadrature_points(l))))

    # ---- minimal_surface/program.cc: Newton's method with line search on -div(grad u / sqrt(1 + |grad u|^2)) = 0
    def minimal_surface_set_boundary_values(self, boundary_values=None):
        """interpolate_boundary_values (:396-408); a callable g(x, y, z) or None = the solver's analytic solution"""
        g = None
        if boundary_values is not None:
            x = self.support_points()
            g = np.ascontiguousarray(boundary_values(x[:, 0], x[:, 1], x[:, 2]), dtype=np.float64)
        lib().orc_mg_ms_set_boundary_values(self.h, _dp(g) if g is not None else None)

    def minimal_surface_step(self, first_time):
        """LaplaceProblem::solve(first_time) (:412-572) -> dict(initial_residual, final_residual, linear_iterations,
        line_search_steps, step_length)"""
        info = np.zeros(5)
        lib().orc_mg_ms_newton_step(self.h, int(first_time), _dp(info))
        return dict(initial_residual=info[0], final_residual=info[1], linear_iterations=int(info[2]),
                    line_search_steps=int(info[3]), step_length=info[4])

    def minimal_surface_solve(self, max_steps=100, tolerance=1e-12):
        """the inner loop of run() (:652-659) -> list of the steps' dicts"""
        steps = []
        for it in range(max_steps):
            steps.append(self.minimal_surface_step(it == 0))
            if steps[-1]["final_residual"] < tolerance:
                break
        return steps

    def minimal_surface_residual(self, u, first_time=False):
        out = np.zeros(self.n_dofs())
        lib().orc_mg_ms_residual(self.h, _dp(np.ascontiguousarray(u, dtype=np.float64)), int(first_time), _dp(out))
        return out

    def minimal_surface_level_solution(self, level):
        out = np.zeros(self.n_dofs(level))
        lib().orc_mg_ms_level_solution(self.h, self._l(level), _dp(out))
        return out

    def solve(self, do_analyze=False):
        trace = np.zeros((self.n_levels, 4))
        r = lib().orc_mg_solve(self.h, int(do_analyze), _dp(trace))
        return (r, trace) if do_analyze else r

    def solve_cg(self):
        its = C.c_uint()
        hist = np.zeros(64)
        rate = lib().orc_mg_solve_cg(self.h, C.byref(its), _dp(hist), 64)
        return its.value, rate, hist[:its.value + 1]

    def compute_l2_error(self, level=-1):
        return lib().orc_mg_l2_error(self.h, self._l(level))

    def vmult(self, src, level=-1, double=True):
        src = np.ascontiguousarray(src, dtype=np.float64)
        dst = np.zeros_like(src)
        lib().orc_mg_vmult(self.h, self._l(level), int(double), _dp(src), _dp(dst))
        return dst

    def vmult_residual(self, rhs, lhs, level=-1, double=True):
        rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        lhs = np.ascontiguousarray(lhs, dtype=np.float64)
        res = np.zeros_like(rhs)
        lib().orc_mg_vmult_residual(self.h, self._l(level), int(double), _dp(rhs), _dp(lhs), _dp(res))
        return res

    def vmult_with_cg_update(self, alpha, beta, r, q, p, x, level=-1):
        r, q, p, x = (np.array(v, dtype=np.float64) for v in (r, q, p, x))
        sums = np.zeros(4)
        lib().orc_mg_vmult_with_cg_update(self.h, self._l(level), alpha, beta, _dp(r), _dp(q), _dp(p), _dp(x), _dp(sums))
        return sums, q, p, x

    def get_vector(self, which, level=-1):
        out = np.zeros(self.n_dofs(level))
        lib().orc_mg_get_vector(self.h, which, self._l(level), _dp(out))
        return out

    def inverse_diagonal_dp(self, level=-1):
        out = np.zeros(self.n_dofs(level))
        lib().orc_mg_inverse_diagonal_dp(self.h, self._l(level), _dp(out))
        return out

    def chebyshev_info(self, level):
        info = np.zeros(6)
        lib().orc_mg_chebyshev_info(self.h, self._l(level), _dp(info))
        return dict(min_eig=info[0], max_eig=info[1], theta=info[2], delta=info[3],
                    degree=int(info[4]), cg_iterations=int(info[5]))

    def smoother_reinit(self, level, smoothing_range, degree, eig_cg_n_iterations):
        """degree < 0: automatic degree (coarse solver); eig_cg_n_iterations < 0: as many as the level has DoFs"""
        lib().orc_mg_smoother_reinit(self.h, self._l(level), smoothing_range, degree, eig_cg_n_iterations)

    def chebyshev_vmult(self, b, level=-1):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros_like(b)
        lib().orc_mg_chebyshev_apply(self.h, self._l(level), 0, _dp(x), _dp(b))
        return x

    def chebyshev_step(self, x, b, level=-1):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.array(x, dtype=np.float64)
        lib().orc_mg_chebyshev_apply(self.h, self._l(level), 1, _dp(x), _dp(b))
        return x

    def prolongate(self, coarse, fine_level, add_to=None):
        coarse = np.ascontiguousarray(coarse, dtype=np.float64)
        fine = np.zeros(self.n_dofs(fine_level)) if add_to is None else np.array(add_to, dtype=np.float64)
        lib().orc_mg_transfer(self.h, self._l(fine_level), 0 if add_to is None else 1, _dp(fine), _dp(coarse))
        return fine

    def prolongate_no_boundary(self, coarse, fine_level):
        coarse = np.ascontiguousarray(coarse, dtype=np.float64)
        fine = np.zeros(self.n_dofs(fine_level))
        lib().orc_mg_transfer(self.h, self._l(fine_level), 3, _dp(fine), _dp(coarse))
        return fine

    def restrict_and_add(self, fine, fine_level, coarse=None):
        fine = np.ascontiguousarray(fine, dtype=np.float64)
        c = np.zeros(self.n_dofs(self._l(fine_level) - 1)) if coarse is None else np.array(coarse, dtype=np.float64)
        lib().orc_mg_transfer(self.h, self._l(fine_level), 2, _dp(c), _dp(fine))
        return c

    def vcycle(self, src):
        src = np.ascontiguousarray(src, dtype=np.float64)
        dst = np.zeros_like(src)
        lib().orc_mg_vcycle(self.h, _dp(dst), _dp(src))
        return dst

    def time_matvec(self, double=True, n_mv=50):
        return lib().orc_mg_time_matvec(self.h, int(double), n_mv)

    def setup_times(self):
        out = np.zeros(3)
        lib().orc_mg_setup_times(self.h, _dp(out))
        return out


def dg_trafo(shear=0.12):
    """vertex map of matvec_dg/program.cc:66-76: I + 0.12 (d+1)(e+1)"""
    return np.eye(3) + shear * np.outer(np.arange(1, 4), np.arange(1, 4))


DG_LOWER = tuple(-1. + 0.05 * (d + 1) for d in range(3))   # matvec_dg/program.cc:55-60
DG_UPPER = tuple(0.95 - 0.06 * d for d in range(3))


class DGOperator:
    """SIP-DG Laplacian of laplace_operator_dg_face.h on the affine box mesh of matvec_dg/program.cc;
    DoFs cell-wise contiguous, cells lexicographic (x fastest)."""

    def __init__(self, degree, ncell, lower=DG_LOWER, upper=DG_UPPER, trafo=None, basis=DG_BASIS_GAUSS_LOBATTO):
        self.degree, self.ncell = degree, tuple(ncell)
        self.tables = dg_basis(degree, basis)
        self.trafo = dg_trafo() if trafo is None else np.ascontiguousarray(trafo, dtype=np.float64)
        nc = (C.c_int * 3)(*ncell)
        lo = (C.c_double * 3)(*lower)
        hi = (C.c_double * 3)(*upper)
        self.lower, self.upper = tuple(lower), tuple(upper)
        self.h = lib().dgo_create(degree, nc, lo, hi, _dp(self.trafo))
        if not self.h:
            raise ValueError("unsupported degree")
        if basis != DG_BASIS_GAUSS_LOBATTO:
            t = self.tables
            lib().dgo_set_basis(self.h, _dp(t["S"]), _dp(t["G"]), _dp(t["f0"]), _dp(t["f1"]), _dp(t["e0"]), _dp(t["e1"]))

    def __del__(self):
        if getattr(self, "h", None):
            lib().dgo_destroy(self.h)
            self.h = None

    def n_dofs(self):
        return lib().dgo_n_dofs(self.h)

    def penalty(self):
        s = np.zeros(3)
        lib().dgo_penalty(self.h, _dp(s))
        return s

    def vmult(self, src):
        src = np.ascontiguousarray(src, dtype=np.float64)
        dst = np.zeros_like(src)
        lib().dgo_vmult(self.h, _dp(src), _dp(dst))
        return dst

    def diagonal(self):
        """exact diagonal by probing with unit vectors, 8 colours of cells (cells two apart do not couple)"""
        n3 = (self.degree + 1) ** 3
        nc = self.ncell
        cells = np.arange(nc[0] * nc[1] * nc[2])
        cx, cy, cz = cells % nc[0], (cells // nc[0]) % nc[1], cells // (nc[0] * nc[1])
        colour = (cx % 2) + 2 * (cy % 2) + 4 * (cz % 2)
        diag = np.zeros(self.n_dofs())
        for col in range(8):
            sel = cells[colour == col]
            if sel.size == 0:
                continue
            for l in range(n3):
                e = np.zeros(self.n_dofs())
                e[sel * n3 + l] = 1.0
                y = self.vmult(e)
                diag[sel * n3 + l] = y[sel * n3 + l]
        return diag

    def support_points(self):
        """physical coordinates of all DoFs, shape (n_dofs, 3)"""
        b = basis(self.degree)
        g = b["gll"]
        n = self.degree + 1
        nc = self.ncell
        h = [(self.upper[d] - self.lower[d]) / nc[d] for d in range(3)]
        cz, cy, cx, k, j, i = np.meshgrid(np.arange(nc[2]), np.arange(nc[1]), np.arange(nc[0]), np.arange(n),
                                          np.arange(n), np.arange(n), indexing="ij")
        X = np.stack([self.lower[0] + (cx + g[i]) * h[0], self.lower[1] + (cy + g[j]) * h[1],
                      self.lower[2] + (cz + g[k]) * h[2]], axis=-1).reshape(-1, 3)
        return X @ self.trafo.T


class JacobiTransformed:
    """multigrid::JacobiTransformed (laplace_operator_dg.h:2028-2256) with LocalBasisTransformer type 0 (:176-222)
    restated with numpy/scipy: V = generalised eigenvectors of the 1-D SIP Laplacian w.r.t. the 1-D mass matrix
    (LAPACK through scipy.linalg.eigh, as the reference's compute_generalized_eigenvalues_symmetric), and per cell
    D = diag(V^T A_KK V) obtained by probing the oracle's face-based operator (8 colours of cells)."""

    def __init__(self, op):
        import scipy.linalg
        self.op = op
        p = op.degree
        n = p + 1
        b = op.tables
        S, G, w = b["S"], b["G"], b["wq"]
        mass = S.T @ np.diag(w) @ S
        lapl = G.T @ np.diag(w) @ G
        pen = float(n * n)  # penalty_factor = 1
        v0, v1, g0, g1 = b["f0"], b["f1"], b["e0"], b["e1"]
        lapl = lapl + pen * np.outer(v0, v0) + 0.5 * (np.outer(g0, v0) + np.outer(v0, g0))
        lapl = lapl + pen * np.outer(v1, v1) - 0.5 * (np.outer(g1, v1) + np.outer(v1, g1))
        self.eigenvalues, vecs = scipy.linalg.eigh(lapl, mass)
        self.V = vecs.T.copy()  # V[k, j] = component j of eigenvector k
        # diagonal of V^T A_KK V per cell by probing
        nc = op.ncell
        n3 = n ** 3
        cells = np.arange(nc[0] * nc[1] * nc[2])
        cx, cy, cz = cells % nc[0], (cells // nc[0]) % nc[1], cells // (nc[0] * nc[1])
        colour = (cx % 2) + 2 * (cy % 2) + 4 * (cz % 2)
        self.diag = np.zeros((cells.size, n3))
        V = self.V
        for k in range(n3):
            k0, k1, k2 = k % n, (k // n) % n, k // (n * n)
            vk = np.einsum("c,b,a->cba", V[k2], V[k1], V[k0]).ravel()
            for col in range(8):
                sel = cells[colour == col]
                if sel.size == 0:
                    continue
                u = np.zeros(op.n_dofs())
                u.reshape(-1, n3)[sel] = vk
                y = op.vmult(u).reshape(-1, n3)
                self.diag[sel, k] = y[sel] @ vk

    def vmult(self, src):
        n = self.op.degree + 1
        V = self.V
        r = np.asarray(src, dtype=np.float64).reshape(-1, n, n, n)      # [cell, z, y, x]
        c = np.einsum("kz,jy,ix,czyx->ckji", V, V, V, r)                 # V^T in every direction
        c = c / self.diag.reshape(-1, n, n, n)
        return np.einsum("kz,jy,ix,ckji->czyx", V, V, V, c).ravel()     # back to the nodal basis


class ChebyshevDG:
    """deal.II PreconditionChebyshev as multigrid_solver_dg.h:294-304 sets it up over the DG operator: eigenvalue
    estimate by eig_cg_n_iterations of CG on P^-1 A from the vector "index mod 11, mean removed", max estimate x 1.2,
    then the three-term recurrence (vmult: zero start, step: given start)."""

    def __init__(self, op, prec, smoothing_range=20., degree=3, eig_cg_n_iterations=15):
        """degree None: the smoother as coarse solver, smoothing_range is then the tolerance and the degree follows
        from it (multigrid_solver_dg_plain.h:204-208)"""
        self.op, self.prec, self.degree = op, prec, degree
        N = op.n_dofs()
        g = (np.arange(N) % 11).astype(np.float64)
        g = -(g - g.mean())
        diag, off = [], []
        h = prec(g)
        d = -h
        gh = g @ h
        alpha_old = beta_old = 0.0
        for it in range(1, eig_cg_n_iterations + 1):
            h = op.vmult(d)
            alpha = gh / (d @ h)
            g = g + alpha * h
            if it == 1:
                diag.append(1. / alpha)
            else:
                diag.append(1. / alpha + beta_old / alpha_old)
                off.append(np.sqrt(beta_old) / alpha_old)
            self.cg_iterations = it
            if np.linalg.norm(g) <= 1e-10:
                break
            h = prec(g)
            gh_new = g @ h
            beta = gh_new / gh
            gh = gh_new
            d = beta * d - h
            alpha_old, beta_old = alpha, beta
        T = np.diag(diag) + np.diag(off, 1) + np.diag(off, -1)
        ev = np.linalg.eigvalsh(T)
        self.min_eig, self.max_eig = ev[0], ev[-1]
        max_est = 1.2 * self.max_eig
        a = max_est / smoothing_range if smoothing_range > 1. else min(0.9 * max_est, self.min_eig)
        if degree is None:
            rng = max_est / a
            sigma = (1. - np.sqrt(1. / rng)) / (1. + np.sqrt(1. / rng))
            eps = smoothing_range
            self.degree = 1 + int(np.log(1. / eps + np.sqrt(1. / eps / eps - 1.)) / np.log(1. / sigma))
        self.theta, self.delta = 0.5 * (max_est + a), 0.5 * (max_est - a)

    def _loop(self, x, x_old, b):
        rhok, sigma = self.delta / self.theta, self.theta / self.delta
        for _ in range(self.degree - 1):
            rhokp = 1. / (2. * sigma - rhok)
            f1, f2 = rhokp * rhok, 2. * rhokp / self.delta
            rhok = rhokp
            x, x_old = x + f1 * (x - x_old) + f2 * self.prec(b - self.op.vmult(x)), x
        return x

    def vmult(self, b):
        return self._loop(self.prec(b) / self.theta, np.zeros_like(b), b)

    def step(self, x, b):
        return self._loop(x + self.prec(b - self.op.vmult(x)) / self.theta, x, b)


class MultigridSolverDG:
    """multigrid::MultigridSolverDG (common/multigrid_solver_dg.h:44-735) composed from the oracle's parts: the
    face-based DG operator (dg_oracle.cpp), JacobiTransformed / ChebyshevDG above, and the FE_Q hierarchy of
    mg_oracle.cpp with the smoother settings of :272-293.  Double precision in the DG part (the reference runs
    the cycle in float); small meshes only."""

    def __init__(self, degree, n_subdivisions, n_refine, lower=(-1.,) * 3, upper=(1.,) * 3, degree_pre=3,
                 degree_post=3, preconditioner=1, basis=DG_BASIS_GAUSS_LOBATTO):
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        assert degree_pre == degree_post
        self.fe = MultigridSolver(degree, n_subdivisions, n_refine, lower, upper, degree_pre, degree_post, 1)
        L = self.fe.n_levels - 1
        self.fe.smoother_reinit(0, 2e-3, -1, -1)
        if L > 0:
            self.fe.smoother_reinit(L, 20., max(1, degree_pre - 1), 15)
        nc = [k << n_refine for k in n_subdivisions]
        self.op = DGOperator(degree, nc, lower=lower, upper=upper, trafo=np.eye(3), basis=basis)
        # LocalBasisTransformer<dim,1,...> (laplace_operator_dg.h:106-141): DG coefficients from the values in the
        # Gauss-Lobatto points of FE_Q (the identity for FE_DGQ)
        self.Binv = np.linalg.inv(self.op.tables["B"])
        if preconditioner == 1:
            self.jt = JacobiTransformed(self.op)
            self.prec = self.jt.vmult
        else:
            inv = 1. / self.op.diagonal()
            self.prec = lambda v: inv * v
        self.smooth = ChebyshevDG(self.op, self.prec, 20., degree_pre, 15)
        p, n = degree, degree + 1
        nd = self.fe.dof_grid(L)
        cz, cy, cx, k, j, i = np.meshgrid(np.arange(nc[2]), np.arange(nc[1]), np.arange(nc[0]), np.arange(n),
                                          np.arange(n), np.arange(n), indexing="ij")
        self.lex = (((cz * p + k) * nd[1] + cy * p + j) * nd[0] + cx * p + i).ravel()
        mask = np.ones((nd[2], nd[1], nd[0]))
        mask[0], mask[-1], mask[:, 0], mask[:, -1], mask[:, :, 0], mask[:, :, -1] = (0,) * 6
        self.mask = mask.ravel()
        # rhs_i = (f, phi_i) with QGauss(p+1)   (multigrid_solver_dg.h:243-265)
        b = self.op.tables
        S, xq, wq = b["S"], b["xq"], b["wq"]
        h = [(upper[d] - lower[d]) / nc[d] for d in range(3)]
        self.h, self.nc, self.lower = h, nc, lower
        X = lower[0] + h[0] * (cx + xq[i])
        Y = lower[1] + h[1] * (cy + xq[j])
        Z = lower[2] + h[2] * (cz + xq[k])
        self.Xq = (X, Y, Z)
        f = 3 * (3 * np.pi) ** 2 * np.sin(3 * np.pi * X) * np.sin(3 * np.pi * Y) * np.sin(3 * np.pi * Z)
        self.W = wq[k] * wq[j] * wq[i] * h[0] * h[1] * h[2]
        self.S = S
        self.rhs = np.einsum("rc,qb,pa,zyxrqp->zyxcba", S, S, S, f * self.W).ravel()
        self.solution = np.zeros_like(self.rhs)

    def n_dofs(self):
        return self.op.n_dofs()

    def vmult(self, src):
        """dg_v_cycle(1) (multigrid_solver_dg.h:605-633)"""
        x = self.smooth.vmult(src)
        res = src - self.op.vmult(x)
        d = np.zeros(self.mask.size)
        np.add.at(d, self.lex, self._basis_change(res, self.Binv.T))       # action 1 epilogue, :1798-1819
        u = self.fe.vcycle(d * self.mask)
        x = x + self._basis_change((u * self.mask)[self.lex], self.Binv)    # prolongate_add_cg_to_dg, :1863-1894
        return self.smooth.step(x, src)

    def _basis_change(self, v, M):
        n = self.op.degree + 1
        r = np.asarray(v).reshape(-1, n, n, n)
        return np.einsum("kz,jy,ix,czyx->ckji", M, M, M, r).ravel()

    def solve_cg(self, tolerance=1e-3):
        """SolverCG with ReductionControl(100, 1e-16, tolerance) (:410-424) -> (estimate, rate, residual history)"""
        x = np.zeros_like(self.rhs)
        g = -self.rhs
        res0 = np.linalg.norm(g)
        hist = [res0]
        it, gh, d = 0, 0., None
        while not (hist[-1] <= 1e-16 or hist[-1] <= tolerance * res0):
            it += 1
            h = self.vmult(g)
            gh_new = g @ h
            d = -h if it == 1 else (gh_new / gh) * d - h
            gh = gh_new
            h = self.op.vmult(d)
            alpha = gh / (d @ h)
            x += alpha * d
            g = g + alpha * h
            hist.append(np.linalg.norm(g))
            assert it < 100
        self.solution = x
        self.last_step = it
        rate = (hist[-1] / res0) ** (1. / it)
        return np.log(tolerance) / np.log(rate), rate, np.array(hist)

    def compute_l2_error(self):
        """:328-366"""
        n = self.op.degree + 1
        u = self.solution.reshape(self.nc[2], self.nc[1], self.nc[0], n, n, n)
        uq = np.einsum("rc,qb,pa,zyxcba->zyxrqp", self.S, self.S, self.S, u)
        X, Y, Z = self.Xq
        ex = np.sin(3 * np.pi * X) * np.sin(3 * np.pi * Y) * np.sin(3 * np.pi * Z)
        return np.sqrt(np.sum((uq - ex) ** 2 * self.W) / np.sum(self.W))


class MultigridSolverDGPlain:
    """multigrid::MultigridSolverDGPlain (common/multigrid_solver_dg_plain.h:54-590) composed from the oracle's parts:
    the face-based DG operator (dg_oracle.cpp) on every level, JacobiTransformed / ChebyshevDG with the settings of
    :192-213, the cell-wise embedding as level transfer, the V-cycle of :456-497 and SolverCG (:304-316).  Double
    precision throughout (the reference runs the cycle in float); small meshes only."""

    def __init__(self, degree, n_subdivisions, n_refine, lower=(-1.,) * 3, upper=(1.,) * 3, degree_pre=3, degree_post=3,
                 basis=DG_BASIS_GAUSS_LOBATTO):
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        assert degree_pre == degree_post
        self.degree, self.n_levels = degree, n_refine + 1
        self.ops, self.jt, self.smooth = [], [], []
        for l in range(self.n_levels):
            nc = [k << l for k in n_subdivisions]
            op = DGOperator(degree, nc, lower=lower, upper=upper, trafo=np.eye(3), basis=basis)
            jt = JacobiTransformed(op)
            if l == 0:
                sm = ChebyshevDG(op, jt.vmult, 1e-5, None, op.n_dofs())
            else:
                sm = ChebyshevDG(op, jt.vmult, 20., degree_pre if l < n_refine else max(1, degree_pre - 1), 15)
            self.ops.append(op)
            self.jt.append(jt)
            self.smooth.append(sm)
        self.op = self.ops[-1]
        b = self.op.tables
        self.E = (b["E0"], b["E1"])
        n = degree + 1
        nc = self.op.ncell
        cz, cy, cx, k, j, i = np.meshgrid(np.arange(nc[2]), np.arange(nc[1]), np.arange(nc[0]), np.arange(n),
                                          np.arange(n), np.arange(n), indexing="ij")
        S, xq, wq = b["S"], b["xq"], b["wq"]
        h = [(upper[d] - lower[d]) / nc[d] for d in range(3)]
        self.nc = nc
        X = lower[0] + h[0] * (cx + xq[i])
        Y = lower[1] + h[1] * (cy + xq[j])
        Z = lower[2] + h[2] * (cz + xq[k])
        self.Xq = (X, Y, Z)
        f = 3 * (3 * np.pi) ** 2 * np.sin(3 * np.pi * X) * np.sin(3 * np.pi * Y) * np.sin(3 * np.pi * Z)
        self.W = wq[k] * wq[j] * wq[i] * h[0] * h[1] * h[2]
        self.S = S
        self.rhs = np.einsum("rc,qb,pa,zyxrqp->zyxcba", S, S, S, f * self.W).ravel()   # :162-185
        self.solution = np.zeros_like(self.rhs)

    def n_dofs(self):
        return self.op.n_dofs()

    def prolongate(self, to_level, src):
        """coarse cell -> its eight children, (E_oz x E_oy x E_ox) per child"""
        n = self.degree + 1
        ncx, ncy, ncz = self.ops[to_level - 1].ncell
        u = np.asarray(src, dtype=np.float64).reshape(ncz, ncy, ncx, n, n, n)
        out = np.zeros((2 * ncz, 2 * ncy, 2 * ncx, n, n, n))
        for oz in (0, 1):
            for oy in (0, 1):
                for ox in (0, 1):
                    out[oz::2, oy::2, ox::2] = np.einsum("kc,jb,ia,zyxcba->zyxkji", self.E[oz], self.E[oy], self.E[ox], u)
        return out.ravel()

    def restrict(self, from_level, src):
        """the transpose of prolongate()"""
        n = self.degree + 1
        ncx, ncy, ncz = self.ops[from_level - 1].ncell
        r = np.asarray(src, dtype=np.float64).reshape(2 * ncz, 2 * ncy, 2 * ncx, n, n, n)
        out = np.zeros((ncz, ncy, ncx, n, n, n))
        for oz in (0, 1):
            for oy in (0, 1):
                for ox in (0, 1):
                    out += np.einsum("kc,jb,ia,zyxkji->zyxcba", self.E[oz], self.E[oy], self.E[ox], r[oz::2, oy::2, ox::2])
        return out.ravel()

    def v_cycle(self, level, defect):
        """:456-497 with my_n_cycles = 1"""
        if level == 0:
            return self.smooth[0].vmult(defect)
        x = self.smooth[level].vmult(defect)
        t = defect - self.ops[level].vmult(x)
        x = x + self.prolongate(level, self.v_cycle(level - 1, self.restrict(level, t)))
        return self.smooth[level].step(x, defect)

    def vmult(self, src):
        return self.v_cycle(self.n_levels - 1, src)

    solve_cg = MultigridSolverDG.solve_cg
    compute_l2_error = MultigridSolverDG.compute_l2_error
