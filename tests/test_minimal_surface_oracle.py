"""The oracle's minimal surface solver (oracle/mg_oracle.cpp ms_*, restating minimal_surface/program.cc:102-198,412-572 in
3-D on the structured meshes of this library).  The reference program is 2-D on a disk and publishes no numbers, so the
restatement is pinned by mathematics: Scherk's surface u = log(cos x / cos y) solves the minimal surface equation (also
as a function of three variables), so with its boundary values the discrete solution converges to it at order p + 1;
Newton converges quadratically; affine data are reproduced in one step; the residual is the derivative of the area."""
import numpy as np
import pytest

scherk = lambda x, y, z: np.log(np.cos(x) / np.cos(y)) + 0 * z


def make(orc, p, nref, **kw):
    s = orc.MultigridSolver(p, 1, nref, lower=(-1,) * 3, upper=(1,) * 3, degree_pre=2, degree_post=2, n_cycles=1, **kw)
    return s


@pytest.mark.parametrize("p", [2, 3])
def test_converges_to_scherk_surface(orc, p):
    errs = []
    for nref in (1, 2, 3):
        s = make(orc, p, nref)
        s.minimal_surface_set_boundary_values(scherk)
        steps = s.minimal_surface_solve(20, 1e-11)
        assert len(steps) <= 6 and steps[-1]["final_residual"] < 1e-11
        assert all(st["line_search_steps"] == 0 for st in steps)
        x = s.support_points()
        errs.append(np.abs(s.get_vector(orc.VEC_SOLUTION) - scherk(x[:, 0], x[:, 1], x[:, 2])).max())
    assert errs[1] / errs[2] > 2 ** p      # order p + 1 asymptotically; at least p on these coarse meshes
    assert errs[2] < (4e-4 if p == 2 else 3e-5)


def test_newton_is_quadratic_and_first_step_is_laplace(orc):
    s = make(orc, 3, 2)
    s.minimal_surface_set_boundary_values(scherk)
    steps = s.minimal_surface_solve(20, 1e-12)
    r = [st["final_residual"] for st in steps]
    assert r[2] < 0.05 * r[1] ** 1.5 and r[3] < r[2] ** 1.5     # super-linear once in the basin
    # the first step solves the Laplace equation with the same data: its initial residual is the Laplace residual
    s2 = make(orc, 3, 2)
    s2.minimal_surface_set_boundary_values(scherk)
    u0 = s2.get_vector(orc.VEC_SOLUTION)
    assert abs(np.linalg.norm(s2.minimal_surface_residual(u0, True)) - steps[0]["initial_residual"]) < 1e-12


def test_affine_data_are_a_fixed_point(orc):
    g = lambda x, y, z: 0.3 * x - 0.7 * y + 0.2 * z + 0.1
    s = make(orc, 2, 2)
    s.minimal_surface_set_boundary_values(g)
    steps = s.minimal_surface_solve(5, 1e-12)
    x = s.support_points()
    assert np.abs(s.get_vector(orc.VEC_SOLUTION) - g(x[:, 0], x[:, 1], x[:, 2])).max() < 1e-6
    assert steps[-1]["final_residual"] < 1e-12


def test_residual_is_the_gradient_of_the_area(orc):
    """r(u) = -dA/du for A(u) = int sqrt(1 + |grad u|^2): checked by central differences of a quadrature of A built from
    the residual itself along a direction (d/dt A(u + t v) = -r(u + t v) . v integrated by Simpson's rule)"""
    s = make(orc, 2, 1)
    s.minimal_surface_set_boundary_values(scherk)
    rng = np.random.default_rng(0)
    N = s.n_dofs()
    u = s.get_vector(orc.VEC_SOLUTION) + 0.1 * rng.standard_normal(N)
    v, w = rng.standard_normal(N), rng.standard_normal(N)
    x = s.support_points()
    inner = (np.abs(x) < 1 - 1e-12).all(axis=1)
    v[~inner] = 0
    w[~inner] = 0
    # symmetry of the Jacobian of r (r is a gradient field): (r(u + e v) - r(u - e v)) . w = (r(u + e w) - r(u - e w)) . v
    e = 1e-5
    a = (s.minimal_surface_residual(u + e * v) - s.minimal_surface_residual(u - e * v)) @ w
    b = (s.minimal_surface_residual(u + e * w) - s.minimal_surface_residual(u - e * w)) @ v
    assert abs(a - b) < 1e-6 * abs(a)


def test_line_search_engages_for_steep_data(orc):
    g = lambda x, y, z: 10 * np.sin(2 * np.pi * (x + y)) + 0 * z     # the reference's data (program.cc:93), steeper
    s = make(orc, 2, 2)
    s.minimal_surface_set_boundary_values(g)
    steps = s.minimal_surface_solve(40, 1e-9)
    assert any(st["line_search_steps"] > 0 for st in steps)
    assert all(st["final_residual"] < st["initial_residual"] for st in steps[1:])
    assert steps[-1]["final_residual"] < 1e-9


def test_level_solutions_interpolate(orc):
    """coefficient_solutions (:417-456): a polynomial of degree p is carried to the coarse levels exactly"""
    g = lambda x, y, z: x * y - 0.5 * z * z + 0.25 * x
    s = make(orc, 2, 2)
    s.minimal_surface_set_boundary_values(g)
    s.minimal_surface_step(True)
    for level in (0, 1):
        xs = s.support_points(level)
        bnd = ~(np.abs(xs) < 1 - 1e-12).all(axis=1)
        got = s.minimal_surface_level_solution(level)
        # level vectors hold the restriction of the iterate BEFORE the step: g on the boundary, zero inside
        assert np.abs(got[bnd] - g(xs[bnd, 0], xs[bnd, 1], xs[bnd, 2])).max() < 1e-6
