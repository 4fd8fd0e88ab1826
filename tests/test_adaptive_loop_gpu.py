"""GPU parity of the adaptive loop's parts (csrc/adaptive_estimate.cuh through the C ABI, mgb_adaptive_*): boundary
values, right-hand side with inhomogeneous Dirichlet data, Kelly error estimator, error norms -- against the geometric
oracle (oracle/adaptive_oracle.py AdaptiveLoopOracle) on the meshes of tests/adaptive_cases.py; and the loop of
poisson_l/program.cc:497-602 cycle by cycle against the oracle's loop (same meshes, errors, indicators)."""
import numpy as np
import pytest

from adaptive_cases import CASES, L_EXTRUDED, make

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("no GPU")
    return t


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("name", sorted(CASES))
def test_parts_match_oracle(mgb, torch, name):
    from oracle import adaptive_oracle as ao
    op = make(mgb, name)
    lo, sz = op.cell_boxes()
    o = ao.AdaptiveLoopOracle(op.degree, lo, sz)
    perm = o.permutation_to(op.dof_points())
    N = op.m()
    # boundary values of the L-shape singularity
    u = torch.zeros(N, dtype=torch.float64, device="cuda")
    op.interpolate_boundary_values(u, op.FUNCTION_L_SINGULARITY)
    ub = o.boundary_values(ao.l_singularity)
    assert np.abs(u.cpu().numpy() - ub[perm]).max() <= 1e-14
    # a vector with boundary data and a rough interior
    rng = np.random.default_rng(2)
    v = ub[perm] + np.where(op.dirichlet_mask(), 0., 0.3 * rng.standard_normal(N))
    vo = np.zeros(N)
    vo[perm] = v
    dv = torch.tensor(v, device="cuda")
    for f in (0., 1.):
        r = torch.zeros_like(dv)
        op.compute_residual(r, dv, f)
        assert rel(r.cpu().numpy(), o.compute_residual(vo, f)[perm]) <= 1e-12, f
    eta = op.estimate_kelly(dv)
    eta_o = o.kelly(vo)
    assert rel(eta, eta_o) <= 2e-7                      # the indicators are float, as the reference's Vector<float>
    for norm, nq in ((0, op.degree + 2), (1, op.degree + 1), (0, op.degree + 1)):
        got, sq = op.integrate_difference(dv, op.FUNCTION_L_SINGULARITY, norm, nq, per_cell=True)
        want, sq_o = o.integrate_difference(vo, ao.l_singularity, ao.l_singularity_gradient, norm, nq)
        assert abs(got - want) <= 1e-12 * want and rel(sq, sq_o) <= 1e-11, (norm, nq)
    z, _ = op.integrate_difference(dv, op.FUNCTION_ZERO, 0, per_cell=True)
    assert abs(z - o.integrate_difference(vo, lambda x, y, w: 0 * x, None, 0, op.degree + 2)[0]) <= 1e-12 * z


def test_adaptive_loop_matches_oracle(mgb, torch):
    """LaplaceProblem::run on the extruded L (the program's default case) with FE_Q(2): four cycles, every cycle the same
    mesh, error norms and indicators as the oracle's loop with a direct solve; CG iteration counts stay bounded"""
    import scipy.sparse.linalg as sla
    from oracle import adaptive_oracle as ao
    p = 2
    op = mgb.LaplaceOperatorAdaptive(p, (2, 2, 1), L_EXTRUDED, (-1., -1., 0.), (1., 1., 1.), 1)
    table = []
    for cycle in range(4):
        lo, sz = op.cell_boxes()
        o = ao.AdaptiveLoopOracle(p, lo, sz)
        perm = o.permutation_to(op.dof_points())
        stats = op.solve_cycle(op.FUNCTION_L_SINGULARITY, 0., 3)
        table.append(stats)
        ub = o.boundary_values(ao.l_singularity)
        free = ~o.dirichlet
        upd = np.zeros_like(ub)
        upd[free] = sla.spsolve(o.K[free][:, free].tocsc(), o.compute_residual(ub)[free])
        sol_o = ub + upd
        sol = op.solution()
        assert rel(sol, sol_o[perm]) <= 1e-7, cycle
        l2, _ = o.integrate_difference(sol_o, ao.l_singularity, ao.l_singularity_gradient, 0, p + 2)
        h1, _ = o.integrate_difference(sol_o, ao.l_singularity, ao.l_singularity_gradient, 1, p + 1)
        assert abs(stats["l2_error"] - l2) <= 1e-5 * l2 and abs(stats["h1_error"] - h1) <= 1e-6 * h1, (cycle, stats, l2, h1)
        assert 0 < stats["iterations"] <= 14 and stats["rate"] < 0.3
        eta = op.estimate_kelly(torch.tensor(sol, device="cuda"))
        eta_o = o.kelly(sol_o)
        assert rel(eta, eta_o) <= 1e-5
        # the same flags and the same new mesh (indicators that tie within rounding could flip a flag: use the library's)
        lev, pos = op.cells()
        cells = [(int(l), int(q[0]), int(q[1]), int(q[2])) for l, q in zip(lev, pos)]
        ref, coa = ao.fixed_number_flags(eta, 0.3, 0.03)
        want = ao.refine_and_coarsen(cells, ref, coa, (2, 2, 1), L_EXTRUDED)
        op.refine_and_coarsen_fixed_number(eta, 0.3, 0.03)
        lev, pos = op.cells()
        assert {(int(l), int(q[0]), int(q[1]), int(q[2])) for l, q in zip(lev, pos)} == want
    assert all(b["h1_error"] < a["h1_error"] for a, b in zip(table, table[1:]))
    assert table[-1]["h1_error"] < 0.45 * table[0]["h1_error"]


def test_hyper_l_case_and_driver(mgb, torch):
    """the 'h' case of the program (GridGenerator::hyper_L, f = 1, zero boundary values; here with the Dirichlet condition
    on the whole boundary) through the loop driver: the estimator drives refinement towards the re-entrant edges, the
    number of CG iterations stays bounded, and the discrete energy increases monotonically towards the solution's"""
    from adaptive_cases import HYPER_L
    op = mgb.LaplaceOperatorAdaptive(2, (2, 2, 2), HYPER_L, (-1.,) * 3, (1.,) * 3, 1)
    table = op.run_adaptive_loop(4, function=op.FUNCTION_ZERO, f=1., n_smoother=3, top_fraction=0.15, bottom_fraction=0.03)
    assert [t["n_cells"] for t in table] == sorted(t["n_cells"] for t in table) and table[-1]["n_cells"] > 2 * table[0]["n_cells"]
    assert all(0 < t["iterations"] <= 15 for t in table)
    # l2_error against the zero function = |u_h|_L2, h1_error = energy norm: both grow monotonically under refinement
    assert all(b["h1_error"] > a["h1_error"] for a, b in zip(table, table[1:]))
    lev, _ = op.cells()
    lo, sz = op.cell_boxes()
    c = lo + 0.5 * sz
    finest = lev == lev.max()
    # the finest cells sit at the re-entrant edges (two coordinates close to zero)
    d = np.sort(np.abs(c[finest]), axis=1)[:, 1]
    assert d.max() < 0.6
