"""Adaptive meshes with hanging nodes, CPU side: the mesh logic of the library (host code, no GPU), the
oracle's properties, and the library's constraint tables against the oracle's geometric restatement."""
import numpy as np
import pytest

from adaptive_cases import CASES, make


@pytest.fixture(scope="module")
def ado():
    from oracle import adaptive_oracle
    return adaptive_oracle


@pytest.fixture(scope="module", params=sorted(CASES))
def case(request, mgb, ado):
    op = make(mgb, request.param)
    lo, sz = op.cell_boxes()
    return request.param, op, ado.AdaptiveOracle(op.degree, lo, sz)


def test_mesh_is_balanced_and_fills_the_domain(case):
    name, op, _ = case
    lo, sz = op.cell_boxes()
    level, _ = op.cells()
    vol = np.prod(sz, axis=1).sum()
    p, nr, mask, ng, rounds = CASES[name]
    n_roots = nr[0] * nr[1] * nr[2] if mask is None else int(mask.sum())
    root_vol = np.prod((np.array(op.upper) - np.array(op.lower)) / np.array(nr))
    assert abs(vol - n_roots * root_vol) < 1e-12 * vol
    assert level.max() > level.min()                      # the case has hanging nodes
    tol = 1e-9
    for c in range(lo.shape[0]):                          # 2:1 over faces, edges and corners
        touch = np.all((lo <= lo[c] + sz[c] + tol) & (lo + sz >= lo[c] - tol), axis=1)
        assert np.abs(level[touch] - level[c]).max() <= 1


def test_oracle_properties(case):
    name, op, o = case
    K = o.K_free
    assert abs(K - K.T).max() <= 1e-12 * abs(K).max()
    assert np.abs(K @ np.ones(o.n_dofs())).max() <= 1e-11 * abs(K).max()     # constants are in the conforming space
    p = op.degree
    # the conforming space contains Q_p: energy of the interpolant of a polynomial is the exact integral
    if p == 1:
        u = o.interpolate(lambda x, y, z: x * y + 2 * z)
        grad2 = lambda x, y, z: y ** 2 + x ** 2 + 4
    else:
        u = o.interpolate(lambda x, y, z: x * x * y - z * y)
        grad2 = lambda x, y, z: (2 * x * y) ** 2 + (x * x - z) ** 2 + y ** 2
    lo, sz = op.cell_boxes()
    gx, gw = np.polynomial.legendre.leggauss(4)
    gx, gw = 0.5 * (gx + 1), 0.5 * gw
    exact = 0.0
    for c in range(lo.shape[0]):
        X, Y, Z = np.meshgrid(lo[c, 0] + sz[c, 0] * gx, lo[c, 1] + sz[c, 1] * gx, lo[c, 2] + sz[c, 2] * gx, indexing="ij")
        W = np.einsum("i,j,k->ijk", gw, gw, gw) * np.prod(sz[c])
        exact += np.sum(W * grad2(X, Y, Z))
    assert abs(u @ (K @ u) - exact) <= 1e-11 * exact
    assert (o.K.diagonal() > 0).all()


def test_library_tables_match_oracle(case):
    name, op, o = case
    assert op.m() == o.n_dofs()
    perm = o.permutation_to(op.dof_points())
    assert np.array_equal(op.dirichlet_mask(), o.dirichlet[perm])
    ci, rs, rd, rw = op.tables()
    info = op.info()
    assert info["n_dirichlet"] == int(o.dirichlet.sum()) and info["hanging_local_nodes"] > 0
    n3 = (op.degree + 1) ** 3
    for c in range(op.n_cells()):
        L = np.zeros((n3, op.m()))
        for l in range(n3):
            e = ci[c, l]
            if e >= 0:
                L[l, e] = 1.0
            elif e <= -2:
                r = -2 - e
                L[l, rd[rs[r]:rs[r + 1]]] = rw[rs[r]:rs[r + 1]]
        assert np.abs(L - o.cell_rows(c)[:, perm]).max() <= 1e-12, (name, c)
    # (with the balance over edges and corners a master is never constrained itself: a refined neighbour of a coarse
    #  face would touch whatever touches that face; rows therefore reach over one coarse face or edge at most)
    assert max(rs[1:] - rs[:-1]) <= (op.degree + 1) ** 2


def test_refinement_api_errors(mgb):
    with pytest.raises(mgb.MgbError):
        mgb.LaplaceOperatorAdaptive(9)
    with pytest.raises(mgb.MgbError):
        mgb.LaplaceOperatorAdaptive(2, (1, 1, 1), np.zeros((1, 1, 1), dtype=np.uint8))
    op = mgb.LaplaceOperatorAdaptive(2, (1, 1, 1), None, n_global_refinements=1)
    assert op.n_cells() == 8 and op.m() == 125 and op.info()["hanging_local_nodes"] == 0
    # node indices must fit the key: 64 roots x 2^8 cells x degree 8 does not
    with pytest.raises(mgb.MgbError):
        deep = mgb.LaplaceOperatorAdaptive(8, (64, 1, 1), None, (-1.,) * 3, (1.,) * 3, 0)
        for _ in range(13):
            lo, sz = deep.cell_boxes()
            deep.refine(np.linalg.norm(lo, axis=1) < 1e-9 + np.linalg.norm(lo, axis=1).min())


@pytest.mark.parametrize("name", ["hyper_l_p2", "extruded_l_p1", "box_point_p2"])
def test_local_smoothing_oracle_properties(mgb, ado, name):
    """the local-smoothing multigrid restatement (poisson_l/program.cc:283-420): every unconstrained active DoF is copied
    from / to exactly one level, the preconditioner is symmetric positive, the edge matrices are transposes of each other
    on the unconstrained / refinement-edge blocks, CG converges at a multigrid rate to the solution of the active system"""
    op = make(mgb, name)
    lo, sz = op.cell_boxes()
    root = (np.array(op.upper) - np.array(op.lower)) / np.array(op.n_roots)
    o = ado.LocalSmoothingOracle(op.degree, lo, sz, root, op.lower, n_smoother=3)
    act = o.active
    free = ~act.dirichlet
    copied = np.concatenate([L["copy_active"] for L in o.levels])
    assert np.array_equal(np.sort(copied), np.where(free)[0])
    assert sum(int(L["edge"].sum()) for L in o.levels) > 0          # the case has refinement edges
    rng = np.random.default_rng(2)
    u, v = (np.where(free, rng.standard_normal(free.size), 0.0) for _ in range(2))
    Mu, Mv = o.vmult(u), o.vmult(v)
    assert abs(v @ Mu - u @ Mv) <= 1e-10 * abs(v @ Mu) and u @ Mu > 0
    for l in range(1, o.n_levels):
        L = o.levels[l]
        a, b = rng.standard_normal(L["n"]), rng.standard_normal(L["n"])
        assert abs(b @ o.interface_down(l, a) - a @ o.interface_up(l, b)) <= 1e-10 * (np.abs(L["K"]).sum() ** 0.5 + 1)
    x, its, hist = o.solve_cg(u)
    assert its <= 14 and (hist[-1] / hist[0]) ** (1.0 / its) < 0.25
    assert np.linalg.norm(act.K @ x - u) <= 1e-8 * np.linalg.norm(u)
