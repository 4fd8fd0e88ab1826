"""Generates tests/golden/vectors.npz: seeded inputs are NOT stored (they are regenerated from the seed), the outputs of the
CPU oracle are.  Run once from the repository root (python tests/golden/make_golden.py); the oracle tests check that today's
oracle still reproduces the file, the GPU tests compare the library with it directly.

  fe_q_p{p}   : y = A x, FE_Q(p) on the cube [-0.9,1]^3, 2 x 2 x 2 cells (p <= 4: 4 x 4 x 4), lexicographic, Dirichlet rows identity
  dg_p{p}     : y = A x, SIP-DG FE_DGQ(p) on the sheared mesh of matvec_dg, 2 x 2 x 2 cells, cell-wise
  adaptive_*  : y = A x on the hanging-node meshes of tests/adaptive_cases.py, in the order of the sorted DoF coordinates
  dg_solver_* : MultigridSolverDG (FE_DGQ(2), 4 x 4 x 4 cells on [-1,1]^3): residual history of solve_cg(1e-8), [L2 error],
                and one preconditioner cycle applied to a seeded vector
  dg_hermite_p{p}, dg_gauss_p{p} : as dg_p{p} on FE_DGQHermite(p) / FE_DGQArbitraryNodes(QGauss(p+1)) (coefficient vectors of
                that basis in and out)
  dg_hermite_solver_* : as dg_solver_* on FE_DGQHermite(2), the basis of poisson_dg (poisson_dg/program.cc:174)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
SEED = 20260
OUT = os.path.join(ROOT, "tests", "golden", "vectors.npz")


def fe_q_case(p):
    return (1, 2) if p <= 4 else (1, 1)      # n_subdivisions, n_refinements


def seeded(n, salt):
    return np.random.default_rng(SEED + salt).standard_normal(n)


def sorted_order(points):
    """deterministic order of DoFs that does not depend on a numbering: sort by (z, y, x) rounded to 1e-9"""
    k = np.round(points / 1e-9).astype(np.int64)
    return np.lexsort((k[:, 0], k[:, 1], k[:, 2]))


def generate():
    from oracle import orc, adaptive_oracle
    from adaptive_cases import CASES
    out = {}
    for p in range(1, 9):
        nsub, nref = fe_q_case(p)
        o = orc.MultigridSolver(p, nsub, nref)
        out["fe_q_p%d" % p] = o.vmult(seeded(o.n_dofs(), p))
        d = orc.DGOperator(p, (2, 2, 2))
        out["dg_p%d" % p] = d.vmult(seeded(d.n_dofs(), 100 + p))
        for name, kind in (("hermite", orc.DG_BASIS_HERMITE), ("gauss", orc.DG_BASIS_GAUSS)):
            d = orc.DGOperator(p, (2, 2, 2), basis=kind)
            out["dg_%s_p%d" % (name, p)] = d.vmult(seeded(d.n_dofs(), 100 + p))
    import multigrid_b200 as mgb          # host side only: mesh and cell list (no GPU needed)
    from adaptive_cases import make
    for i, name in enumerate(sorted(CASES)):
        op = make(mgb, name)
        lo, sz = op.cell_boxes()
        o = adaptive_oracle.AdaptiveOracle(op.degree, lo, sz)
        order = sorted_order(o.points)
        x = np.zeros(o.n_dofs())
        x[order] = seeded(o.n_dofs(), 200 + i)
        out["adaptive_" + name] = o.vmult(x)[order]
    s = orc.MultigridSolverDG(*DG_SOLVER_CASE)
    out["dg_solver_cycle"] = s.vmult(seeded(s.n_dofs(), 300))
    _, _, hist = s.solve_cg(1e-8)
    out["dg_solver_history"] = np.asarray(hist)
    out["dg_solver_l2_error"] = np.array([s.compute_l2_error()])
    out["dg_solver_solution"] = s.solution.copy()
    s = orc.MultigridSolverDG(*DG_SOLVER_CASE, basis=orc.DG_BASIS_HERMITE)
    out["dg_hermite_solver_cycle"] = s.vmult(seeded(s.n_dofs(), 300))
    _, _, hist = s.solve_cg(1e-8)
    out["dg_hermite_solver_history"] = np.asarray(hist)
    out["dg_hermite_solver_l2_error"] = np.array([s.compute_l2_error()])
    out["dg_hermite_solver_solution"] = s.solution.copy()
    return out


DG_SOLVER_CASE = (2, (1, 1, 1), 2)      # degree, n_subdivisions, n_refinements


if __name__ == "__main__":
    res = generate()
    np.savez_compressed(OUT, **res)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(res), "vectors")
