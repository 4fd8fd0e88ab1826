"""poisson_dg protocol (poisson_dg/program.cc:227-263): MultigridSolverDG on the doubling mesh, minimum of 4 CG solves
to the given tolerance; DoF/s = DG DoFs / solve time.
usage: python tools/time_dg_solver.py [cycle=18] [degree=4] [tolerance=1e-3] [basis=0: FE_DGQ, 1: FE_DGQHermite, 2: Gauss]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb


def doubling_mesh(cycle):
    """poisson_dg/program.cc:357-376"""
    rem = cycle % 3
    nsub = tuple(2 if d < rem else 1 for d in range(3))
    upper = tuple(3. if d < rem else 1. for d in range(3))
    return nsub, cycle // 3, (-1.,) * 3, upper


def run(cycle=18, degree=4, tolerance=1e-3, n_solves=4, preconditioner=1, basis=0):
    nsub, nref, lower, upper = doubling_mesh(cycle)
    t0 = time.perf_counter()
    s = mgb.MultigridSolverDG(degree, nsub, nref, lower, upper, 3, 3, preconditioner, basis=basis)
    s.sync()
    setup = time.perf_counter() - t0
    n = s.n_dofs()
    best = 1e30
    l0 = s.launch_count()
    for _ in range(n_solves):
        s.sync()
        t0 = time.perf_counter()
        est, rate = s.solve_cg(tolerance)
        s.sync()
        best = min(best, time.perf_counter() - t0)
    launches = (s.launch_count() - l0) // n_solves
    host = torch.empty(n, dtype=torch.float64).pin_memory()
    e2e = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        s.solve_cg_host(tolerance, host)
        e2e = min(e2e, time.perf_counter() - t0)
    out = dict(n_dofs=n, cells=[k << nref for k in nsub], degree=degree, tolerance=tolerance, cg_iterations=s.last_step,
               reduction_rate=rate, iterations_estimate=est, solve_ms=best * 1e3, dofs_per_s=n / best,
               e2e_ms=e2e * 1e3, e2e_dofs_per_s=n / e2e, d2h_bytes=n * 8, l2_error=s.compute_l2_error(),
               setup_s=setup, launches_per_solve=launches, smoother=s.smoother_info())
    s.close()
    return out


if __name__ == "__main__":
    cyc = int(sys.argv[1]) if len(sys.argv) > 1 else 18
    deg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    tol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-3
    basis = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    print('basis', basis)
    for k, v in run(cyc, deg, tol, basis=basis).items():
        print(k, v)
