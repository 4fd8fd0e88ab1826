"""Small run of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

for p, nsub, nref in ((4, (1, 1, 1), 2), (3, (1, 2, 1), 1), (2, (3, 1, 2), 1)):
    s = mgb.MultigridSolver(p, nsub, nref)
    s.solve(True)
    its, _ = s.solve_cg()
    print("FE_Q p=%d its %d l2 %.3e" % (p, its, s.compute_l2_error()))
    s.close()
s = mgb.MultigridSolver(4, 1, 1, problem=mgb.PROBLEM_SHELL_COEF)
print("varcoef its", s.solve_cg()[0])
s.close()
for p in (1, 4, 6):
    for prec in (mgb.F64, mgb.F32):
        g = mgb.LaplaceOperatorDG(p, (2, 1, 2), 1, precision=prec)
        a, b, c = (torch.rand(g.m(), dtype=g.dtype, device="cuda") for _ in range(3))
        g.vmult(b, a)
        g.vmult_residual(c, a, b)
        g.vmult_with_chebyshev_update(c, 2, 0.6, 0.2, a, b)
        g.vmult_with_cg_update(0.1, 0.2, a, b, c, torch.zeros_like(a))
        g.close()
# ---- round 2: mapped geometry, minimal surface, DG solvers, adaptive loop, allocation-free dot
import numpy as np
s = mgb.MultigridSolver(3, 1, 1, problem=mgb.PROBLEM_SHELL_COEF, geometry=mgb.GEOMETRY_SHELL_SECTOR)
print("shell its", s.solve_cg()[0])
s.close()
s = mgb.MultigridSolver(2, 1, 2, lower=(-1.,) * 3, upper=(1.,) * 3, degree_pre=2, degree_post=2, n_cycles=1)
s.minimal_surface_set_boundary_values(lambda x, y, z: np.log(np.cos(x) / np.cos(y)) + 0 * z)
print("minimal surface", [st["linear_iterations"] for st in s.minimal_surface_solve(3, 1e-8)])
s.close()
for p in (2, 4):
    for basis in (0, 1, 2):
        for cls in (mgb.MultigridSolverDG, mgb.MultigridSolverDGPlain):
            d = cls(p, (1, 1, 1), 2, basis=basis)
            d.solve_cg(1e-3)
            print(cls.__name__, p, basis, d.last_step)
            d.close()
g = mgb.LaplaceOperatorDG(3, (2, 1, 1), 1, precision=mgb.F64, basis=mgb.DG_BASIS_GAUSS)
D = g.lumped_mass_inverse()
b = torch.rand(g.m(), dtype=g.dtype, device="cuda")
x = g.initialize_dof_vector()
print("solver_dg", g.solve_cg(x, b, D, 5, True), g.solve_cg(x, b, D, 5, False))
g.close()
mask = np.ones((1, 2, 2), dtype=np.uint8)
mask[0, 1, 1] = 0
for p in (1, 3):
    a = mgb.LaplaceOperatorAdaptive(p, (2, 2, 1), mask, (-1., -1., 0.), (1., 1., 1.), 1)
    print("adaptive loop", [(t["n_dofs"], t["iterations"]) for t in a.run_adaptive_loop(3)])
    a.close()
v = torch.rand(100000, dtype=torch.float64, device="cuda")
out = torch.zeros(1, dtype=torch.float64, device="cuda")
mgb.noalloc_dot(v, v, out)
mgb.noalloc_dot(v[:5000], v[:5000], out)
torch.cuda.synchronize()
print("noalloc", float(out[0]))
print("sanitize run ok")
