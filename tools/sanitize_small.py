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
print("sanitize run ok")
