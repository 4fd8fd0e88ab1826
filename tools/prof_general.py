"""Workload for ncu captures of the general cell kernel: variable-coefficient operator on the 64^3-cell box (17 M DoF),
or with `shell` on the curved shell sector; two double and two float matvecs, one float Chebyshev step."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

shell = len(sys.argv) > 1 and sys.argv[1] == "shell"
s = mgb.MultigridSolver(4, 1, 6, problem=mgb.PROBLEM_SHELL_COEF, degree_pre=3, degree_post=3,
                        geometry=mgb.GEOMETRY_SHELL_SECTOR if shell else mgb.GEOMETRY_BOX)
L = s.n_levels - 1
s.do_matvec()
s.do_matvec_smoother()
x = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
b = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
s.sync()
torch.cuda.profiler.start()
s.do_matvec()
s.do_matvec_smoother()
mgb._check(mgb.lib().mgb_smoother_step(s.h, L, mgb._ptr(x), mgb._ptr(b)))
s.sync()
torch.cuda.profiler.stop()
