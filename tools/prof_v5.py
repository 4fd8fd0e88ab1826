"""Workload for ncu captures of the float tile kernel on the finest level (135 M DoF): two float matvecs (mode SET),
one smoother step (Chebyshev modes)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

s = mgb.MultigridSolver(4, 1, int(sys.argv[1]) if len(sys.argv) > 1 else 7, degree_pre=3, degree_post=3)
L = s.n_levels - 1
s.sync()
torch.cuda.profiler.start()
s.do_matvec_smoother()
s.do_matvec_smoother()
x = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
b = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
mgb._check(mgb.lib().mgb_smoother_step(s.h, L, mgb._ptr(x), mgb._ptr(b)))
s.sync()
torch.cuda.profiler.stop()
