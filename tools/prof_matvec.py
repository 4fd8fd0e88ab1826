"""Small driver for ncu: a few finest-level operator applies (fp64 vmult, fp32 vmult, fp32 Chebyshev step)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 128
s = mgb.MultigridSolver(4, 1, cells.bit_length() - 1, degree_pre=3, degree_post=3)
s.do_matvec()
s.do_matvec_smoother()
s.sync()
torch.cuda.profiler.start()  # ncu --profile-from-start off
for _ in range(3):
    s.do_matvec()
for _ in range(3):
    s.do_matvec_smoother()
L = s.n_levels - 1
x = torch.rand(s.n_dofs(), dtype=torch.float32, device="cuda")
b = torch.rand(s.n_dofs(), dtype=torch.float32, device="cuda")
s.smoother_step(L, x, b)
s.sync()
torch.cuda.profiler.stop()
print("ok")
