"""Driver for the ncu launch list: one solve_cg() between cudaProfilerStart/Stop (ncu --profile-from-start off)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 128
s = mgb.MultigridSolver(4, 1, cells.bit_length() - 1, degree_pre=3, degree_post=3)
s.solve_cg()
s.sync()
torch.cuda.profiler.start()
its, rate = s.solve_cg()
s.sync()
torch.cuda.profiler.stop()
print("ok", its, rate)
