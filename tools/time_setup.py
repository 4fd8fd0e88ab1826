"""Set-up cost of the solver: wall time of mgb_solver_create for 128^3 cells, first and second creation in
one process (the first one pays CUDA context creation and module loading)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

nref = int(sys.argv[1]) if len(sys.argv) > 1 else 7
torch.cuda.init()
torch.zeros(1, device="cuda")
torch.cuda.synchronize()
for i in range(2):
    t0 = time.perf_counter()
    s = mgb.MultigridSolver(4, 1, nref, degree_pre=3, degree_post=3)
    s.sync()
    t1 = time.perf_counter()
    print("create %d: %.3f s" % (i, t1 - t0), flush=True)
    s.close()
