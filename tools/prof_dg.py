"""ncu driver: a few DG operator applies (double vmult, float Chebyshev update) on 16 M DoF.
usage: python tools/prof_dg.py [basis=0: FE_DGQ, 1: FE_DGQHermite, 2: Gauss]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

BASIS = int(sys.argv[1]) if len(sys.argv) > 1 else 0
for prec in (mgb.F64, mgb.F32):
    g = mgb.LaplaceOperatorDG(4, (2, 2, 1), 5, precision=prec, basis=BASIS)
    n = g.m()
    a, b, c = (torch.rand(n, dtype=g.dtype, device="cuda") for _ in range(3))
    g.get_matrix_diagonal_inverse()
    g.vmult(b, a)
    torch.cuda.profiler.start()
    g.vmult(b, a)
    g.vmult_with_chebyshev_update(c, 2, 0.6, 0.2, a, b)
    torch.cuda.profiler.stop()
print("ok")
