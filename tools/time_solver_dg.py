"""solver_dg/program.cc on the GPU: time per CG iteration of the cell-based and the interleaved loop, DoFs/s."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

import multigrid_b200 as mgb

p = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 21        # n_cell_steps of the program: refinements = steps / 3
n_it = int(sys.argv[3]) if len(sys.argv) > 3 else 100
nsub = [2 if c < steps % 3 else 1 for c in range(3)]
trafo = np.eye(3) + 0.12 * np.outer(np.arange(1, 4), np.arange(1, 4))
out = {}
for prec, name in ((mgb.F64, "f64"), (mgb.F32, "f32")):
    g = mgb.LaplaceOperatorDG(p, nsub, steps // 3, lower=[-1 + 0.05 * (d + 1) for d in range(3)],
                              upper=[0.95 - 0.06 * d for d in range(3)], trafo=trafo, precision=prec, basis=mgb.DG_BASIS_GAUSS)
    D = g.lumped_mass_inverse()
    b = torch.rand(g.m(), dtype=g.dtype, device="cuda")
    x = g.initialize_dof_vector()
    y = g.initialize_dof_vector()
    best = 1e10
    for _ in range(3):
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(20):
            g.vmult(y, b)
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t) / 20)
    out[f"{name}_vmult"] = dict(n_dofs=g.m(), ms=best * 1e3)
    for interleaved in (False, True):
        best = 1e10
        for _ in range(3):
            torch.cuda.synchronize()
            t = time.perf_counter()
            its, res = g.solve_cg(x, b, D, n_it, interleaved)
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t) / its)
        out[f"{name}_{'interleaved' if interleaved else 'cell_based'}"] = dict(
            n_dofs=g.m(), ms_per_iteration=best * 1e3, dofs_per_s=g.m() / best, residual=res)
print(json.dumps(out, indent=1))
