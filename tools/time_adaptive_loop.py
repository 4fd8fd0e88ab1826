"""poisson_l/program.cc:497-602 on the GPU (default case: extruded L, exact solution r^(2/3) sin(2 phi / 3)): the
convergence table of the program with wall times of the parts per cycle."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

import multigrid_b200 as mgb

p = int(sys.argv[1]) if len(sys.argv) > 1 else 2
max_size = int(sys.argv[2]) if len(sys.argv) > 2 else 2000000
mask = np.ones((1, 2, 2), dtype=np.uint8)
mask[0, 1, 1] = 0
op = mgb.LaplaceOperatorAdaptive(p, (2, 2, 1), mask, (-1., -1., 0.), (1., 1., 1.), 3)      # refine_global(3), :513
rows = []
for cycle in range(20):
    row = {}
    if cycle > 0:
        t = time.perf_counter()
        u = torch.tensor(op.solution(), device="cuda")
        eta = op.estimate_kelly(u)
        row["estimate_s"] = time.perf_counter() - t
        t = time.perf_counter()
        op.refine_and_coarsen_fixed_number(eta, 0.3, 0.03)
        row["refine_s"] = time.perf_counter() - t
    t = time.perf_counter()
    n = op.m()
    row["numbering_s"] = time.perf_counter() - t
    t = time.perf_counter()
    row.update(op.solve_cycle(op.FUNCTION_L_SINGULARITY, 0., 3))
    row["solve_cycle_s"] = time.perf_counter() - t
    # the solve alone, again (the program times the second solve as well, :400-404)
    b = torch.zeros(n, dtype=torch.float64, device="cuda")
    ub = torch.zeros_like(b)
    op.interpolate_boundary_values(ub, op.FUNCTION_L_SINGULARITY)
    op.compute_residual(b, ub, 0.)
    x = torch.zeros_like(b)
    torch.cuda.synchronize()
    t = time.perf_counter()
    its, rate, hist = op.solve_cg(x, b)
    torch.cuda.synchronize()
    row["cg_s"] = time.perf_counter() - t
    lev, _ = op.cells()
    row["levels"] = int(lev.max()) + 1
    rows.append(row)
    print(json.dumps(row), flush=True)
    if n > max_size:
        break
