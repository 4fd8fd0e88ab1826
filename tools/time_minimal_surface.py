"""minimal_surface/program.cc on the GPU in 3-D: Newton steps on [-1,1]^3 with Scherk's surface as Dirichlet data.
Prints per step: residuals, CG iterations, wall time; and the error against the exact solution."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

import multigrid_b200 as mgb

p = int(sys.argv[1]) if len(sys.argv) > 1 else 4
nref = int(sys.argv[2]) if len(sys.argv) > 2 else 5
scherk = lambda x, y, z: np.log(np.cos(x) / np.cos(y)) + 0 * z
s = mgb.MultigridSolver(p, 1, nref, lower=(-1.,) * 3, upper=(1.,) * 3, degree_pre=2, degree_post=2, n_cycles=1)
t0 = time.perf_counter()
s.minimal_surface_set_boundary_values(scherk)
t_bc = time.perf_counter() - t0
steps = []
for it in range(20):
    torch.cuda.synchronize()
    t = time.perf_counter()
    st = s.minimal_surface_step(it == 0)
    st["ms"] = (time.perf_counter() - t) * 1e3
    steps.append(st)
    if st["final_residual"] < 1e-10:
        break
x = s.support_points()
err = float(np.abs(s.get_solution() - scherk(x[:, 0], x[:, 1], x[:, 2])).max())
print(json.dumps(dict(degree=p, n_refinements=nref, n_dofs=s.n_dofs(), boundary_values_s=t_bc, steps=steps,
                      total_ms=sum(st["ms"] for st in steps), max_error_vs_scherk=err), indent=1))
