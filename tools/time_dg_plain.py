"""poisson_dg_plain/program.cc:232-330 on the GPU: solve_cg(1e-3) timings of the pure DG h-multigrid solver for the
three bases, beside the DG/FE_Q solver on the same mesh."""
import json
import sys
import time

import os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

import multigrid_b200 as mgb

nref = int(sys.argv[1]) if len(sys.argv) > 1 else 5
out = {}
for name, basis in (("fe_dgq_hermite", mgb.DG_BASIS_HERMITE), ("fe_dgq", mgb.DG_BASIS_GAUSS_LOBATTO), ("gauss", mgb.DG_BASIS_GAUSS)):
    for kind, cls in (("dg_plain", mgb.MultigridSolverDGPlain), ("dg_over_fe_q", mgb.MultigridSolverDG)):
        s = cls(4, (1, 1, 1), nref, basis=basis)
        best = 1e10
        for _ in range(4):
            torch.cuda.synchronize()
            t = time.perf_counter()
            est, rate = s.solve_cg(1e-3)
            s.sync()
            best = min(best, time.perf_counter() - t)
        out[f"{kind}_{name}"] = dict(n_dofs=s.n_dofs(), cg_ms=best * 1e3, its=s.last_step, its_estimate=est, rate=rate,
                                     l2_error=s.compute_l2_error())
        s.close()
print(json.dumps(out, indent=1))
