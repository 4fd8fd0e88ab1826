"""Timing of the adaptive-mesh operator (SURVEY row H, config 5's pattern): FE_Q(3) on hyper_L, refined globally and
then repeatedly towards the re-entrant edge; double and float vmult with the hanging-node constraints applied in
the kernel.  Algorithmic bytes per DoF: 2 N (source and result vector) + the 4-byte index table entry per local node
((p+1)^3 / p^3 per DoF).
usage: python tools/time_adaptive.py [degree=3] [n_global=4] [rounds=3]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import multigrid_b200 as mgb


def run(degree=3, n_global=4, rounds=3, reps=20):
    mask = np.ones((2, 2, 2), dtype=np.uint8)
    mask[1, 1, 1] = 0
    t0 = time.perf_counter()
    op = mgb.LaplaceOperatorAdaptive(degree, (2, 2, 2), mask, n_global_refinements=n_global)
    r = 0.5
    for _ in range(rounds):
        lo, sz = op.cell_boxes()
        c = lo + 0.5 * sz
        op.refine(np.hypot(c[:, 0], c[:, 1]) < r)
        r *= 0.5
    n = op.m()
    info = op.info()
    setup = time.perf_counter() - t0
    out = dict(n_dofs=n, n_cells=op.n_cells(), degree=degree, levels=[int(op.cells()[0].min()), int(op.cells()[0].max())],
               hanging_local_nodes=info["hanging_local_nodes"], constraint_rows=info["constraint_rows"], host_setup_s=setup)
    n3 = (degree + 1) ** 3
    for name, prec, dt, nb in (("vmult_f64", mgb.F64, torch.float64, 8), ("vmult_f32", mgb.F32, torch.float32, 4)):
        x = torch.rand(n, dtype=dt, device="cuda")
        y = torch.zeros_like(x)
        op.vmult(y, x, prec)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            for _ in range(reps):
                mgb._check(mgb.lib().mgb_adaptive_vmult(op.h, prec, mgb._ptr(y), mgb._ptr(x)))
            op.sync()
            best = min(best, (time.perf_counter() - t0) / reps * 1e3)
        bytes_per_dof = 2 * nb + 4.0 * n3 * op.n_cells() / n
        out[name] = dict(ms=best, dofs_per_s=n / best * 1e3, bytes_per_dof=bytes_per_dof, gbs=n * bytes_per_dof / best / 1e6)
    # the solve of poisson_l (program.cc:337-420): local-smoothing V-cycle (float) inside CG (double), right-hand side 1
    t0 = time.perf_counter()
    op.setup_multigrid(3)
    op.sync()
    mg_setup = time.perf_counter() - t0
    free = torch.tensor(~op.dirichlet_mask(), device="cuda")
    b = torch.where(free, torch.ones(n, dtype=torch.float64, device="cuda"), torch.zeros(n, dtype=torch.float64, device="cuda"))
    x = torch.zeros_like(b)
    op.solve_cg(x, b)
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        its, rate, hist = op.solve_cg(x, b)
        best = min(best, time.perf_counter() - t0)
    out["solve_cg"] = dict(ms=best * 1e3, iterations=its, reduction_per_iteration=rate, dofs_per_s=n / best,
                           n_levels=op.n_levels, multigrid_setup_s=mg_setup,
                           level_dofs=[int(op.mg_level_info(l)["n_dofs"]) for l in range(op.n_levels)],
                           level_edge_dofs=[int(op.mg_level_info(l)["n_edge_dofs"]) for l in range(op.n_levels)])
    op.close()
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    print(run(*a))
