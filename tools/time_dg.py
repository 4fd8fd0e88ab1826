"""DG benchmark protocols: matvec_dg/program.cc:139-204 (double vmult, GB/s at 3 N bytes per DoF) and
matvec_dg_cheby/program.cc:110-190 (float Chebyshev update, iteration index 2, factors 0.6 / 0.2, 5 N bytes per DoF).
usage: python tools/time_dg.py [n_cell_steps=20] [degree=4] [basis=0: FE_DGQ, 1: FE_DGQHermite, 2: Gauss]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb


def run(steps=20, degree=4, reps=20, basis=0):
    nsub = tuple(2 if c < steps % 3 else 1 for c in range(3))   # matvec_dg/program.cc:61-64
    out = {}
    for name, prec in (("vmult_f64", mgb.F64), ("vmult_f32", mgb.F32), ("cheby_f32", mgb.F32), ("cheby_f64", mgb.F64)):
        g = mgb.LaplaceOperatorDG(degree, nsub, steps // 3, precision=prec, basis=basis)
        n = g.m()
        a = torch.rand(n, dtype=g.dtype, device="cuda")
        b = torch.rand(n, dtype=g.dtype, device="cuda")
        c = torch.rand(n, dtype=g.dtype, device="cuda")
        if name.startswith("vmult"):
            fn = lambda: g.vmult_async(b, a)
            nvec = 3
        else:
            g.set_preconditioner(g.JACOBI_TRANSFORMED)   # as matvec_dg_cheby/program.cc:108
            fn = lambda: g.chebyshev_update_async(c, 2, 0.6, 0.2, a, b)
            nvec = 5
        fn()
        g.sync()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(3):
            g.timer_start()
            for _ in range(reps):
                fn()
            best = min(best, g.timer_stop() / reps)
        nb = 8 if prec == mgb.F64 else 4
        out[name] = dict(n_dofs=n, ms=best, dofs_per_s=n / best * 1e3, gbs=n * nb * nvec / best / 1e6, bytes_per_dof=nb * nvec)
        g.close()
        del a, b, c
    return out


if __name__ == "__main__":
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    deg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    basis = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    for k, v in run(steps, deg, basis=basis).items():
        print("DG basis=%d p=%d %-10s n=%d  %.3f ms  %.3e DoF/s  %.0f GB/s (%d B/DoF)" % (basis, deg, k, v["n_dofs"], v["ms"], v["dofs_per_s"], v["gbs"], v["bytes_per_dof"]))
