"""Variable-coefficient operator (per-quadrature-point merged coefficient, laplace_operator.h:388-430,493-522;
the streaming pattern of poisson_shell) on the box mesh: matvec timing, protocol of poisson_cube/program.cc:343-380.
usage: python tools/time_varcoef.py [cells=64] [degree=4]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrid_b200 as mgb


def run(cells=64, degree=4, reps=10):
    s = mgb.MultigridSolver(degree, 1, cells.bit_length() - 1, problem=mgb.PROBLEM_SHELL_COEF, degree_pre=3, degree_post=3)
    n = s.n_dofs()
    out = {"n_dofs": n, "cells": cells ** 3}
    nq = (degree + 1) ** 3 / degree ** 3
    for name, fn, nb in (("fp64", s.do_matvec, 8), ("fp32", s.do_matvec_smoother, 4)):
        fn()
        s.sync()
        best = 1e30
        for _ in range(3):
            s.timer_start()
            for _ in range(reps):
                fn()
            best = min(best, s.timer_stop() / reps)
        bpd = 2 * nb + 6 * nb * nq + 4.0 * 27 / degree ** 3   # vectors + coefficient stream + indices
        out[name] = dict(ms=best, dofs_per_s=n / best * 1e3, bytes_per_dof=bpd, gbs=n * bpd / best / 1e6)
    s.close()
    return out


if __name__ == "__main__":
    cells = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    deg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    r = run(cells, deg)
    for k in ("fp64", "fp32"):
        v = r[k]
        print("varcoef p=%d %s n=%d  %.3f ms  %.3e DoF/s  %.0f GB/s (%.1f B/DoF)" % (deg, k, r["n_dofs"], v["ms"], v["dofs_per_s"], v["gbs"], v["bytes_per_dof"]))
