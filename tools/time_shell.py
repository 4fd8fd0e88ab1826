"""BASELINE config 3 on the structured curved mesh: FE_Q(4) on one coarse cell of hyper_shell(0, 0.5, 1, 6) (a cube face
blown up between the spheres r = 0.5 and r = 1, poisson_shell/program.cc:418-431) refined `nref` times, coefficient
1 + 1e6 prod cos^2(2 pi x_e + 0.1 e) and solution sin(2 pi (x + y)) (:97-225): the six-entry merged coefficient from the
Jacobian of the polynomial cell geometry streams once per matvec (laplace_operator.h:388-430, :493-522).  Reports the
operator (protocol of poisson_cube/program.cc:343-380), solve_cg and the FMG solve.
usage: python tools/time_shell.py [n_refinements=6] [degree=4]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrid_b200 as mgb


def run(nref=6, degree=4, reps=10):
    t0 = time.perf_counter()
    s = mgb.MultigridSolver(degree, 1, nref, problem=mgb.PROBLEM_SHELL_COEF, geometry=mgb.GEOMETRY_SHELL_SECTOR,
                            degree_pre=3, degree_post=3)
    s.sync()
    setup = time.perf_counter() - t0
    n = s.n_dofs()
    out = {"n_dofs": n, "cells": 8 ** nref, "setup_s": setup}
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
        bpd = 2 * nb + 6 * nb * nq     # vectors + coefficient stream (SURVEY 8(d) without the index table this kernel has not)
        out[name] = dict(ms=best, dofs_per_s=n / best * 1e3, bytes_per_dof=bpd, gbs=n * bpd / best / 1e6)
    s.solve_cg()
    best = 1e30
    for _ in range(3):
        s.timer_start()
        its, rate = s.solve_cg()
        best = min(best, s.timer_stop())
    out["solve_cg"] = dict(ms=best, iterations=its, reduction_per_iteration=rate, l2_error=s.compute_l2_error(),
                           dofs_per_s=n / best * 1e3)
    s.timer_start()
    red = s.solve(False)
    out["solve_fmg"] = dict(ms=s.timer_stop(), reduction=red, l2_error=s.compute_l2_error())
    s.close()
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    r = run(*a)
    print(r)
