"""Timing of solve_cg() on the cube (min and mean of `reps` solves) plus the README-style per-level table.
usage: [MGB200_*=...] python tools/time_solve.py [cells=128] [reps=6]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrid_b200 as mgb

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
s = mgb.MultigridSolver(4, 1, cells.bit_length() - 1, degree_pre=3, degree_post=3)
for _ in range(2):
    its, rate = s.solve_cg()
times = []
for _ in range(reps):
    s.timer_start()
    its, rate = s.solve_cg()
    times.append(s.timer_stop())
print("solve_cg %d^3 cells: min %.3f ms, mean %.3f ms, its %d, rate %.5f, L2 %.6e"
      % (cells, min(times), sum(times) / len(times), its, rate, s.compute_l2_error()))
s.profile(True)
s.timings()
s.solve_cg()
t = s.timings() * 1e3
s.profile(False)
print("level smoother   mg_mv  mg_vec restrict prolong   [ms per solve]")
for l in range(s.n_levels):
    print("L%-2d  %8.3f %7.3f %7.3f %8.3f %7.3f" % (l, t[l][5], t[l][0], t[l][4], t[l][1], t[l][2]))
print("sum of table: %.2f ms" % t.sum())
