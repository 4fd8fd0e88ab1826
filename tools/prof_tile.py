"""Workload for ncu captures of the finest-level tile kernels: set-up of the 135 M DoF solver, two float matvecs,
one CG solve.  tools/prof_tile.sh lists the launches first and then captures the finest-level ones with --set full."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrid_b200 as mgb

s = mgb.MultigridSolver(4, 1, int(sys.argv[1]) if len(sys.argv) > 1 else 7, degree_pre=3, degree_post=3)
s.do_matvec_smoother()
s.do_matvec_smoother()
s.do_matvec()
s.solve_cg()
s.sync()
