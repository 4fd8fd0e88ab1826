"""Times the finest-level operator applies with the solver's CUDA-event stopwatch.
usage: python tools/time_matvec.py [cells] [degree]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import multigrid_b200 as mgb

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 128
deg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
s = mgb.MultigridSolver(deg, 1, cells.bit_length() - 1, degree_pre=3, degree_post=3)
n = s.n_dofs()
L = s.n_levels - 1
x = torch.rand(n, dtype=torch.float32, device="cuda")
b = torch.rand(n, dtype=torch.float32, device="cuda")


def bench(fn, reps=20):
    fn()
    s.sync()
    best = 1e30
    for _ in range(3):
        s.timer_start()
        for _ in range(reps):
            fn()
        best = min(best, s.timer_stop() / reps)
    return best


t64 = bench(s.do_matvec)
t32 = bench(s.do_matvec_smoother)
s.profile(True)
s.class_times(L)
for _ in range(5):
    s.smoother_step(L, x, b)
c = s.class_times(L)
tch = c["apply_f32_chebyshev"][0] / max(c["apply_f32_chebyshev"][1], 1)
s.profile(False)
print("%s p=%d n=%d  fp64 %.3f ms (%.3g DoF/s, %.0f GB/s)  fp32 %.3f ms (%.3g DoF/s)  cheb32 %.3f ms (%.0f GB/s)"
      % (os.environ.get("MGB200_LIB", "default"), deg, n, t64, n / t64 * 1e3, 16 * n / t64 / 1e6, t32, n / t32 * 1e3,
         tch, 20 * n / tch / 1e6))
