"""A/B timing of the finest-level tile kernels (135 M DoF): double matvec, float matvec, float Chebyshev step.
usage: [MGB200_LIB=...] python tools/time_tile.py [n_refinements=7] [reps=40]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrid_b200 as mgb

nref = int(sys.argv[1]) if len(sys.argv) > 1 else 7
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
s = mgb.MultigridSolver(4, 1, nref, degree_pre=3, degree_post=3)
L = s.n_levels - 1


def timed(fn):
    fn()
    s.sync()
    best = 1e30
    for _ in range(5):
        s.timer_start()
        for _ in range(reps):
            fn()
        best = min(best, s.timer_stop() / reps)
    return best


out = dict(f64=timed(s.do_matvec), f32=timed(s.do_matvec_smoother))
import torch
x = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
b = torch.rand(s.n_dofs(L), dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
xp, bp = mgb._ptr(x), mgb._ptr(b)
out["cheb"] = timed(lambda: mgb.lib().mgb_smoother_step(s.h, L, xp, bp)) / 3   # degree 3: three fused launches
out = {k: round(v, 4) for k, v in out.items()}
print(os.environ.get("MGB200_LIB", "default"), os.environ.get("MGB200_F32X2", "1"), out)
