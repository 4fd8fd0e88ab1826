"""One rank of the multi-GPU parity test (tests/test_multi_gpu.py), run under torchrun with one process
per GPU: the z-slab solver against the GLOBAL CPU oracle on the same mesh.  Every rank checks its owned
entries; sums, iteration counts and errors must agree with the oracle's.
usage: mgpu_worker.py p nx,ny,nz n_refinements"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import multigrid_b200 as mgb  # noqa: E402
from oracle import orc  # noqa: E402  (the checker)


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


class Slab:
    """local <-> global-lexicographic views of one level of this rank"""

    def __init__(self, s, level):
        self.s, self.level = s, s._l(level)
        self.q = s.partition(level)
        self.lex = s.lex_map(level)
        nd = s.dof_grid(level)
        self.pitch = nd[0] * nd[1]
        self.g0 = self.q["plane0"] * self.pitch
        self.nloc = self.q["n_local"]
        own = np.zeros(self.nloc, dtype=bool)
        own[self.q["owned_begin"]:self.q["owned_begin"] + self.q["n_owned"]] = True
        self.own = own

    def to_dev(self, x_glob, double=True, poison=True):
        v = np.zeros(self.nloc)
        v[self.lex] = x_glob[self.g0:self.g0 + self.nloc]
        if poison and self.q["distributed"]:
            v[~self.own] = 1e30  # ghosts must be refreshed by the library, never trusted
        return torch.tensor(v, dtype=torch.float64 if double else torch.float32, device="cuda")

    def owned_pair(self, t, y_glob):
        """(library values, oracle values) on the owned entries"""
        ref = np.zeros(self.nloc)
        ref[self.lex] = y_glob[self.g0:self.g0 + self.nloc]
        return t.double().cpu().numpy()[self.own], ref[self.own]


def main():
    p, nsub, nref = int(sys.argv[1]), tuple(int(v) for v in sys.argv[2].split(",")), int(sys.argv[3])
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rank, R = dist.get_rank(), dist.get_world_size()
    comm = mgb.comm_from_torch_distributed()
    s = mgb.MultigridSolver(p, nsub, nref, device=local_rank, comm=comm)
    o = orc.MultigridSolver(p, nsub, nref, degree_pre=3, degree_post=3)
    L = s.n_levels - 1
    assert s.partition(L)["distributed"], "finest level must be distributed in this test"
    rng = np.random.default_rng(42)
    n_checked = 0
    for level in range(s.n_levels):
        sl = Slab(s, level)
        N = o.n_dofs(level)
        assert sl.q["n_global"] == N
        x, b = rng.random(N), rng.random(N)
        for double in (True, False):
            tol = 1e-12 if double else 1e-5
            op = s.matrix_dp(level) if double else s.matrix(level)
            src, dst = sl.to_dev(x, double), sl.to_dev(b, double)
            op.vmult(dst, src)
            got, ref = sl.owned_pair(dst, o.vmult(x, level, double))
            assert rel(got, ref) < tol, ("vmult", level, double, rel(got, ref))
            res = torch.zeros_like(src)
            op.vmult_residual(sl.to_dev(b, double), src, res)
            got, ref = sl.owned_pair(res, o.vmult_residual(b, x, level, double))
            assert rel(got, ref) < tol, ("vmult_residual", level, double, rel(got, ref))
            n_checked += 2
        # sums of vmult_with_cg_update run over all ranks
        r_, q_, p_, x_ = (rng.random(N) for _ in range(4))
        sums = s.matrix_dp(level).vmult_with_cg_update(0.3, 0.7, *(sl.to_dev(v) for v in (r_, q_, p_, x_)))
        sums_ref = o.vmult_with_cg_update(0.3, 0.7, r_, q_, p_, x_, level)[0]
        assert np.allclose(sums, sums_ref, rtol=1e-11), ("cg_update sums", level, sums, sums_ref)
        # smoother: same eigenvalue estimate, same step
        gi, oi = s.smoother_info(level), o.chebyshev_info(level)
        assert abs(gi["max_eig"] - oi["max_eig"]) < 2e-4 * oi["max_eig"], (level, gi, oi)
        assert gi["degree"] == oi["degree"]
        if level > 0:
            xs = sl.to_dev(x, False)
            s.smoother_step(level, xs, sl.to_dev(b, False))
            got, ref = sl.owned_pair(xs, o.chebyshev_step(x, b, level))
            assert rel(got, ref) < 2e-3, ("smoother step", level, rel(got, ref))
            # transfer between this level and the one below
            sc = Slab(s, level - 1)
            Nc = o.n_dofs(level - 1)
            xc = rng.random(Nc)
            fine = sl.to_dev(np.zeros(N), False, poison=False)
            s.prolongate(level, fine, sc.to_dev(xc, False))
            got, ref = sl.owned_pair(fine, o.prolongate(xc, level))
            assert rel(got, ref) < 1e-5, ("prolongate", level, rel(got, ref))
            coarse = sc.to_dev(np.zeros(Nc), False, poison=False)
            s.restrict_and_add(level, coarse, sl.to_dev(x, False))
            got, ref = sc.owned_pair(coarse, o.restrict_and_add(x, level))
            assert rel(got, ref) < 1e-5, ("restrict", level, rel(got, ref))
            n_checked += 3
    # compress(add) / zero_out_ghost_values of the vector interface (mgb_vec_compress_add): ghost planes are added
    # to their owners on the neighbouring slabs, then zeroed
    sl = Slab(s, L)
    for dt, prec in ((torch.float64, mgb.F64), (torch.float32, mgb.F32)):
        v = torch.full((sl.nloc,), 2.0, dtype=dt, device="cuda")
        v[torch.tensor(sl.own, device="cuda")] = 1.0
        mgb._check(mgb.lib().mgb_vec_compress_add(s.h, L, prec, mgb._ptr(v)))
        s.sync()
        got = v.double().cpu().numpy()
        ob, no = sl.q["owned_begin"], sl.q["n_owned"]
        want = np.zeros(sl.nloc)
        want[ob:ob + no] = 1.0
        if rank > 0:                       # the plane shared with the top ghost plane of the rank below
            want[ob:ob + sl.pitch] += 2.0
        if rank + 1 < R:                   # the p planes the rank above holds as ghosts
            want[ob + no - p * sl.pitch:ob + no] += 2.0
        assert np.array_equal(got, want), ("compress_add", dt, np.abs(got - want).max())
        v.fill_(5.0)
        mgb._check(mgb.lib().mgb_vec_zero_out_ghost_values(s.h, L, prec, mgb._ptr(v)))
        s.sync()
        got = v.double().cpu().numpy()
        assert np.all(got[sl.own] == 5.0) and np.all(got[~sl.own] == 0.0), "zero_out_ghost_values"
    n_checked += 4
    # V-cycle as preconditioner, FMG, CG
    N = o.n_dofs(L)
    nd = s.dof_grid(L)
    mask = np.ones((nd[2] if not sl.q["distributed"] else sl.q["n_planes_global"], nd[1], nd[0]))
    mask[0], mask[-1], mask[:, 0], mask[:, -1], mask[:, :, 0], mask[:, :, -1] = (0,) * 6
    r_ = rng.random(N) * mask.ravel()  # zero on the Dirichlet boundary, as a CG residual is
    dst = torch.zeros(sl.nloc, dtype=torch.float64, device="cuda")
    s.vmult(dst, sl.to_dev(r_))
    got, ref = sl.owned_pair(dst, o.vcycle(r_))
    assert rel(got, ref) < 5e-3, ("vcycle", rel(got, ref))
    rate, trace = s.solve(True)
    rate_o, trace_o = o.solve(True)
    assert np.allclose(trace[1:], trace_o[1:], rtol=2e-2), (trace, trace_o)
    e_fmg, e_fmg_o = s.compute_l2_error(), o.compute_l2_error()
    assert abs(e_fmg - e_fmg_o) < 5e-3 * e_fmg_o, (e_fmg, e_fmg_o)
    its, red = s.solve_cg()
    its_o, red_o, hist_o = o.solve_cg()
    assert its == its_o, (its, its_o)
    hist = s.cg_history()
    assert np.allclose(hist, hist_o, rtol=0.05), (hist, hist_o)
    e, e_o = s.compute_l2_error(), o.compute_l2_error()
    assert abs(e - e_o) < 1e-3 * e_o, (e, e_o)
    sol = torch.tensor(s.get_solution(), device="cuda")
    got, ref = sl.owned_pair(sol, o.get_vector(orc.VEC_SOLUTION))
    assert rel(got, ref) < 1e-6, ("solution", rel(got, ref))
    assert s.exchange_count() > 0
    dist.barrier()
    print("mgpu worker ok rank %d/%d: %d vector checks, cg its %d (oracle %d), L2 %.4e (oracle %.4e), "
          "%d exchanges, %d launches" % (rank, R, n_checked, its, its_o, e, e_o, s.exchange_count(), s.launch_count()),
          flush=True)
    s.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
