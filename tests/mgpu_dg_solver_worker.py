"""One rank of the multi-GPU parity test of the DG multigrid solver (tests/test_multi_gpu.py), run under torchrun
with one process per GPU: MultigridSolverDG on z-slabs (DG operator and finest FE_Q level share the slabs and one
NCCL communicator) against the GLOBAL oracle composition (oracle/orc.py MultigridSolverDG).
usage: mgpu_dg_solver_worker.py p nx,ny,nz n_refinements"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import multigrid_b200 as mgb  # noqa: E402
from oracle import orc  # noqa: E402  (the checker)


def main():
    p, nsub, nref = int(sys.argv[1]), tuple(int(v) for v in sys.argv[2].split(",")), int(sys.argv[3])
    basis = int(sys.argv[4]) if len(sys.argv) > 4 else 0      # DG_BASIS_*: 0 FE_DGQ, 1 FE_DGQHermite, 2 Gauss
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lower, upper = (-1.,) * 3, tuple(-1. + 2. * s for s in nsub)
    o = orc.MultigridSolverDG(p, nsub, nref, lower, upper, basis=basis)      # the whole mesh
    g = mgb.MultigridSolverDG(p, nsub, nref, lower, upper, device=local, comm=mgb.comm_from_torch_distributed(),
                              basis=basis)
    q = g.partition()
    assert q["n_global"] == o.n_dofs() and q["n_owned"] * dist.get_world_size() == o.n_dofs()
    assert g.n_dofs() == q["n_local"]
    layer = q["n_owned"] // q["n_layers"]
    g0 = q["first_layer"] * layer - q["owned_offset"]            # global index of local entry 0
    own = slice(q["owned_offset"], q["owned_offset"] + q["n_owned"])
    gown = slice(g0 + own.start, g0 + own.stop)

    rhs = g.get_vector(1)
    assert np.abs(rhs[own] - o.rhs[gown]).max() <= 1e-12 * np.abs(o.rhs).max()
    gi = g.smoother_info()
    assert gi["degree"] == o.smooth.degree and gi["cg_iterations"] == o.smooth.cg_iterations
    assert abs(gi["max_eig"] - o.smooth.max_eig) <= 2e-3 * o.smooth.max_eig, (gi, o.smooth.max_eig)

    # one cycle as preconditioner on the same vector
    rng = np.random.default_rng(3)
    src = rng.standard_normal(o.n_dofs())
    loc = np.full(q["n_local"], 1e30)                             # ghosts must be refreshed by the library
    loc[own] = src[gown]
    dst = torch.zeros(q["n_local"], dtype=torch.float64, device="cuda")
    g.vmult(dst, torch.tensor(loc, device="cuda"))
    ref = o.vmult(src)
    assert np.abs(dst.cpu().numpy()[own] - ref[gown]).max() <= 2e-4 * np.abs(ref).max()

    for tol in (1e-3, 1e-8):
        est_o, rate_o, hist_o = o.solve_cg(tol)
        est_g, rate_g = g.solve_cg(tol)
        hist_g = g.residual_history()
        assert g.last_step == o.last_step, (hist_g, hist_o)
        assert np.all(np.abs(hist_g - hist_o) <= 0.1 * hist_o + 1e-6 * hist_o[0]), (hist_g, hist_o)
        assert abs(rate_g - rate_o) <= 3e-2 * rate_o
        eg, eo = g.compute_l2_error(), o.compute_l2_error()
        assert abs(eg - eo) <= 1e-3 * eo, (eg, eo)
    xs = g.get_solution()
    assert np.abs(xs[own] - o.solution[gown]).max() <= 1e-6 * np.abs(o.solution).max()
    dist.barrier()
    print("mgpu dg solver worker ok rank %d its %d l2 %.4e" % (rank, g.last_step, eg), flush=True)
    g.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
