"""One rank of the multi-GPU DG parity test (tests/test_multi_gpu.py), run under torchrun with one process per GPU:
the z-slab SIP-DG operator (ghost cell layers exchanged over NCCL) against the GLOBAL face-based CPU oracle.
usage: mgpu_dg_worker.py p nx,ny,nz n_refinements"""
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


def main():
    p, nsub, nref = int(sys.argv[1]), tuple(int(v) for v in sys.argv[2].split(",")), int(sys.argv[3])
    basis = int(sys.argv[4]) if len(sys.argv) > 4 else 0      # DG_BASIS_*: 0 FE_DGQ, 1 FE_DGQHermite, 2 Gauss
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    o = orc.DGOperator(p, [s << nref for s in nsub], basis=basis)          # the whole mesh, sheared as matvec_dg
    rng = np.random.default_rng(5)
    x, b, x_old = rng.standard_normal(o.n_dofs()), rng.standard_normal(o.n_dofs()), rng.standard_normal(o.n_dofs())
    Ax = o.vmult(x)
    for precision, dtype, tol in ((mgb.F64, torch.float64, 1e-12), (mgb.F32, torch.float32, 1e-5)):
        comm = mgb.comm_from_torch_distributed()   # a unique id makes one communicator: a fresh one per operator
        g = mgb.LaplaceOperatorDG(p, nsub, nref, trafo=orc.dg_trafo(), device=int(os.environ["LOCAL_RANK"]),
                                  precision=precision, comm=comm, basis=basis)
        q = g.partition()
        assert q["n_global"] == o.n_dofs() and q["n_owned"] * dist.get_world_size() == o.n_dofs()
        layer = q["n_owned"] // q["n_layers"]
        g0 = q["first_layer"] * layer - q["owned_offset"]        # global index of local entry 0
        own = slice(q["owned_offset"], q["owned_offset"] + q["n_owned"])
        gown = slice(g0 + own.start, g0 + own.stop)

        def to_dev(v, poison=True):
            loc = np.full(q["n_local"], 1e30 if poison else 0.0)   # ghosts must be refreshed by the library
            loc[own] = v[gown]
            return torch.tensor(loc, dtype=dtype, device="cuda")

        src, dst = to_dev(x), to_dev(np.zeros_like(x), poison=False)
        g.vmult(dst, src)
        assert rel(dst.double().cpu().numpy()[own], Ax[gown]) <= tol, ("vmult", precision)
        res = to_dev(np.zeros_like(x), poison=False)
        g.vmult_residual(to_dev(b), src, res)
        assert rel(res.double().cpu().numpy()[own], (b - Ax)[gown]) <= tol * 10, ("residual", precision)
        # CG update with its four global sums
        r_, q_, p_, x_ = to_dev(b), to_dev(x_old), to_dev(x), to_dev(b)
        sums = g.vmult_with_cg_update(0.3, 0.7, r_, q_, p_, x_)
        p_new = 0.7 * x + x_old
        q_new = o.vmult(p_new)
        ref = np.array([q_new @ p_new, b @ b, q_new @ b, q_new @ q_new])
        assert np.abs(np.array(sums) - ref).max() <= (1e-11 if precision == mgb.F64 else 2e-4) * np.abs(ref).max(), (sums, ref)
        assert rel(q_.double().cpu().numpy()[own], q_new[gown]) <= tol
        # Chebyshev update with JacobiTransformed (the smoother of poisson_dg)
        g.set_preconditioner(g.JACOBI_TRANSFORMED)
        jt = orc.JacobiTransformed(o) if precision == mgb.F64 and o.n_dofs() <= 20000 else None
        if jt is not None:
            sol, old = to_dev(x), to_dev(x_old)
            g.vmult_with_chebyshev_update(to_dev(b), 2, 0.6, 0.2, sol, old)
            want = x + 0.6 * (x - x_old) + 0.2 * jt.vmult(b - Ax)
            got = old.double().cpu().numpy()[own]   # the update lands in solution_old (the caller swaps, as deal.II does)
            assert rel(got, want[gown]) <= 1e-11, "chebyshev / JacobiTransformed"
        assert g.exchange_count() > 0
        g.close()
    dist.barrier()
    print("mgpu dg worker ok rank %d" % rank, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
