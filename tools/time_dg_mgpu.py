"""Weak scaling of the DG operator (matvec_dg protocol per GPU: 128 x 128 x 64 cells, FE_DGQ(4), 131 M DoF per GPU):
z-slabs of cell layers, ghost layers over NCCL.  Run under torchrun, one process per GPU.
usage: python -m torch.distributed.run --nproc-per-node N tools/time_dg_mgpu.py [reps=10] [basis=0: FE_DGQ, 1: FE_DGQHermite, 2: Gauss]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import multigrid_b200 as mgb


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    basis = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    out = {}
    for name, prec in (("vmult_f64", mgb.F64), ("cheby_f32", mgb.F32)):
        comm = mgb.comm_from_torch_distributed() if world > 1 else None
        g = mgb.LaplaceOperatorDG(4, (2, 2, world), 6, precision=prec, device=lr, comm=comm, basis=basis)
        q = g.partition()
        a = torch.rand(q["n_local"], dtype=g.dtype, device="cuda")
        b = torch.rand(q["n_local"], dtype=g.dtype, device="cuda")
        c = torch.rand(q["n_local"], dtype=g.dtype, device="cuda")
        if name.startswith("vmult"):
            fn = lambda: g.vmult_async(b, a)
        else:
            g.set_preconditioner(g.JACOBI_TRANSFORMED)
            fn = lambda: g.chebyshev_update_async(c, 2, 0.6, 0.2, a, b)
        fn()
        g.sync()
        best = 1e30
        for _ in range(3):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            g.timer_start()
            for _ in range(reps):
                fn()
            ms = torch.tensor([g.timer_stop() / reps], device="cuda")
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            best = min(best, float(ms))
        out[name] = dict(n_dofs=q["n_global"], ms=best, dofs_per_s=q["n_global"] / best * 1e3, exchanges=g.exchange_count())
        g.close()
        del a, b, c
    if int(os.environ.get("RANK", "0")) == 0:
        print("DG weak scaling, %d GPU(s): %s" % (world, out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
