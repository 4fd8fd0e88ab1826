"""Weak scaling of poisson_dg (poisson_dg/program.cc:227-263) with one process per GPU: doubling-mesh cycle
`base + log2(n_gpus)`, so every GPU keeps the DoFs of cycle `base` (the doubling goes x, y, z; the z-slabs cut z).
Minimum of 4 CG solves, time = max over ranks (barrier on both sides).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/time_dg_solver_mgpu.py
       [base_cycle=21] [degree=4] [tolerance=1e-3]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import multigrid_b200 as mgb
from time_dg_solver import doubling_mesh


def main():
    base = int(sys.argv[1]) if len(sys.argv) > 1 else 21
    degree = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    tol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-3
    basis = int(sys.argv[4]) if len(sys.argv) > 4 else 0   # 0: FE_DGQ, 1: FE_DGQHermite, 2: Gauss
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = mgb.comm_from_torch_distributed()
    cycle = base + (world.bit_length() - 1)
    nsub, nref, lower, upper = doubling_mesh(cycle)
    s = mgb.MultigridSolverDG(degree, nsub, nref, lower, upper, 3, 3, 1, device=local, comm=comm, basis=basis)
    q = s.partition()
    best = 1e30
    for _ in range(4):
        s.sync()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        est, rate = s.solve_cg(tol)
        s.sync()
        if world > 1:
            dist.barrier()
        best = min(best, time.perf_counter() - t0)
    err = s.compute_l2_error()
    if world > 1:
        t = torch.tensor([best], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = float(t)
    if rank == 0:
        print(json.dumps(dict(n_gpus=world, cycle=cycle, cells=[k << nref for k in nsub], degree=degree, tolerance=tol,
                              n_dofs=q["n_global"], dofs_per_gpu=q["n_owned"], cg_iterations=s.last_step,
                              reduction_rate=rate, solve_ms=best * 1e3, dofs_per_s=q["n_global"] / best, l2_error=err)),
              flush=True)
    s.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    main()
