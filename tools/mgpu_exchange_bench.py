"""Latency of one update_ghost_values per level (mgb_vec_update_ghost_values), under torchrun, one rank per GPU.
usage: torchrun ... tools/mgpu_exchange_bench.py [cells]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
import multigrid_b200 as mgb

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 128
lr = int(os.environ.get("LOCAL_RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
comm = mgb.comm_from_torch_distributed()
nsub, nref, lo, hi = bench.weak_scaling_mesh(cells, world)
s = mgb.MultigridSolver(4, nsub, nref, lower=lo, upper=hi, device=lr, comm=comm)
L = mgb.lib()
for level in range(s.n_levels):
    q = s.partition(level)
    if not q["distributed"]:
        continue
    for prec, dt in ((mgb.F32, torch.float32), (mgb.F64, torch.float64)):
        v = torch.zeros(q["n_local"], dtype=dt, device="cuda")
        torch.cuda.synchronize()
        for reps in (5, 50):
            dist.barrier()
            s.timer_start()
            for _ in range(reps):
                mgb._check(L.mgb_vec_update_ghost_values(s.h, level, prec, C.c_void_p(v.data_ptr())))
            ms = s.timer_stop() / reps
        nd = s.dof_grid(level)
        if lr == 0:
            print("P2P=%s ranks %d level %d %s: plane %d values, %.1f us per exchange (%.2f MB up)" % (
                os.environ.get("MGB200_P2P", "1"), world, level, "f32" if prec == mgb.F32 else "f64", nd[0] * nd[1], ms * 1e3,
                4 * nd[0] * nd[1] * (4 if prec == mgb.F32 else 8) / 1e6), flush=True)
dist.barrier()
dist.destroy_process_group()
