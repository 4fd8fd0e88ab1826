"""Per-level wall-time table (print_wall_times, multigrid_solver.h:348-371) of solve_cg on the bench mesh,
under torchrun with one rank per GPU.  usage: torchrun ... tools/mgpu_probe.py [cells]"""
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
comm = None
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    comm = mgb.comm_from_torch_distributed()
nsub, nref, lo, hi = bench.weak_scaling_mesh(cells, world)
s = mgb.MultigridSolver(4, nsub, nref, lower=lo, upper=hi, device=lr, comm=comm)
for _ in range(3):
    s.solve_cg()
x0 = s.exchange_count()
s.timer_start()
for _ in range(3):
    its, _ = s.solve_cg()
ms = s.timer_stop() / 3
s.profile(True)
s.timings()
s.solve_cg()
t = s.timings() * 1e3
s.profile(False)
if lr == 0:
    print("ranks %d  min_layers %s overlap %s: %.2f ms/solve, its %d, exchanges/solve %d" % (
        world, os.environ.get("MGB200_MIN_SLAB_LAYERS", "2"), os.environ.get("MGB200_OVERLAP", "1"), ms, its,
        (s.exchange_count() - x0) // 3))
    print("level dist  smoother   mg_mv  mg_vec restrict prolong   [ms per solve]")
    for l in range(s.n_levels):
        print("L%-2d   %d   %8.3f %7.3f %7.3f %8.3f %7.3f" % (l, s.partition(l)["distributed"], t[l][5], t[l][0], t[l][4],
                                                            t[l][1], t[l][2]))
    print("sum of table: %.2f ms" % t.sum())
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
