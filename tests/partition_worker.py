"""One rank of the world_size-2 gloo test (tests/test_partition.py): carries out update_ghost_values,
compress(add) and an owned-range dot product on numpy arrays by the exchange plan of mgb_host_partition,
and checks them against the global field.  usage (under torchrun): partition_worker.py p nx,ny,nz level"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import multigrid_b200 as mgb  # noqa: E402


def main():
    p, nsub, level = int(sys.argv[1]), tuple(int(v) for v in sys.argv[2].split(",")), int(sys.argv[3])
    dist.init_process_group("gloo")
    rank, R = dist.get_rank(), dist.get_world_size()
    q = mgb.host_partition(p, nsub, level, R, rank, want_map=True)
    assert q["distributed"], "test case must give a distributed level"
    nd0, nd1 = (nsub[0] << level) * p + 1, (nsub[1] << level) * p + 1
    pitch = nd0 * nd1
    lex = q["lex_to_dof"]  # local lexicographic point -> local index
    # global field: value = global lexicographic index + 1
    nloc = q["n_local"]
    glob_of_local = np.zeros(nloc, dtype=np.int64)
    glob_of_local[lex] = np.arange(nloc) + q["plane0"] * pitch
    field = (glob_of_local + 1).astype(np.float64)

    # ---- update_ghost_values: owned entries known, ghosts poisoned
    v = np.full(nloc, np.nan)
    ob, no = q["owned_begin"], q["n_owned"]
    v[ob:ob + no] = field[ob:ob + no]
    t = torch.from_numpy(v)

    def seg(name):
        o, n = q[name]
        return t[o:o + n]

    reqs = []
    if q["ghost_plane_top"]:
        reqs.append(dist.isend(seg("send_up").clone(), rank + 1, tag=1))
        reqs.append(dist.irecv(seg("recv_up"), rank + 1, tag=2))
    if q["ghost_layers_below"]:
        reqs.append(dist.isend(seg("send_dn").clone(), rank - 1, tag=2))
        reqs.append(dist.irecv(seg("recv_dn"), rank - 1, tag=1))
    for r in reqs:
        r.wait()
    assert not np.isnan(v).any(), "a ghost entry was not filled"
    assert np.array_equal(v, field), "ghost values differ from the owner's"

    # ---- compress(add): every rank adds 1000*(rank+1) on its top ghost plane, the owner accumulates it
    w = np.zeros(nloc)
    w[ob:ob + no] = 1.0
    tw = torch.from_numpy(w)
    reqs, buf = [], torch.zeros(pitch, dtype=torch.float64)
    if q["ghost_plane_top"]:
        o, n = q["recv_up"]
        tw[o:o + n] = 1000.0 * (rank + 1)
        reqs.append(dist.isend(tw[o:o + n].clone(), rank + 1, tag=3))
    if q["ghost_layers_below"]:
        reqs.append(dist.irecv(buf, rank - 1, tag=3))
    for r in reqs:
        r.wait()
    if q["ghost_layers_below"]:
        o, n = q["send_dn"]
        tw[o:o + n] += buf
        assert np.all(w[o:o + n] == 1.0 + 1000.0 * rank)
        assert np.all(w[o + n:ob + no] == 1.0)

    # ---- sums run over the owned ranges of all ranks: sum of the field = N(N+1)/2
    s = torch.tensor([field[ob:ob + no].sum()], dtype=torch.float64)
    dist.all_reduce(s)
    N = q["n_global"]
    assert s.item() == N * (N + 1) / 2, (s.item(), N)
    dist.barrier()
    print("partition worker ok rank %d/%d: %d local, %d owned of %d" % (rank, R, nloc, no, N), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
