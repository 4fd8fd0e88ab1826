"""Host logic of the z-slab decomposition over several GPUs (csrc/numbering.h through mgb_host_partition):
pure checks of the partition, and a world_size-2 `gloo` run on the CPU that carries out
update_ghost_values / compress(add) / the owned-range sums exactly by the library's exchange plan."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [  # p, n_subdivisions, level, n_ranks
    (4, (1, 1, 1), 3, 2), (4, (1, 1, 1), 3, 8), (4, (1, 1, 1), 2, 8), (2, (1, 2, 4), 1, 4), (3, (5, 5, 5), 2, 2),
    (3, (5, 5, 5), 1, 2), (1, (3, 3, 3), 3, 4), (4, (7, 7, 7), 2, 4), (4, (1, 1, 6), 0, 2), (5, (1, 1, 3), 2, 3),
]


@pytest.mark.parametrize("p,nsub,level,R", CASES)
def test_slabs_tile_the_level(mgb, p, nsub, level, R):
    parts = [mgb.host_partition(p, nsub, level, R, r) for r in range(R)]
    ncz = nsub[2] << level
    nd01 = ((nsub[0] << level) * p + 1) * ((nsub[1] << level) * p + 1)
    dist = parts[0]["distributed"]
    assert all(q["distributed"] == dist for q in parts)
    assert all(q["n_global"] == nd01 * (ncz * p + 1) for q in parts)
    if not dist:
        # replicated level: every rank holds everything, nothing to exchange
        for q in parts:
            assert q["n_local"] == q["n_global"] == q["n_owned"] and q["owned_begin"] == 0
            assert q["send_up"][1] == q["send_dn"][1] == q["recv_up"][1] == q["recv_dn"][1] == 0
        return
    assert ncz % R == 0
    # owned layers and owned planes tile the level without gaps or overlaps
    assert parts[0]["layer_begin"] == 0 and parts[-1]["layer_end"] == ncz
    planes = 0
    for r, q in enumerate(parts):
        assert q["layer_end"] - q["layer_begin"] == ncz // R
        if r:
            assert q["layer_begin"] == parts[r - 1]["layer_end"]
        assert q["ghost_layers_below"] == (1 if r else 0) and q["ghost_plane_top"] == (1 if r < R - 1 else 0)
        assert q["n_owned"] % nd01 == 0 and q["owned_begin"] == q["ghost_layers_below"] * p * nd01
        first_owned_plane = q["plane0"] + q["owned_begin"] // nd01
        assert first_owned_plane == planes
        planes += q["n_owned"] // nd01
        assert q["n_local"] == q["n_planes"] * nd01
        # exchange plan: what one rank sends is what its neighbour receives
        if r < R - 1:
            assert q["send_up"][1] == parts[r + 1]["recv_dn"][1] == p * nd01
            assert q["recv_up"][1] == parts[r + 1]["send_dn"][1] == nd01
            assert q["send_up"][0] + q["send_up"][1] == q["recv_up"][0] == q["n_local"] - nd01
        else:
            assert q["send_up"][1] == q["recv_up"][1] == 0
        if r:
            assert q["recv_dn"][0] == 0 and q["send_dn"][0] == q["owned_begin"]
    assert planes == ncz * p + 1


@pytest.mark.parametrize("p,nsub,R", [(4, (1, 1, 1), 2), (4, (1, 1, 1), 8), (3, (5, 5, 5), 2), (2, (3, 3, 3), 4),
                                      (4, (7, 7, 7), 4), (4, (1, 1, 3), 3)])
def test_slabs_nest_between_levels(mgb, p, nsub, R):
    """children of a rank's layers are its layers on the finer level; once distributed, always distributed"""
    was = False
    for level in range(0, 6):
        for r in range(R):
            q = mgb.host_partition(p, nsub, level, R, r)
            if q["distributed"]:
                f = mgb.host_partition(p, nsub, level + 1, R, r)
                assert f["distributed"]
                assert (f["layer_begin"], f["layer_end"]) == (2 * q["layer_begin"], 2 * q["layer_end"])
            elif r == 0 and was:
                raise AssertionError("a level above a distributed one is replicated")
        f = mgb.host_partition(p, nsub, level + 1, R, 0)
        if f["distributed"] and not mgb.host_partition(p, nsub, level, R, 0)["distributed"]:
            for r in range(R):  # first distributed level: the slabs must be cut between parent layers
                g = mgb.host_partition(p, nsub, level + 1, R, r)
                assert g["layer_begin"] % 2 == 0 and g["layer_end"] % 2 == 0
        was = mgb.host_partition(p, nsub, level, R, 0)["distributed"]


def test_partition_rejects_bad_rank(mgb):
    with pytest.raises(mgb.MgbError):
        mgb.host_partition(4, (1, 1, 1), 2, 2, 2)


@pytest.mark.parametrize("p,nsub,level", [(4, (1, 1, 1), 2), (3, (1, 2, 1), 2), (2, (3, 3, 2), 1)])
def test_ghost_exchange_gloo_world_size_2(mgb, p, nsub, level):
    """two CPU processes over gloo run the exchange plan on real arrays (tests/partition_worker.py)"""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + (os.getpid() + 7 * p + level) % 300), os.path.join(ROOT, "tests", "partition_worker.py"),
           str(p), ",".join(map(str, nsub)), str(level)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("partition worker ok") == 2, r.stdout[-2000:]
