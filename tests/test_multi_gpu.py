"""Multi-GPU parity (z-slabs over the GPUs of one node, NCCL ghost exchange): one process per GPU under
torchrun, every rank checks its owned entries against the global CPU oracle (tests/mgpu_worker.py).
Skipped on boxes with fewer GPUs than ranks."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_ranks(n, args, timeout=600, p2p=False, worker="mgpu_worker.py", ok="mgpu worker ok"):
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)
    # the test meshes are far below the size from which the library cuts a level by default
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MGB200_MIN_SLAB_DOFS="0", MGB200_P2P="1" if p2p else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % n, "--master-addr",
           "127.0.0.1", "--master-port", str(29900 + os.getpid() % 90), os.path.join(ROOT, "tests", worker)] + args
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(ok) == n, r.stdout[-3000:]
    return r.stdout


@pytest.mark.parametrize("p,nsub,nref", [(4, "1,1,1", 3), (3, "1,2,3", 2), (2, "1,1,1", 4)])
def test_two_gpus_against_oracle(mgb, orc, p, nsub, nref):
    run_ranks(2, [str(p), nsub, str(nref)])


def test_two_gpus_peer_to_peer_exchange(mgb, orc):
    """the optional exchange over IPC-mapped peer memory (kernels store straight into the neighbour's buffers)"""
    out = run_ranks(2, ["4", "1,1,1", "3"], p2p=True)
    assert "exchanges" in out


def test_four_gpus_against_oracle(mgb, orc):
    run_ranks(4, ["4", "1,1,1", "3"])


def test_eight_gpus_against_oracle(mgb, orc):
    run_ranks(8, ["4", "1,1,1", "4"])


@pytest.mark.parametrize("p,nsub,nref", [(4, "1,1,1", 2), (2, "2,1,3", 1), (3, "1,2,1", 2)])
def test_two_gpus_dg_operator_against_oracle(mgb, orc, p, nsub, nref):
    """SIP-DG operator on z-slabs of cell layers, ghost layers over NCCL (the MPI face exchange of
    laplace_operator_dg.h:981-1058) against the global face-based oracle"""
    run_ranks(2, [str(p), nsub, str(nref)], worker="mgpu_dg_worker.py", ok="mgpu dg worker ok")


@pytest.mark.parametrize("basis", [1, 2])
def test_two_gpus_dg_operator_other_bases(mgb, orc, basis):
    """the same on FE_DGQHermite (1) and the Gauss basis (2): ghost cell layers carry coefficients of that basis"""
    run_ranks(2, ["4", "1,1,1", "2", str(basis)], worker="mgpu_dg_worker.py", ok="mgpu dg worker ok")


def test_four_gpus_dg_operator_against_oracle(mgb, orc):
    run_ranks(4, ["4", "1,1,1", "2"], worker="mgpu_dg_worker.py", ok="mgpu dg worker ok")


@pytest.mark.parametrize("p,nsub,nref", [(2, "1,1,1", 2), (3, "1,1,2", 2)])
def test_two_gpus_dg_multigrid_solver_against_oracle(mgb, orc, p, nsub, nref):
    """MultigridSolverDG (common/multigrid_solver_dg.h) with one process per GPU: same CG iteration count, residual
    history, L2 error and solution as the global oracle composition"""
    run_ranks(2, [str(p), nsub, str(nref)], worker="mgpu_dg_solver_worker.py", ok="mgpu dg solver worker ok")


@pytest.mark.parametrize("basis", [1, 2])
def test_two_gpus_dg_multigrid_solver_other_bases(mgb, orc, basis):
    """poisson_dg's own basis FE_DGQHermite (1) and the Gauss basis (2): the change of basis of the DG <-> FE_Q transfer
    runs on every local cell layer, ghost layers included"""
    run_ranks(2, ["2", "1,1,1", "2", str(basis)], worker="mgpu_dg_solver_worker.py", ok="mgpu dg solver worker ok")


def test_four_gpus_dg_multigrid_solver_against_oracle(mgb, orc):
    run_ranks(4, ["2", "1,1,1", "3"], worker="mgpu_dg_solver_worker.py", ok="mgpu dg solver worker ok")
