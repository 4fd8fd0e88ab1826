"""CPU tests of the library's host logic: the C-ABI library loads and exports every symbol that
include/mgb200.h declares, the 1-D ingredients agree with the oracle's, and the patch-entity
numbering is a permutation with the right block structure.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(mgb):
    header = open(os.path.join(ROOT, "include", "mgb200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(mgb_[a-zA-Z0-9_]+)\s*\(", header))
    assert len(declared) >= 40
    L = C.CDLL(mgb.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(mgb.SYMBOLS), declared ^ set(mgb.SYMBOLS)
    assert mgb.lib().mgb_version() >= 100


def test_no_gpu_fails_loudly(mgb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mgb.MgbError, match="no CUDA device|CUDA"):
        mgb.MultigridSolver(2, 1, 1)


def test_error_reporting(mgb):
    with pytest.raises(mgb.MgbError, match="degree"):
        mgb.host_basis(11)


def test_dg_options_are_checked_before_the_device(mgb):
    """an unknown DG basis is an error of the options (return code + message), with or without a GPU; a valid basis
    without a GPU fails loudly - there is no CPU fallback"""
    with pytest.raises(mgb.MgbError, match="basis"):
        mgb.LaplaceOperatorDG(2, (1, 1, 1), 0, basis=7)
    with pytest.raises(mgb.MgbError, match="basis"):
        mgb.MultigridSolverDG(2, (1, 1, 1), 1, basis=-1)
    import torch
    if not torch.cuda.is_available():
        for kind in (mgb.DG_BASIS_GAUSS_LOBATTO, mgb.DG_BASIS_HERMITE, mgb.DG_BASIS_GAUSS):
            with pytest.raises(mgb.MgbError, match="no CUDA device|CUDA"):
                mgb.LaplaceOperatorDG(2, (1, 1, 1), 0, basis=kind)
    assert [mgb.dg_basis_of_type(t) for t in (0, 1, 2)] == [mgb.DG_BASIS_HERMITE, mgb.DG_BASIS_GAUSS_LOBATTO, mgb.DG_BASIS_GAUSS]


@pytest.mark.parametrize("p", range(1, 9))
def test_basis_matches_oracle(mgb, orc, p):
    a, b = mgb.host_basis(p), orc.basis(p)
    for k in ("S", "G", "D", "P1", "xq", "wq"):
        assert np.abs(a[k] - b[k]).max() < 1e-13, k
    assert np.abs(a["nodes"] - b["gll"]).max() < 1e-15
    # M1 = S^T W S, K1 = G^T W G; G = D S
    assert np.allclose(a["M1"], a["S"].T @ np.diag(a["wq"]) @ a["S"], atol=1e-14)
    assert np.allclose(a["K1"], a["G"].T @ np.diag(a["wq"]) @ a["G"], atol=1e-12)
    assert np.allclose(a["G"], a["D"] @ a["S"], atol=1e-11)
    assert abs(a["wq"].sum() - 1) < 1e-15
    assert np.allclose(a["K1"].sum(axis=0), 0, atol=1e-11)


@pytest.mark.parametrize("p,ncell,cw", [(1, (4, 4, 4), 0), (2, (4, 8, 4), 0), (4, (4, 4, 8), 0), (4, (2, 6, 2), 0),
                                        (3, (3, 5, 7), 0), (4, (8, 8, 8), 2), (8, (2, 2, 4), 0), (5, (4, 4, 4), 0),
                                        (4, (9, 5, 1), 0), (1, (40, 17, 3), 0)])
def test_numbering(mgb, p, ncell, cw):
    """tile numbering of csrc/level.h: a permutation; z-plane major over the level (a slab of planes is one
    contiguous range), inside a plane tile by tile, every tile one contiguous row-major chunk; Dirichlet
    DoFs = the boundary points"""
    r = mgb.host_numbering(p, ncell, cw)
    nd = [c * p + 1 for c in ncell]
    N = nd[0] * nd[1] * nd[2]
    assert r["n_dofs"] == N
    lex = r["lex_to_dof"].astype(np.int64)
    assert np.array_equal(np.sort(lex), np.arange(N))  # a permutation
    CW = r["tile_cells"]
    TW = CW * p
    assert (cw == 0 and 12 <= TW <= 16) or CW == cw
    ntx, nty = -(-nd[0] // TW), -(-nd[1] // TW)
    assert r["n_tiles"] == ntx * nty
    grid = lex.reshape(nd[2], nd[1], nd[0])
    base = 0
    for gk in (0, nd[2] // 2, nd[2] - 1):
        base = gk * nd[0] * nd[1]
        for ty in range(nty):
            for tx in range(ntx):
                wx, wy = min(TW, nd[0] - tx * TW), min(TW, nd[1] - ty * TW)
                chunk = grid[gk, ty * TW:ty * TW + wy, tx * TW:tx * TW + wx]
                assert np.array_equal(chunk.ravel(), base + np.arange(wy * wx))
                base += wy * wx
        assert base == (gk + 1) * nd[0] * nd[1]
    n_boundary = N - (nd[0] - 2) * (nd[1] - 2) * (nd[2] - 2)
    assert r["n_constrained"] == n_boundary


def test_numbering_rejects_bad_degree(mgb):
    with pytest.raises(mgb.MgbError):
        mgb.host_numbering(4, (1 << 9, 1 << 9, 1 << 9))  # > 2^32 DoFs on one device
