"""The C++ header adapters (include/multigrid/*.h: the reference's class and method names over the C ABI) compile
with plain g++ against libmgb200.so; on the GPU the example driver reproduces the oracle's numbers."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "adapter_example")


def build_example(mgb):
    mgb.lib()
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    libdir = os.path.join(ROOT, "multigrid_b200")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "adapter_example.cc"), "-o", EXE, "-L", libdir, "-lmgb200",
           "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_adapters_compile_and_link(mgb):
    build_example(mgb)
    assert os.path.exists(EXE)


def test_header_is_plain_c(mgb):
    """include/mgb200.h must be consumable from C (the drop-in boundary has no C++ types)"""
    src = os.path.join(ROOT, "tests", "cpp", "_build", "c_check.c")
    os.makedirs(os.path.dirname(src), exist_ok=True)
    open(src, "w").write('#include "mgb200.h"\nint main(void) { mgb_options o; mgb_default_options(&o); return o.degree == 4 ? 0 : 1; }\n')
    libdir = os.path.join(ROOT, "multigrid_b200")
    exe = src[:-2]
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), src, "-o", exe, "-L", libdir,
                        "-lmgb200", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert subprocess.run([exe]).returncode == 0   # host-only call, no GPU needed


@pytest.mark.gpu
def test_example_driver_matches_oracle(mgb, orc):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    build_example(mgb)
    r = subprocess.run([EXE, "3"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    out = dict(re.findall(r"(\w+) ([-+0-9.e]+)", r.stdout))
    o = orc.MultigridSolver(4, 1, 3, degree_pre=3, degree_post=3)
    rate_o = o.solve(False)
    e_fmg_o = o.compute_l2_error()
    its_o, red_o, _ = o.solve_cg()
    assert int(out["cg_its"]) == its_o
    assert abs(float(out["fmg_l2"]) - e_fmg_o) <= 5e-3 * e_fmg_o
    assert abs(float(out["cg_l2"]) - o.compute_l2_error()) <= 1e-3 * o.compute_l2_error()
    assert abs(float(out["fmg_reduction"]) - rate_o) <= 2e-2 * rate_o
    ones = np.ones(o.n_dofs())
    assert abs(float(out["pAp"]) - ones @ o.vmult(ones)) <= 1e-10 * abs(ones @ o.vmult(ones))
    assert float(out["residual_norm"]) <= 1e-10
    # operator concept extras: el, converting assignment, interface matrices, compute_residual with a caller's function
    inv = o.inverse_diagonal_dp()
    assert abs(float(out["diag_inv_norm"]) - np.linalg.norm(inv)) <= 1e-10 * np.linalg.norm(inv)
    assert float(out["convert_roundtrip"]) == 0.0 and float(out["interface_down_norm"]) == 0.0
    assert float(out["el_mid"]) > 0
    o.set_problem(right_hand_side=lambda x, y, z: 1. + x * y - z)
    load = o.get_vector(orc.VEC_RHS).sum()
    assert abs(float(out["load_sum"]) - load) <= 1e-10 * abs(load)
    # the third problem of the example (defined only in tests/cpp/adapter_example.cc) against the oracle with the same
    # functions handed over as samples
    u3 = lambda x, y, z: np.exp(x) * np.sin(2 * y) * np.cos(z)
    c3 = lambda x, y, z: 1 + x * x + 0.5 * np.sin(3 * y)

    def f3(x, y, z):
        ux, uy = u3(x, y, z), 2 * np.exp(x) * np.cos(2 * y) * np.cos(z)
        return -(c3(x, y, z) * (ux - 5 * u3(x, y, z)) + 2 * x * ux + 1.5 * np.cos(3 * y) * uy)

    o3 = orc.MultigridSolver(4, 1, 2, degree_pre=3, degree_post=3, geometry=1)
    o3.set_problem(c3, f3, u3, u3)
    its3, rate3, _ = o3.solve_cg()
    assert int(out["third_cg_its"]) == its3
    assert abs(float(out["third_cg_rate"]) - rate3) <= 0.05 * rate3
    assert abs(float(out["third_l2"]) - o3.compute_l2_error()) <= 1e-3 * o3.compute_l2_error()
    d = orc.DGOperator(4, (4, 4, 4))
    x = np.ones(d.n_dofs())
    assert abs(float(out["xAx"]) - x @ d.vmult(x)) <= 1e-10 * abs(x @ d.vmult(x))
    assert abs(float(out["penalty_x"]) - d.penalty()[0]) <= 1e-6 * d.penalty()[0]  # printed with 7 digits
    for name, kind in (("xAx_hermite", orc.DG_BASIS_HERMITE), ("xAx_gauss", orc.DG_BASIS_GAUSS)):
        dk = orc.DGOperator(4, (4, 4, 4), basis=kind)
        z = 0.1 + (np.arange(dk.n_dofs()) % 7) / 7.
        want = z @ dk.vmult(z)
        assert abs(float(out[name]) - want) <= 1e-10 * abs(want), name
    jt = orc.JacobiTransformed(d)
    assert float(out["xPx"]) > 0 and abs(float(out["xPx"]) - x @ jt.vmult(x)) <= 1e-9 * abs(x @ jt.vmult(x))
    # the adaptive operator of the example: same mesh through the Python mirror, energy against the geometric oracle
    from oracle import adaptive_oracle
    mask = np.ones((2, 2, 2), dtype=np.uint8)
    mask[1, 1, 1] = 0
    a = mgb.LaplaceOperatorAdaptive(2, (2, 2, 2), mask, n_global_refinements=1)
    lo, sz = a.cell_boxes()
    a.refine(np.hypot(*(lo + 0.5 * sz)[:, :2].T) < 0.55)
    lo, sz = a.cell_boxes()
    oa = adaptive_oracle.AdaptiveOracle(2, lo, sz)
    assert int(out["adaptive_cells"]) == a.n_cells() and int(out["adaptive_m"]) == oa.n_dofs()
    e_ref = float(np.ones(oa.n_dofs()) @ oa.vmult(np.ones(oa.n_dofs())))
    assert abs(float(out["adaptive_xAx"]) - e_ref) <= 1e-10 * abs(e_ref)
    # the local-smoothing solve through the adapter: A x = A 1 (identity rows on the boundary), so x = 1
    assert int(out["adaptive_levels"]) == 3 and 0 < int(out["adaptive_cg_its"]) <= 14
    assert float(out["adaptive_cg_rate"]) < 0.25 and float(out["adaptive_solution_error"]) <= 1e-7
    ds = orc.MultigridSolverDG(4, (1, 1, 1), 2, basis=orc.DG_BASIS_HERMITE)   # the adapter's default, as the reference
    _, rate_dg, _ = ds.solve_cg(1e-3)
    assert int(out["dg_cg_its"]) == ds.last_step
    assert abs(float(out["dg_cg_rate"]) - rate_dg) <= 3e-2 * rate_dg
    assert abs(float(out["dg_l2"]) - ds.compute_l2_error()) <= 1e-3 * ds.compute_l2_error()
    dp = orc.MultigridSolverDGPlain(4, (1, 1, 1), 2, basis=orc.DG_BASIS_GAUSS)
    _, rate_p, _ = dp.solve_cg(1e-3)
    assert int(out["plain_levels"]) == 3 and int(out["plain_cg_its"]) == dp.last_step
    assert abs(float(out["plain_cg_rate"]) - rate_p) <= 3e-2 * rate_p
    assert abs(float(out["plain_l2"]) - dp.compute_l2_error()) <= 1e-3 * dp.compute_l2_error()
    ones = np.ones(dp.n_dofs())
    rMr = ones @ dp.vmult(ones)
    assert abs(float(out["plain_rMr"]) - rMr) <= 1e-4 * rMr
    assert abs(float(out["plain_rMr"]) - float(out["plain_rMr_check"])) <= 1e-9 * rMr
    scherk = lambda x, y, z: np.log(np.cos(x) / np.cos(y)) + 0 * z
    om = orc.MultigridSolver(3, 1, 2, lower=(-1,) * 3, upper=(1,) * 3, degree_pre=2, degree_post=2, n_cycles=1)
    om.minimal_surface_set_boundary_values(scherk)
    steps = om.minimal_surface_solve(20, 1e-11)
    xs = om.support_points()
    err = np.abs(om.get_vector(orc.VEC_SOLUTION) - scherk(xs[:, 0], xs[:, 1], xs[:, 2])).max()
    assert int(out["minimal_steps"]) == len(steps) and float(out["minimal_residual"]) < 1e-11
    assert abs(int(out["minimal_linear_its"]) - sum(st["linear_iterations"] for st in steps)) <= 1
    assert abs(float(out["minimal_error"]) - err) <= 1e-3 * err
    # the adaptive loop through the adapter against the Python mirror's loop on the same forest
    from adaptive_cases import L_EXTRUDED
    lp = mgb.LaplaceOperatorAdaptive(2, (2, 2, 1), L_EXTRUDED, (-1., -1., 0.), (1., 1., 1.), 1)
    table = lp.run_adaptive_loop(4, function=lp.FUNCTION_L_SINGULARITY, f=0., n_smoother=3)
    for c, t in enumerate(table):
        assert int(out["loop%d_cells" % c]) == t["n_cells"] and int(out["loop%d_dofs" % c]) == t["n_dofs"]
        assert int(out["loop%d_its" % c]) == t["iterations"]
        assert abs(float(out["loop%d_h1" % c]) - t["h1_error"]) <= 1e-5 * t["h1_error"]
        assert abs(float(out["loop%d_l2" % c]) - t["l2_error"]) <= 1e-4 * t["l2_error"]
    assert float(out["loop3_h1"]) < 0.45 * float(out["loop0_h1"])
