#!/usr/bin/env python
"""bench.py -- the reference's headline benchmark (poisson_cube, FE_Q<3>(4)) on B200.

A *step* is one pass of the hot path over the synthetic mesh: one `solve_cg()` of
common/multigrid_solver.h:483-493 -- double-precision CG to 1e-9 preconditioned by one float
V-cycle (Chebyshev/point-Jacobi smoother fused with the operator apply, matrix-free transfer).
`value` = DoFs solved per second with everything resident in HBM; `e2e` = the same through
mgb_solver_solve_cg_host() with the right-hand side coming from and the solution going back to
pinned HOST buffers inside the timed region.  The matvec DoFs/s figures of the metric
(poisson_cube/program.cc:343-380) are measured in the same run and reported under "matvec".

    python bench.py --gpus N --steps K --warmup W            (native arm)
    python bench.py --impl reference --gpus N --steps K ...  (reference arm: CPU restatement, host cores)

Only the `cpu_baseline` leg and `--impl reference` touch oracle/ (as the thing that is timed beside
the GPU path, never inside it).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DEGREE = 4
BASELINE_CG_DOFS_PER_S = 135005697 / 16.49       # README.md:159 (12-core Broadwell), BASELINE.md section 1
BASELINE_MATVEC_DOFS_PER_S = 135005697 / 0.15455  # README.md:122-127,159


def load_peaks():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def flops_per_dof(p):
    """reference op model (matvec_dg/program.cc:176-178) for the FE_Q cell."""
    n = p + 1
    ops1 = 2 * ((p + 1) // 2) * 2 + (p + 1) + 2 * (((p - 1) * (p + 1)) // 2)
    return (4 * 3 * ops1 * n * n + n ** 3 * 18) / p ** 3


def bytes_per_dof(p, nbytes, n_vectors):
    """n_vectors * N + the 27 compressed indices of 4 bytes per cell (laplace_operator.h:233-237)."""
    return n_vectors * nbytes + 4.0 * 27 / p ** 3


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        # samples under load: the upper half of the clock samples (idle samples sit at the floor)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU side: the oracle (oracle/mg_oracle.cpp, a CPU restatement of the reference's algorithm; the
# reference itself needs deal.II + p4est + MPI and cannot be built here: DESIGN.md) timed on the
# host cores with all OpenMP threads.
def cpu_sample_refinements(budget_s):
    cores = os.cpu_count() or 1
    try:
        mem_gb = os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES") / 2 ** 30
    except (ValueError, OSError):
        mem_gb = 16
    # 64^3 cells (17 M DoF) needs ~4 GB and ~2e6 DoF/s per 8 cores; otherwise 32^3 cells (2.1 M DoF)
    return 6 if (cores >= 16 and mem_gb >= 24 and budget_s >= 20) else 5


def run_cpu_solver(steps, warmup, n_refine, n_matvec=5):
    from oracle import orc
    orc.build()
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    s = orc.MultigridSolver(DEGREE, 1, n_refine, degree_pre=3, degree_post=3, n_cycles=2)
    n = s.n_dofs()
    its = 0
    for _ in range(warmup):
        its, _, _ = s.solve_cg()
    t0 = time.perf_counter()
    for _ in range(steps):
        its, _, _ = s.solve_cg()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    mv64 = s.time_matvec(True, n_matvec)
    mv32 = s.time_matvec(False, n_matvec)
    return dict(n_dofs=n, cells=8 ** n_refine, seconds_per_solve=dt, iterations=its, l2_error=s.compute_l2_error(),
                matvec_fp64_dofs_per_s=n / mv64, matvec_fp32_dofs_per_s=n / mv32, cores=cores)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nref = cpu_sample_refinements(60)
    r = run_cpu_solver(args.steps, args.warmup, nref)
    value = r["n_dofs"] / r["seconds_per_solve"]
    sample = ("solve_cg on %d^3 cells = %d DoF (same FE_Q(4) cube problem, smaller mesh; multigrid cost is linear "
              "in DoFs), oracle/mg_oracle.cpp with %d OpenMP threads" % (round(r["cells"] ** (1 / 3)), r["n_dofs"], r["cores"]))
    line = {
        "impl": "reference", "metric": "cg_vcycle_solve_dofs_per_s", "value": value, "unit": "DoF/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds_per_solve"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": value / BASELINE_CG_DOFS_PER_S, "dtype": "f64 outer CG / f32 V-cycle", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": "DoF/s", "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "DoF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cg_iterations": r["iterations"], "l2_error": r["l2_error"],
        "matvec": {"fp64_dofs_per_s": r["matvec_fp64_dofs_per_s"], "fp32_dofs_per_s": r["matvec_fp32_dofs_per_s"]},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def weak_scaling_mesh(cells, n_gpus):
    """The reference's weak-scaling series (poisson_cube/program.cc:508-527, `use_doubling_mesh`): the box is
    doubled direction by direction, 128^3 -> 128x128x256 -> 128x256x256 -> 256^3 cells, here z first so that
    every GPU's z-slab is 128^3 cells' worth of work.  Returns (n_subdivisions, n_refinements, lower, upper)."""
    nref = cells.bit_length() - 1
    assert 1 << nref == cells, "--cells must be a power of two"
    if n_gpus == 1:
        return (1, 1, 1), nref, (-0.9,) * 3, (1.0,) * 3
    if n_gpus == 8:  # the cube again (sizes[] = 256: n_subdiv 1, one more refinement)
        return (1, 1, 1), nref + 1, (-0.9,) * 3, (1.0,) * 3
    assert n_gpus in (2, 4), "--gpus must be 1, 2, 4 or 8"
    nsub = (1, 1, 2) if n_gpus == 2 else (1, 2, 2)
    return nsub, nref, (-1.0,) * 3, tuple(0.9 + 1.9 * (k - 1) for k in nsub)


def workload_config(args, n_gpus):
    cells = args.cells
    nsub, nref, lo, hi = weak_scaling_mesh(cells, n_gpus)
    shape = [k << nref for k in nsub]
    dofs = 1
    for c in shape:
        dofs *= DEGREE * c + 1
    return {"workload": "poisson_cube FE_Q<3>(%d), %dx%dx%d cells = %d DoF on %d GPU(s) (%d^3 cells per GPU), box %s..%s, "
                        "2 V-cycles/FMG level, Chebyshev 2+2 (deal.II >= 9.5 degree 3), CG to 1e-9"
                        % (DEGREE, shape[0], shape[1], shape[2], dofs, n_gpus, cells, list(lo), list(hi)),
            "degree": DEGREE, "cells": shape, "cells_per_gpu": cells ** 3, "total_dofs": dofs,
            "levels": nref + 1,
            "parallelism": "1 GPU" if n_gpus == 1 else
            "%d z-slabs of whole cell layers, one per GPU; NCCL ghost-plane exchange per operator apply, "
            "scalar all-reduces in CG; coarse levels replicated" % n_gpus,
            "l2_flush": "inputs larger than L2 (every finest-level vector is >= 540 MB per GPU)"}


# ------------------------------------------------------------------------------------------------
def native_arm(args):
    import numpy as np
    import torch
    import __graft_entry__ as entry
    entry.build_library()
    import multigrid_b200 as mgb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    nsub, nref, lo, hi = weak_scaling_mesh(args.cells, world)
    comm = mgb.comm_from_torch_distributed() if dist is not None else None
    # CUDA context creation and the first touch of the device (seconds on a fresh box) are not part of the solver set-up
    torch.zeros(1, device="cuda")
    torch.cuda.synchronize()
    t_setup = time.perf_counter()
    s = mgb.MultigridSolver(DEGREE, nsub, nref, lower=lo, upper=hi, degree_pre=3, degree_post=3, n_cycles=2,
                            device=local_rank, comm=comm)
    s.sync()
    t_setup = time.perf_counter() - t_setup
    n = s.n_dofs()  # local vector length (owned + ghost planes)
    L = s.n_levels - 1
    part = s.partition(L)
    hbm_gbs, peak_src = load_peaks()

    # ---- device-resident solve: W warm-up steps, K timed steps
    its = 0
    for _ in range(args.warmup):
        its, rate = s.solve_cg()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    s.profile(True)
    s.class_times(L, reset=True)
    launches0 = s.launch_count()
    barrier()
    s.timer_start()
    for _ in range(args.steps):
        its, rate = s.solve_cg()
    ms = s.timer_stop()
    barrier()
    launches = s.launch_count() - launches0
    exchanges = s.exchange_count()
    classes = s.class_times(L, reset=True)
    s.profile(False)
    ms_step = max_over_ranks(ms / args.steps)
    l2_error = s.compute_l2_error()

    # ---- end to end: right-hand side from pinned host memory, solution back to pinned host memory
    rhs_host = torch.empty(n, dtype=torch.float64).pin_memory()
    sol_host = torch.empty(n, dtype=torch.float64).pin_memory()
    rhs_ptr, _ = s.vector_ptr("rhs", L)
    mgb._check(mgb.lib().mgb_copy_to_host(s.h, rhs_host.data_ptr(), rhs_ptr, n * 8))
    s.solve_cg_host(rhs_host, sol_host)
    barrier()
    s.timer_start()
    for _ in range(args.steps):
        s.solve_cg_host(rhs_host, sol_host)
    ms_e2e = max_over_ranks(s.timer_stop() / args.steps)
    barrier()

    # ---- matvec DoFs/s, protocol of poisson_cube/program.cc:343-380 (min over repeats of the mean)
    n_mv = 50 if n >= 1e7 else 200
    mv = {}
    for name, fn in (("fp64", s.do_matvec), ("fp32", s.do_matvec_smoother)):
        best = 1e30
        for _ in range(3):
            barrier()
            s.timer_start()
            for _ in range(n_mv):
                fn()
            best = min(best, max_over_ranks(s.timer_stop() / n_mv))
        mv[name] = best
    clocks = sampler.stop() if rank == 0 else None
    return dict(s=s, n=n, L=L, its=its, rate=rate, ms_step=ms_step, ms_e2e=ms_e2e, mv=mv, classes=classes,
                launches=launches, l2_error=l2_error, clocks=clocks, t_setup=t_setup, hbm_gbs=hbm_gbs,
                peak_src=peak_src, rank=rank, world=world, dist=dist, part=part, exchanges=exchanges)


def report(args, r):
    world = r["world"]
    n = r["part"]["n_owned"]             # DoFs this GPU owns (what its kernels produce per launch)
    total_dofs = r["part"]["n_global"]   # one global problem, cut into z-slabs
    value = total_dofs / (1e-3 * r["ms_step"])
    e2e_value = total_dofs / (1e-3 * r["ms_e2e"])
    hbm = r["hbm_gbs"]
    cls = r["classes"]

    def kernel(name, n_vectors, nbytes):
        ms, cnt = cls[name]
        if cnt == 0:
            return None
        per_launch_ms = ms / cnt
        alg_bytes = bytes_per_dof(DEGREE, nbytes, n_vectors) * n
        achieved = alg_bytes / (1e-3 * per_launch_ms) / 1e9
        return {"launches": cnt, "ms_per_launch": per_launch_ms, "share_of_step": ms / (r["ms_step"] * args.steps),
                "algorithmic_bytes_per_dof": bytes_per_dof(DEGREE, nbytes, n_vectors), "achieved_gbs": achieved,
                "frac_hbm": achieved / hbm, "dofs_per_s": n / (1e-3 * per_launch_ms)}

    kernels = {"apply_f32_chebyshev": kernel("apply_f32_chebyshev", 5, 4),
               "apply_f32_residual": kernel("apply_f32", 3, 4),
               "apply_f64": kernel("apply_f64", 2, 8)}
    epi_ms, epi_cnt = cls["apply_epilogue"]
    dom = kernels["apply_f32_chebyshev"]
    # dram__bytes_read.sum + dram__bytes_write.sum of one finest-level launch on 128^3 cells per GPU, from the
    # `ncu --set full` capture summarised in profiles/r1_tile_f32x2_ncu_full_summary.txt (2.215 GB + 0.523 GB)
    traffic = 2.215254e9 + 0.522742e9 if n == 135005697 else None
    roofline = {"bound": "hbm", "kernel": "k_apply_tile_f32x2<4, APPLY_CHEBYSHEV> (float operator apply with fused "
                                          "Chebyshev update, finest level)",
                "achieved": dom["achieved_gbs"], "peak": hbm, "unit": "GB/s", "frac": dom["frac_hbm"],
                "traffic": traffic, "traffic_source": "profiles/r1_tile_f32x2_ncu_full_summary.txt",
                "peak_source": r["peak_src"], "ms_per_launch": dom["ms_per_launch"],
                "share_of_step": dom["share_of_step"],
                "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_dof"] * n}
    fl = flops_per_dof(DEGREE)
    matvec = {}
    for name, nbytes, peak_tf in (("fp64", 8, 37.2), ("fp32", 4, 74.4)):
        sec = 1e-3 * r["mv"][name]
        dps = total_dofs / sec
        gbs = bytes_per_dof(DEGREE, nbytes, 2) * n / sec / 1e9
        matvec[name] = {"dofs_per_s": dps, "ms": r["mv"][name], "achieved_gbs_per_gpu": gbs, "frac_hbm": gbs / hbm,
                        "tflops_per_gpu": fl * n / sec / 1e12, "frac_cuda_core_peak": fl * n / sec / 1e12 / peak_tf}
    matvec["fp64"]["vs_readme_12core_broadwell"] = matvec["fp64"]["dofs_per_s"] / BASELINE_MATVEC_DOFS_PER_S
    line = {
        "metric": "cg_vcycle_solve_dofs_per_s", "value": value, "unit": "DoF/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": value / BASELINE_CG_DOFS_PER_S,
        "dtype": "f64 outer CG / f32 V-cycle", "data": "synthetic", "config": workload_config(args, world),
        "e2e": {"value": e2e_value, "unit": "DoF/s", "ms_per_step": r["ms_e2e"], "h2d_bytes_per_step": 8 * r["n"] * world,
                "d2h_bytes_per_step": 8 * r["n"] * world, "api": "mgb_solver_solve_cg_host (pinned host rhs in, solution out)"},
        "exchanges_total": r["exchanges"],
        "gpu_launches": r["launches"], "cg_iterations": r["its"], "cg_reduction_per_iteration": r["rate"],
        "l2_error": r["l2_error"], "total_dofs": total_dofs, "setup_s": r["t_setup"],
        "matvec": matvec, "roofline": roofline, "kernels_finest_level": kernels,
        "epilogue_kernel_ms_per_launch": epi_ms / epi_cnt if epi_cnt else None,
        "clocks": r["clocks"],
    }
    return line


def dg_lines(r):
    """BASELINE config 4 (secondary lines, one GPU): FE_DGQ<3>(4) SIP Laplacian on the 131 M DoF mesh of
    matvec_dg/program.cc (s = 20), double vmult with the reference's 3 N bytes/DoF model (:204) and the float
    fused Chebyshev update of matvec_dg_cheby/program.cc (5 N bytes/DoF, :178)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import time_dg
    r["s"].close()          # free the FE_Q solver's vectors first
    torch.cuda.empty_cache()
    out = time_dg.run(20, DEGREE, reps=10)
    for v in out.values():
        v["frac_hbm"] = v["gbs"] / r["hbm_gbs"]
    out["workload"] = "matvec_dg s=20: 128x128x64 cells, FE_DGQ(4), affine sheared mesh, 131072000 DoF"
    # the same protocol on FE_DGQArbitraryNodes(QGauss) (matvec_dg/program.cc:48-49, type 2): collocation kernel variant
    gauss = time_dg.run(20, DEGREE, reps=10, basis=2)
    for v in gauss.values():
        v["frac_hbm"] = v["gbs"] / r["hbm_gbs"]
    out["gauss_basis"] = gauss
    # poisson_dg/program.cc:227-263 on its doubling-mesh cycle 21 (128^3 cells, 262 M DG DoF), default tolerance 1e-3:
    # MultigridSolverDG = CG(double) + Chebyshev/JacobiTransformed + FE_Q V-cycle (float), minimum of 4 solves
    import time_dg_solver
    torch.cuda.empty_cache()
    out["poisson_dg_solve"] = time_dg_solver.run(21, DEGREE, 1e-3)
    out["poisson_dg_solve"]["workload"] = "poisson_dg cycle 21: [-1,1]^3, 128^3 cells, FE_DGQ(4), solve_cg(1e-3)"
    # the same solve on the reference's own basis, FE_DGQHermite (poisson_dg/program.cc:174): vectors are Hermite
    # coefficients, the DG <-> FE_Q transfer carries the change of basis
    torch.cuda.empty_cache()
    out["poisson_dg_solve_hermite"] = time_dg_solver.run(21, DEGREE, 1e-3, basis=1)
    out["poisson_dg_solve_hermite"]["workload"] = "poisson_dg cycle 21: [-1,1]^3, 128^3 cells, FE_DGQHermite(4), solve_cg(1e-3)"
    return out


def varcoef_lines(r):
    """BASELINE config 3's streaming pattern (secondary lines, one GPU): FE_Q<3>(4) operator with a merged coefficient
    per quadrature point (laplace_operator.h:388-430,493-522) and the coefficient of poisson_shell/program.cc:97-118,
    on the 64^3-cell box (17 M DoF, 1.6 GB coefficient stream); 111.4 / 56.6 bytes per DoF."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import time_varcoef
    out = time_varcoef.run(64, DEGREE, reps=10)
    for k in ("fp64", "fp32"):
        out[k]["frac_hbm"] = out[k]["gbs"] / r["hbm_gbs"]
    out["workload"] = "FE_Q(4) variable-coefficient matvec, 64^3 cells, per-q-point merged coefficient streamed once"
    return out


def adaptive_lines(r):
    """BASELINE config 5's access pattern (secondary lines, one GPU): FE_Q(3) on hyper_L refined towards the re-entrant
    edge, hanging-node and Dirichlet constraints applied in the kernel (poisson_l/program.cc:232-260)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import time_adaptive
    out = time_adaptive.run(3, 4, 3, reps=20)
    for k in ("vmult_f64", "vmult_f32"):
        out[k]["frac_hbm"] = out[k]["gbs"] / r["hbm_gbs"]
    out["workload"] = "FE_Q(3) on hyper_L, 4 global refinements + 3 rounds towards the re-entrant edge, hanging nodes in-kernel"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cells", type=int, default=128, help="cells per direction and GPU (power of two)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dg", action="store_true", help="skip the DG operator lines (BASELINE config 4)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    if args.impl == "reference":
        reference_arm(args)
        return
    r = native_arm(args)
    if r["rank"] == 0:
        line = report(args, r)
        if r["world"] == 1 and not args.no_cpu_baseline:
            c = run_cpu_solver(2, 1, 5)
            line["cpu_baseline"] = {
                "value": c["n_dofs"] / c["seconds_per_solve"], "unit": "DoF/s", "cores": c["cores"], "kind": "port",
                "sample": "solve_cg on 32^3 cells = %d DoF, 2 timed solves, oracle/mg_oracle.cpp, %d OpenMP threads"
                          % (c["n_dofs"], c["cores"]),
                "matvec_fp64_dofs_per_s": c["matvec_fp64_dofs_per_s"], "cg_iterations": c["iterations"]}
        if r["world"] == 1 and not args.no_dg:
            line["dg"] = dg_lines(r)
            line["variable_coefficient"] = varcoef_lines(r)
            line["adaptive"] = adaptive_lines(r)
        print(json.dumps(line), flush=True)
    if r["dist"] is not None:
        r["dist"].barrier()
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
