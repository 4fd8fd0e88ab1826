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


# L2 error of the CPU oracle's solve_cg on the same mesh (run once in the build container, 8 iterations)
ORACLE_L2 = {135005697: 4.196178e-10}


def readme_row(n_dofs):
    """the reference's published row for this mesh (tests/golden/readme_table.json = README.md:135-159)"""
    try:
        tab = json.load(open(os.path.join(ROOT, "tests", "golden", "readme_table.json")))
    except OSError:
        return None
    for row in tab["rows"]:
        d = dict(zip(tab["_columns"], row))
        if d["dofs"] == n_dofs:
            return {"cg_its": d["cg_its"], "l2_cg": d["cg_L2error"], "cg_reduction": d["cg_reduction"]}
    return None


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


def bytes_per_dof(p, nbytes, n_vectors, index_table=False):
    """Algorithmic bytes per DoF of one operator apply: n_vectors * N for the tile kernels (closed-form
    numbering, no index table: DESIGN.md section 4); with `index_table` the reference's model of SURVEY.md
    8(d), which adds the 27 compressed indices of 4 bytes per cell (laplace_operator.h:233-237)."""
    return n_vectors * nbytes + (4.0 * 27 / p ** 3 if index_table else 0.0)


def executed_flops_per_dof(p):
    """Flops the Cartesian tile kernel executes per DoF (FMA = 2): seven 1-D products with the assembled
    grid-line matrices (x: M, K on the haloed rows; y: My mx, Ky mx, My kx; z: Mz Q, Kz P per cell layer),
    DESIGN.md section 4.1.  p = 4: about 102."""
    n = p + 1
    cw = {1: 16, 2: 8, 3: 5, 4: 4, 5: 3}.get(p, 2)
    tw = cw * p
    row = ((p - 1) * n + (2 * n - 1)) / p          # multiply-adds per point of one assembled 1-D product
    halo = (tw + p + 1) / tw
    return 2 * (2 * row * halo + 3 * row + 2 + 2 * n * n / p + 1)


def ncu_traffic(summary_file, kernel_substring):
    """dram__bytes_read.sum + dram__bytes_write.sum of the first launch of `kernel_substring` in a summary
    written by tools/ncu_summary.py (profiles/); None if the file or the kernel is not there."""
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        lines = open(os.path.join(ROOT, summary_file)).read().splitlines()
    except OSError:
        return None
    inside, total, found = False, 0.0, 0
    for ln in lines:
        f = ln.split()
        if len(f) >= 2 and f[0] == "Kernel" and f[1] == "Name":
            if found == 2:
                break
            inside = kernel_substring in ln
            continue
        if inside and f and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and len(f) >= 3:
            total += float(f[1]) * scale.get(f[2], 1.0)
            found += 1
    return total if found == 2 else None


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
def mem_available_gb():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) / 2 ** 20
    except OSError:
        pass
    return 0.0


def cpu_sample_refinements(full_size):
    """Refinements of the cube the CPU arm solves: 7 (128^3 cells, 135 M DoF = the native arm's mesh, 11 GB and about
    20 s per solve on 32 cores) when asked for the full size and the box has the memory and cores for it, else 6
    (64^3 cells, 17 M DoF), else 5."""
    cores, mem = os.cpu_count() or 1, mem_available_gb()
    if full_size and cores >= 16 and mem >= 32:
        return 7
    return 6 if (cores >= 16 and mem >= 24) else 5


def run_cpu_solver(steps, warmup, n_refine, n_matvec=5, budget_s=None):
    from oracle import orc
    orc.build()
    # torch.distributed.run exports OMP_NUM_THREADS=1: set the thread count in the OpenMP runtime itself and
    # report what the runtime really uses
    threads = orc.omp_threads(os.cpu_count() or 1)
    s = orc.MultigridSolver(DEGREE, 1, n_refine, degree_pre=3, degree_post=3, n_cycles=2)
    n = s.n_dofs()
    its, t_warm = 0, None
    for _ in range(warmup):
        t0 = time.perf_counter()
        its, _, _ = s.solve_cg()
        t_warm = time.perf_counter() - t0
    timed = steps
    if budget_s is not None and t_warm is not None:  # bounded sample: as many of the K steps as fit the budget
        timed = max(1, min(steps, int(budget_s / max(t_warm, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(timed):
        its, _, _ = s.solve_cg()
    dt = (time.perf_counter() - t0) / max(timed, 1)
    mv64 = s.time_matvec(True, n_matvec)
    mv32 = s.time_matvec(False, n_matvec)
    return dict(n_dofs=n, cells=8 ** n_refine, seconds_per_solve=dt, iterations=its, l2_error=s.compute_l2_error(),
                matvec_fp64_dofs_per_s=n / mv64, matvec_fp32_dofs_per_s=n / mv32, cores=threads, steps_timed=timed)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nref = cpu_sample_refinements(args.cells >= 128)
    r = run_cpu_solver(args.steps, min(args.warmup, 1) if nref >= 7 else args.warmup, nref,
                       budget_s=400 if nref >= 7 else 240)
    value = r["n_dofs"] / r["seconds_per_solve"]
    nc = round(r["cells"] ** (1 / 3))
    sample = ("solve_cg on %d^3 cells = %d DoF (%s), %d of the %d steps timed, oracle/mg_oracle.cpp with %d OpenMP "
              "threads (omp_get_max_threads)"
              % (nc, r["n_dofs"], "the native arm's mesh" if nc == args.cells else
                 "same FE_Q(4) cube problem on a smaller mesh: the box lacks the memory or cores for 128^3; multigrid "
                 "cost is linear in DoFs", r["steps_timed"], args.steps, r["cores"]))
    cfg = workload_config(args, 1, cells=nc)
    cfg["same_mesh_as_native_arm"] = nc == args.cells
    line = {
        "impl": "reference", "metric": "cg_vcycle_solve_dofs_per_s", "value": value, "unit": "DoF/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds_per_solve"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": value / BASELINE_CG_DOFS_PER_S, "dtype": "f64 outer CG / f32 V-cycle", "data": "synthetic",
        "config": cfg, "steps_timed": r["steps_timed"],
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


def workload_config(args, n_gpus, cells=None):
    cells = args.cells if cells is None else cells
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
        bpd = bytes_per_dof(DEGREE, nbytes, n_vectors)
        bpd_ref = bytes_per_dof(DEGREE, nbytes, n_vectors, index_table=True)
        achieved = bpd * n / (1e-3 * per_launch_ms) / 1e9
        return {"launches": cnt, "ms_per_launch": per_launch_ms, "share_of_step": ms / (r["ms_step"] * args.steps),
                "algorithmic_bytes_per_dof": bpd, "achieved_gbs": achieved, "frac_hbm": achieved / hbm,
                "survey_8d_bytes_per_dof_with_index_table": bpd_ref,
                "frac_hbm_survey_8d_model": achieved / hbm * bpd_ref / bpd,
                "dofs_per_s": n / (1e-3 * per_launch_ms)}

    kernels = {"apply_f32_chebyshev": kernel("apply_f32_chebyshev", 5, 4),
               "apply_f32_residual": kernel("apply_f32", 3, 4),
               "apply_f64": kernel("apply_f64", 2, 8)}
    epi_ms, epi_cnt = cls["apply_epilogue"]
    dom = kernels["apply_f32_chebyshev"]
    # dram__bytes_read.sum + dram__bytes_write.sum of one finest-level launch on 128^3 cells per GPU, read from the
    # committed `ncu --set full` summary of this kernel (newest round first)
    traffic, traffic_src = None, None
    if n == 135005697:
        for f in ("profiles/r2_tile_f32x2_ncu_full_summary.txt", "profiles/r1_tile_f32x2_ncu_full_summary.txt"):
            traffic = ncu_traffic(f, "k_apply_tile_f32x2<4, 3>")
            if traffic is not None:
                traffic_src = f
                break
    roofline = {"bound": "hbm", "kernel": "k_apply_tile_f32x2<4, APPLY_CHEBYSHEV> (float operator apply with fused "
                                          "Chebyshev update, finest level)",
                "achieved": dom["achieved_gbs"], "peak": hbm, "unit": "GB/s", "frac": dom["frac_hbm"],
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": r["peak_src"], "ms_per_launch": dom["ms_per_launch"],
                "share_of_step": dom["share_of_step"],
                "algorithmic_bytes_per_dof": dom["algorithmic_bytes_per_dof"],
                "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_dof"] * n,
                "bytes_model": "5 vectors x 4 B = 20 B/DoF (x_k, b, D^-1, x_k-1 read, x_k+1 written; the tile kernel reads "
                               "no index table); SURVEY 8(d)'s 21.69 B/DoF variant, which bills the reference's 27 "
                               "indices per cell, is kept as frac_survey_8d_model",
                "frac_survey_8d_model": dom["frac_hbm_survey_8d_model"]}
    fl_ref, fl_exec = flops_per_dof(DEGREE), executed_flops_per_dof(DEGREE)
    matvec = {}
    for name, nbytes, peak_tf in (("fp64", 8, 37.2), ("fp32", 4, 74.4)):
        sec = 1e-3 * r["mv"][name]
        dps = total_dofs / sec
        gbs = bytes_per_dof(DEGREE, nbytes, 2) * n / sec / 1e9
        gbs_ref = bytes_per_dof(DEGREE, nbytes, 2, index_table=True) * n / sec / 1e9
        matvec[name] = {"dofs_per_s": dps, "ms": r["mv"][name], "algorithmic_bytes_per_dof": 2 * nbytes,
                        "achieved_gbs_per_gpu": gbs, "frac_hbm": gbs / hbm,
                        "frac_hbm_survey_8d_model": gbs_ref / hbm,
                        "tflops_per_gpu_executed": fl_exec * n / sec / 1e12,
                        "frac_cuda_core_peak_executed": fl_exec * n / sec / 1e12 / peak_tf,
                        "executed_flops_per_dof": fl_exec,
                        "tflops_per_gpu_reference_flop_model": fl_ref * n / sec / 1e12,
                        "reference_flop_model_per_dof": fl_ref,
                        "flop_note": "executed = the Kronecker-form tile kernel's own count; reference model = "
                                     "161.7 flop/DoF of the reference's sum-factorisation cell kernel (SURVEY 8(d)), "
                                     "quoted for comparability only"}
    matvec["fp64"]["vs_readme_12core_broadwell"] = matvec["fp64"]["dofs_per_s"] / BASELINE_MATVEC_DOFS_PER_S
    # parity of this very run against the reference's README row for the mesh (README.md:135-159), asserted
    parity = None
    golden = readme_row(total_dofs)
    if golden is not None:
        rel = abs(r["l2_error"] - golden["l2_cg"]) / golden["l2_cg"]
        parity = {"cg_iterations": r["its"], "cg_iterations_README": golden["cg_its"],
                  "l2_error": r["l2_error"], "l2_error_README": golden["l2_cg"], "l2_rel_diff_vs_README": rel,
                  "l2_error_oracle_same_mesh": ORACLE_L2.get(total_dofs),
                  "cg_reduction": r["rate"], "cg_reduction_README": golden["cg_reduction"],
                  "asserted": "iterations equal; L2 error within 5e-3 relative of the README row (the CPU oracle on "
                              "the same 135 M DoF mesh gives 4.196e-10, as this run; README prints 4.207e-10)"}
        assert r["its"] == golden["cg_its"], "CG iterations %d != README %d" % (r["its"], golden["cg_its"])
        assert rel < 5e-3, "L2 error %g differs from README %g" % (r["l2_error"], golden["l2_cg"])
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
        "parity": parity, "matvec": matvec, "roofline": roofline, "kernels_finest_level": kernels,
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
    # the same operator and the solves on the curved mesh: one coarse cell of hyper_shell(0, 0.5, 1, 6) refined 6 times
    # (262144 cells, 17 M DoF), six-entry coefficient from the Jacobian of the MappingQ(4) cell geometry
    import time_shell
    sh = time_shell.run(6, DEGREE, reps=10)
    for k in ("fp64", "fp32"):
        sh[k]["frac_hbm"] = sh[k]["gbs"] / r["hbm_gbs"]
    sh["workload"] = ("poisson_shell functions on one hyper_shell(0, 0.5, 1, 6) coarse cell, 6 refinements: curved cells, "
                      "coefficient contrast 1e6; operator, solve_cg to 1e-9, FMG solve")
    out["shell_sector"] = sh
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
            c = run_cpu_solver(8, 1, cpu_sample_refinements(False), budget_s=25)
            nc = round(c["cells"] ** (1 / 3))
            line["cpu_baseline"] = {
                "value": c["n_dofs"] / c["seconds_per_solve"], "unit": "DoF/s", "cores": c["cores"], "kind": "port",
                "sample": "solve_cg on %d^3 cells = %d DoF, %d timed solves (about 25 s of CPU work), oracle/mg_oracle.cpp, "
                          "%d OpenMP threads (omp_get_max_threads)" % (nc, c["n_dofs"], c["steps_timed"], c["cores"]),
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
