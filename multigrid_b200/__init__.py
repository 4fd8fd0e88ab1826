"""multigrid_b200 -- host-side mirror of the reference's solver/operator interface over the C ABI.

The numerical path lives entirely in ``libmgb200.so`` (hand-written sm_100a CUDA behind
``include/mgb200.h``).  This module is plumbing: it binds the C ABI with ctypes and exposes
classes that carry the reference's names and call signatures

    MultigridSolver   <-> multigrid::MultigridSolver<3,p,float,double>   common/multigrid_solver.h:96-782
    LaplaceOperator   <-> multigrid::LaplaceOperator<3,p,number>         common/laplace_operator.h:56-164
    Smoother          <-> dealii::PreconditionChebyshev                  multigrid_solver.h:271-288
    Transfer          <-> dealii::MGTransferMatrixFree                   multigrid_solver.h:213-221

PyTorch is used only for device memory (vectors are ``torch.Tensor`` on ``cuda``).  There is no
CPU fallback: loading fails loudly when the extension is missing, and creating a solver fails
when no GPU is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MGB200_LIB", os.path.join(_HERE, "libmgb200.so"))  # override: kernel experiments

F32, F64 = 0, 1
PROBLEM_CUBE_SINE, PROBLEM_SHELL_COEF = 0, 1
KERNEL_AUTO, KERNEL_GENERAL = 0, 1
GEOMETRY_BOX, GEOMETRY_SHELL_SECTOR, GEOMETRY_AFFINE = 0, 1, 2


class MgbError(RuntimeError):
    """Raised where the reference would AssertThrow (poisson_cube/program.cc:717-737)."""


class Options(C.Structure):
    _fields_ = [
        ("degree", C.c_int),
        ("n_subdivisions", C.c_int * 3),
        ("n_refinements", C.c_int),
        ("lower", C.c_double * 3),
        ("upper", C.c_double * 3),
        ("degree_pre", C.c_int),
        ("degree_post", C.c_int),
        ("n_cycles", C.c_int),
        ("problem", C.c_int),
        ("kernel", C.c_int),
        ("device", C.c_int),
        ("tile_cells", C.c_int),
        ("comm_rank", C.c_int),
        ("comm_size", C.c_int),
        ("comm_id", C.c_ubyte * 128),
        ("geometry", C.c_int),
        ("geometry_params", C.c_double * 12),
    ]


class DGOptions(C.Structure):
    _fields_ = [
        ("degree", C.c_int),
        ("n_subdivisions", C.c_int * 3),
        ("n_refinements", C.c_int),
        ("lower", C.c_double * 3),
        ("upper", C.c_double * 3),
        ("trafo", C.c_double * 9),
        ("device", C.c_int),
        ("comm_rank", C.c_int),
        ("comm_size", C.c_int),
        ("comm_id", C.c_ubyte * 128),
        ("basis", C.c_int),
    ]


class DGSolverOptions(C.Structure):
    _fields_ = [
        ("degree", C.c_int),
        ("n_subdivisions", C.c_int * 3),
        ("n_refinements", C.c_int),
        ("lower", C.c_double * 3),
        ("upper", C.c_double * 3),
        ("degree_pre", C.c_int),
        ("degree_post", C.c_int),
        ("preconditioner", C.c_int),
        ("device", C.c_int),
        ("comm_rank", C.c_int),
        ("comm_size", C.c_int),
        ("comm_id", C.c_ubyte * 128),
        ("basis", C.c_int),
    ]


_lib = None

# name -> (restype, argtypes); the full export list of include/mgb200.h
_vp, _i, _d, _u = C.c_void_p, C.c_int, C.c_double, C.c_uint
_dp, _ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
SYMBOLS = {
    "mgb_last_error": (C.c_char_p, []),
    "mgb_version": (_i, []),
    "mgb_default_options": (None, [C.POINTER(Options)]),
    "mgb_solver_create": (_i, [C.POINTER(Options), C.POINTER(_vp)]),
    "mgb_solver_destroy": (_i, [_vp]),
    "mgb_solver_sync": (_i, [_vp]),
    "mgb_solver_n_levels": (_i, [_vp]),
    "mgb_solver_n_dofs": (C.c_int64, [_vp, _i]),
    "mgb_solver_partition": (_i, [_vp, _i, C.POINTER(C.c_int64)]),
    "mgb_solver_exchange_count": (C.c_int64, [_vp]),
    "mgb_comm_unique_id": (_i, [C.POINTER(C.c_ubyte)]),
    "mgb_solver_dof_grid": (_i, [_vp, _i, _ip]),
    "mgb_solver_tile_cells": (_i, [_vp, _i]),
    "mgb_solver_lex_map": (_i, [_vp, _i, C.POINTER(C.c_uint32)]),
    "mgb_solver_solve": (_i, [_vp, _i, _dp, _dp]),
    "mgb_solver_solve_cg": (_i, [_vp, C.POINTER(_u), _dp]),
    "mgb_solver_solve_cg_host": (_i, [_vp, _vp, _vp, C.POINTER(_u), _dp]),
    "mgb_solver_cg_history": (_i, [_vp, _dp, _ip]),
    "mgb_solver_l2_error": (_i, [_vp, _i, _dp]),
    "mgb_solver_vcycle": (_i, [_vp, _vp, _vp]),
    "mgb_solver_vmult_with_residual_update": (_i, [_vp, _vp, _vp, _d, C.POINTER(C.c_double)]),
    "mgb_solver_n_quadrature_points": (C.c_int64, [_vp, _i]),
    "mgb_solver_quadrature_points": (_i, [_vp, _i, _dp]),
    "mgb_solver_support_points": (_i, [_vp, _i, _dp]),
    "mgb_solver_set_coefficient_values": (_i, [_vp, _i, _dp]),
    "mgb_solver_reinit_smoothers": (_i, [_vp]),
    "mgb_solver_set_rhs_values": (_i, [_vp, _i, _dp, _dp]),
    "mgb_solver_set_exact_solution_values": (_i, [_vp, _i, _dp]),
    "mgb_solver_minimal_surface_set_boundary_values": (_i, [_vp, _dp]),
    "mgb_solver_minimal_surface_step": (_i, [_vp, _i, _dp]),
    "mgb_solver_minimal_surface_residual": (_i, [_vp, _vp, _vp, _i, _dp]),
    "mgb_solver_minimal_surface_level_solution": (_i, [_vp, _i, C.POINTER(_vp)]),
    "mgb_solver_do_matvec": (_i, [_vp]),
    "mgb_solver_do_matvec_smoother": (_i, [_vp]),
    "mgb_solver_timings": (_i, [_vp, _dp, _i]),
    "mgb_solver_profile": (_i, [_vp, _i]),
    "mgb_solver_class_times": (_i, [_vp, _i, _dp, C.POINTER(C.c_int64), _i]),
    "mgb_timer_start": (_i, [_vp]),
    "mgb_timer_stop": (_i, [_vp, _dp]),
    "mgb_op_vmult_host": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_solver_vector": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "mgb_solver_get_solution": (_i, [_vp, _vp]),
    "mgb_op_vmult": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_op_vmult_add": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_op_Tvmult": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_op_vmult_residual": (_i, [_vp, _i, _i, _vp, _vp, _vp]),
    "mgb_op_vmult_with_cg_update": (_i, [_vp, _i, _i, _d, _d, _vp, _vp, _vp, _vp, _dp]),
    "mgb_op_inverse_diagonal": (_i, [_vp, _i, _i, _vp]),
    "mgb_op_set_kernel": (_i, [_vp, _i]),
    "mgb_smoother_reinit": (_i, [_vp, _i, _d, _i, _i]),
    "mgb_solver_vcycle_inplace": (_i, [_vp]),
    "mgb_smoother_info": (_i, [_vp, _i, _dp]),
    "mgb_smoother_vmult": (_i, [_vp, _i, _vp, _vp]),
    "mgb_smoother_step": (_i, [_vp, _i, _vp, _vp]),
    "mgb_transfer_prolongate": (_i, [_vp, _i, _vp, _vp]),
    "mgb_transfer_prolongate_and_add": (_i, [_vp, _i, _vp, _vp]),
    "mgb_transfer_restrict_and_add": (_i, [_vp, _i, _vp, _vp]),
    "mgb_transfer_prolongate_no_boundary": (_i, [_vp, _i, _vp, _vp]),
    "mgb_vec_dot": (_i, [_vp, _i, _i, _vp, _vp, _dp]),
    "mgb_vec_convert": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_op_compute_residual": (_i, [_vp, _i, _vp, _vp, _dp]),
    "mgb_op_el": (_i, [_vp, _i, _i, C.c_int64, _dp]),
    "mgb_adaptive_mg_level_vmult": (_i, [_vp, _i, _i, _vp, _vp]),
    "mgb_noalloc_dot": (_i, [_i, C.c_int64, _vp, _vp, _vp, _vp]),
    "mgb_vec_sadd": (_i, [_vp, _i, _i, _vp, _d, _d, _vp]),
    "mgb_vec_set": (_i, [_vp, _i, _i, _vp, _d]),
    "mgb_vec_update_ghost_values": (_i, [_vp, _i, _i, _vp]),
    "mgb_adaptive_mg_setup": (_i, [_vp, _i]),
    "mgb_adaptive_mg_n_levels": (_i, [_vp]),
    "mgb_adaptive_mg_level_info": (_i, [_vp, _i, _dp]),
    "mgb_adaptive_mg_vmult": (_i, [_vp, _vp, _vp]),
    "mgb_adaptive_solve_cg": (_i, [_vp, _vp, _vp, _d, C.POINTER(C.c_uint), _dp, _dp, _ip]),
    "mgb_vec_compress_add": (_i, [_vp, _i, _i, _vp]),
    "mgb_vec_zero_out_ghost_values": (_i, [_vp, _i, _i, _vp]),
    "mgb_device_alloc": (_i, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "mgb_device_free": (_i, [_vp, _vp]),
    "mgb_copy_to_device": (_i, [_vp, _vp, _vp, C.c_size_t]),
    "mgb_copy_to_host": (_i, [_vp, _vp, _vp, C.c_size_t]),
    "mgb_solver_launch_count": (C.c_int64, [_vp]),
    # DG operator
    "mgb_dg_default_options": (None, [C.POINTER(DGOptions)]),
    "mgb_dg_create": (_i, [C.POINTER(DGOptions), C.POINTER(_vp)]),
    "mgb_dg_destroy": (_i, [_vp]),
    "mgb_dg_sync": (_i, [_vp]),
    "mgb_dg_n_dofs": (C.c_int64, [_vp]),
    "mgb_dg_n_cells": (_i, [_vp, _ip]),
    "mgb_dg_penalty": (_i, [_vp, _dp]),
    "mgb_dg_partition": (_i, [_vp, C.POINTER(C.c_int64)]),
    "mgb_dg_exchange_count": (C.c_int64, [_vp]),
    "mgb_dg_vmult": (_i, [_vp, _i, _vp, _vp]),
    "mgb_dg_vmult_residual": (_i, [_vp, _i, _vp, _vp, _vp]),
    "mgb_dg_vmult_with_cg_update": (_i, [_vp, _i, _d, _d, _vp, _vp, _vp, _vp, _dp]),
    "mgb_dg_inverse_diagonal": (_i, [_vp, _i, _vp]),
    "mgb_dg_vmult_with_chebyshev_update": (_i, [_vp, _i, _vp, _i, _d, _d, _vp, _vp]),
    "mgb_dg_prolongate_add_cg_to_dg": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "mgb_dg_vmult_residual_and_restrict_to_cg": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "mgb_dg_set_preconditioner": (_i, [_vp, _i]),
    "mgb_dg_preconditioner_vmult": (_i, [_vp, _i, _vp, _vp]),
    "mgb_dg_vmult_host": (_i, [_vp, _i, _vp, _vp]),
    "mgb_dg_lumped_mass_inverse": (_i, [_vp, _i, _vp]),
    "mgb_dg_solve_cg": (_i, [_vp, _i, _u, _i, _vp, _vp, _vp, _dp]),
    "mgb_dg_timer_start": (_i, [_vp]),
    "mgb_dg_timer_stop": (_i, [_vp, _dp]),
    "mgb_dg_launch_count": (C.c_int64, [_vp]),
    # DG multigrid solver
    "mgb_dg_solver_default_options": (None, [C.POINTER(DGSolverOptions)]),
    "mgb_dg_solver_create": (_i, [C.POINTER(DGSolverOptions), C.POINTER(_vp)]),
    "mgb_dg_solver_destroy": (_i, [_vp]),
    "mgb_dg_solver_n_dofs": (C.c_int64, [_vp]),
    "mgb_dg_solver_parts": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "mgb_dg_solver_solve_cg": (_i, [_vp, _d, _dp, _dp, C.POINTER(_u)]),
    "mgb_dg_solver_residual_history": (_i, [_vp, _dp, _i, _ip]),
    "mgb_dg_solver_vmult": (_i, [_vp, _vp, _vp]),
    "mgb_dg_solver_vmult_with_residual_update": (_i, [_vp, _vp, _vp, _d, _dp]),
    "mgb_dg_solver_l2_error": (_i, [_vp, _dp]),
    "mgb_dg_solver_vector": (_i, [_vp, _i, C.POINTER(_vp)]),
    "mgb_dg_solver_get_solution": (_i, [_vp, _vp]),
    "mgb_dg_solver_smoother_info": (_i, [_vp, _dp]),
    "mgb_dg_solver_do_matvec": (_i, [_vp]),
    "mgb_dg_solver_do_matvec_smoother": (_i, [_vp]),
    "mgb_dg_solver_sync": (_i, [_vp]),
    "mgb_dg_solver_solve_cg_host": (_i, [_vp, _d, _vp, _dp, _dp, C.POINTER(_u)]),
    "mgb_dg_solver_launch_count": (C.c_int64, [_vp]),
    "mgb_dg_plain_solver_create": (_i, [C.POINTER(DGSolverOptions), C.POINTER(_vp)]),
    "mgb_dg_plain_solver_destroy": (_i, [_vp]),
    "mgb_dg_plain_solver_n_dofs": (C.c_int64, [_vp]),
    "mgb_dg_plain_solver_n_levels": (_i, [_vp]),
    "mgb_dg_plain_solver_level_operator": (_i, [_vp, _i, C.POINTER(_vp)]),
    "mgb_dg_plain_solver_solve_cg": (_i, [_vp, _d, _dp, _dp, C.POINTER(_u)]),
    "mgb_dg_plain_solver_residual_history": (_i, [_vp, _dp, _i, _ip]),
    "mgb_dg_plain_solver_vmult": (_i, [_vp, _vp, _vp]),
    "mgb_dg_plain_solver_vmult_with_residual_update": (_i, [_vp, _vp, _vp, _d, _dp]),
    "mgb_dg_plain_solver_l2_error": (_i, [_vp, _dp]),
    "mgb_dg_plain_solver_vector": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "mgb_dg_plain_solver_get_vector": (_i, [_vp, _i, _i, _vp]),
    "mgb_dg_plain_solver_get_solution": (_i, [_vp, _vp]),
    "mgb_dg_plain_solver_smoother_info": (_i, [_vp, _i, _dp]),
    "mgb_dg_plain_solver_prolongate": (_i, [_vp, _i, _vp, _vp, _i]),
    "mgb_dg_plain_solver_restrict": (_i, [_vp, _i, _vp, _vp, _i]),
    "mgb_dg_plain_solver_do_matvec": (_i, [_vp]),
    "mgb_dg_plain_solver_do_matvec_smoother": (_i, [_vp]),
    "mgb_dg_plain_solver_sync": (_i, [_vp]),
    "mgb_dg_plain_solver_solve_cg_host": (_i, [_vp, _d, _vp, _dp, _dp, C.POINTER(_u)]),
    "mgb_dg_plain_solver_launch_count": (C.c_int64, [_vp]),
    # adaptive meshes with hanging nodes
    "mgb_adaptive_create": (_i, [_i, _ip, C.POINTER(C.c_ubyte), _dp, _dp, _i, _i, C.POINTER(_vp)]),
    "mgb_adaptive_destroy": (_i, [_vp]),
    "mgb_adaptive_n_cells": (C.c_int64, [_vp]),
    "mgb_adaptive_cells": (_i, [_vp, _ip, _ip]),
    "mgb_adaptive_refine": (_i, [_vp, C.POINTER(C.c_ubyte)]),
    "mgb_adaptive_n_dofs": (C.c_int64, [_vp]),
    "mgb_adaptive_info": (_i, [_vp, C.POINTER(C.c_int64)]),
    "mgb_adaptive_dof_points": (_i, [_vp, _dp]),
    "mgb_adaptive_dirichlet": (_i, [_vp, C.POINTER(C.c_ubyte)]),
    "mgb_adaptive_tables": (_i, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), _dp]),
    "mgb_adaptive_vmult": (_i, [_vp, _i, _vp, _vp]),
    "mgb_adaptive_inverse_diagonal": (_i, [_vp, _i, _vp]),
    "mgb_adaptive_vmult_host": (_i, [_vp, _i, _vp, _vp]),
    "mgb_adaptive_sync": (_i, [_vp]),
    "mgb_adaptive_launch_count": (C.c_int64, [_vp]),
    "mgb_adaptive_refine_and_coarsen": (_i, [_vp, C.POINTER(C.c_ubyte), C.POINTER(C.c_ubyte)]),
    "mgb_adaptive_refine_and_coarsen_fixed_number": (_i, [_vp, C.POINTER(C.c_float), _d, _d]),
    "mgb_adaptive_face_neighbours": (_i, [_vp, C.POINTER(C.c_byte), C.POINTER(C.c_int32)]),
    "mgb_adaptive_plain_table": (_i, [_vp, C.POINTER(C.c_int32)]),
    "mgb_adaptive_interpolate_boundary_values": (_i, [_vp, _i, _vp]),
    "mgb_adaptive_compute_residual": (_i, [_vp, _vp, _vp, _d]),
    "mgb_adaptive_estimate_kelly": (_i, [_vp, _vp, C.POINTER(C.c_float)]),
    "mgb_adaptive_integrate_difference": (_i, [_vp, _vp, _i, _i, _i, _dp, _dp]),
    "mgb_adaptive_solve_cycle": (_i, [_vp, _i, _d, _i, _dp, C.POINTER(_vp)]),
    # host-only helpers
    "mgb_host_basis": (_i, [_i] + [_dp] * 9),
    "mgb_host_numbering": (_i, [_i, _ip, _i, C.POINTER(C.c_int64), _ip, C.POINTER(C.c_int64),
                                C.POINTER(C.c_int64), C.POINTER(C.c_uint32)]),
    "mgb_host_partition": (_i, [_i, _ip, _i, _i, _i, _i, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                C.POINTER(C.c_uint32)]),
}


def lib():
    """Load libmgb200.so; there is deliberately no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MgbError(
                "libmgb200.so is missing (%s): build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` or `make -C multigrid_b200/csrc -j8`; there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(code):
    if code != 0:
        raise MgbError(lib().mgb_last_error().decode())


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def host_basis(p):
    n = p + 1
    out = {k: np.zeros((n, n)) for k in ("S", "G", "D", "M1", "K1")}
    out["P1"] = np.zeros((2 * p + 1, n))
    for k in ("nodes", "xq", "wq"):
        out[k] = np.zeros(n)
    f = lambda a: a.ctypes.data_as(_dp)
    _check(lib().mgb_host_basis(p, f(out["S"]), f(out["G"]), f(out["D"]), f(out["M1"]), f(out["K1"]),
                                f(out["P1"]), f(out["nodes"]), f(out["xq"]), f(out["wq"])))
    return out


def host_numbering(p, ncell, tile_cells=0, want_map=True):
    nc = (C.c_int * 3)(*ncell)
    nd, nt, ncon, cw = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int()
    lex = None
    _check(lib().mgb_host_numbering(p, nc, tile_cells, C.byref(nd), C.byref(cw), C.byref(nt), C.byref(ncon), None))
    if want_map:
        lex = np.zeros(nd.value, dtype=np.uint32)
        _check(lib().mgb_host_numbering(p, nc, tile_cells, C.byref(nd), C.byref(cw), C.byref(nt), C.byref(ncon),
                                        lex.ctypes.data_as(C.POINTER(C.c_uint32))))
    return dict(n_dofs=nd.value, tile_cells=cw.value, n_tiles=nt.value, n_constrained=ncon.value, lex_to_dof=lex)


PARTITION_KEYS = ("distributed", "n_global", "n_local", "owned_begin", "n_owned", "plane0", "n_planes",
                  "n_planes_global", "ghost_layers_below", "ghost_plane_top", "layer_begin", "layer_end")


def host_partition(p, n_subdivisions, level, n_ranks, rank, tile_cells=0, want_map=False, min_dofs=0):
    """z-slab of `rank` on one multigrid level and its ghost-exchange plan (host logic of csrc/numbering.h)."""
    nsub = (C.c_int * 3)(*n_subdivisions)
    info = (C.c_int64 * 12)()
    plan = (C.c_int64 * 8)()
    _check(lib().mgb_host_partition(p, nsub, level, tile_cells, n_ranks, rank, min_dofs, info, plan, None))
    out = dict(zip(PARTITION_KEYS, info))
    for i, k in enumerate(("send_up", "recv_dn", "send_dn", "recv_up")):
        out[k] = (plan[2 * i], plan[2 * i + 1])
    if want_map:
        lex = np.zeros(out["n_local"], dtype=np.uint32)
        _check(lib().mgb_host_partition(p, nsub, level, tile_cells, n_ranks, rank, min_dofs, info, plan,
                                        lex.ctypes.data_as(C.POINTER(C.c_uint32))))
        out["lex_to_dof"] = lex.astype(np.int64)
    return out


def comm_unique_id():
    """128-byte NCCL id made by rank 0 (mgb_comm_unique_id); the caller distributes it to all ranks."""
    buf = (C.c_ubyte * 128)()
    _check(lib().mgb_comm_unique_id(buf))
    return bytes(buf)


NOALLOC_DOT_MAX_N = 268435456


def noalloc_dot(x, y, out, stream=None):
    """mgb_noalloc_dot: out[0] = x . y (accumulated in double) for two CUDA tensors of dtype float32 or float64; `out` is
    a float64 CUDA tensor with one entry.  Allocates nothing (the scratch is a static device array in the library), needs
    no solver, is queued on `stream` (a torch.cuda.Stream; default: torch's current stream) and does not synchronise.
    At most NOALLOC_DOT_MAX_N entries; calls that overlap on one GPU race on the shared scratch array."""
    import torch
    if x.dtype != y.dtype or x.dtype not in (torch.float32, torch.float64):
        raise MgbError("noalloc_dot: float32 or float64 tensors of the same dtype")
    if x.numel() != y.numel() or not (x.is_cuda and y.is_cuda and out.is_cuda) or not (x.is_contiguous() and y.is_contiguous()):
        raise MgbError("noalloc_dot: contiguous CUDA tensors of the same length")
    if out.dtype != torch.float64 or out.numel() < 1:
        raise MgbError("noalloc_dot: out must be a float64 CUDA tensor")
    if stream is None:
        stream = torch.cuda.current_stream(x.device)
    with torch.cuda.device(x.device):
        _check(lib().mgb_noalloc_dot(F64 if x.dtype == torch.float64 else F32, x.numel(), C.c_void_p(x.data_ptr()),
                                     C.c_void_p(y.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(stream.cuda_stream)))
    return out


def comm_from_torch_distributed():
    """(rank, size, id) for MultigridSolver(comm=...) from an initialised torch.distributed process group:
    rank 0 creates the id and broadcasts it (plumbing only; the data path talks NCCL itself)."""
    import torch
    import torch.distributed as dist
    rank, size = dist.get_rank(), dist.get_world_size()
    if size == 1:
        return None
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(t, 0)
    return rank, size, bytes(t.cpu().numpy().tobytes())


class LaplaceOperator:
    """View of ``matrix[level]`` (float) or ``matrix_dp[level]`` (double) of a solver."""

    def __init__(self, solver, level, precision):
        self.solver, self.level, self.precision = solver, level, precision
        import torch
        self.dtype = torch.float64 if precision == F64 else torch.float32

    def m(self):
        """global number of DoFs (MatrixFreeOperators::Base::m)"""
        return self.solver.partition(self.level)["n_global"]

    n = m

    def local_size(self):
        """length of this rank's vectors: owned planes plus ghost planes (all DoFs on one GPU)"""
        return self.solver.n_dofs(self.level)

    def initialize_dof_vector(self):
        import torch
        return torch.zeros(self.local_size(), dtype=self.dtype, device=self.solver.device)

    def _chk(self, *ts):
        for t in ts:
            if t.dtype != self.dtype or t.numel() != self.local_size() or not t.is_cuda or not t.is_contiguous():
                raise MgbError("vector does not match the operator (dtype/size/device)")

    def vmult(self, dst, src):
        self._chk(dst, src)
        self.solver._torch_sync()
        _check(lib().mgb_op_vmult(self.solver.h, self.level, self.precision, _ptr(dst), _ptr(src)))
        self.solver.sync()

    def Tvmult(self, dst, src):
        self._chk(dst, src)
        self.solver._torch_sync()
        _check(lib().mgb_op_Tvmult(self.solver.h, self.level, self.precision, _ptr(dst), _ptr(src)))
        self.solver.sync()

    def vmult_add(self, dst, src):
        self._chk(dst, src)
        self.solver._torch_sync()
        _check(lib().mgb_op_vmult_add(self.solver.h, self.level, self.precision, _ptr(dst), _ptr(src)))
        self.solver.sync()

    Tvmult_add = vmult_add

    def vmult_residual(self, rhs, lhs, residual):
        self._chk(rhs, lhs, residual)
        self.solver._torch_sync()
        _check(lib().mgb_op_vmult_residual(self.solver.h, self.level, self.precision, _ptr(rhs), _ptr(lhs),
                                           _ptr(residual)))
        self.solver.sync()

    def compute_residual(self, dst, src, rhs_function=None):
        """laplace_operator.h:103-106: dst = (f, phi) - K src on the rows without Dirichlet condition (double operator);
        rhs_function(x, y, z) on arrays, None = the right-hand side of the solver's problem"""
        self._chk(dst, src)
        f = None
        if rhs_function is not None:
            xq = self.solver.quadrature_points(self.level)
            f = np.ascontiguousarray(rhs_function(xq[:, 0], xq[:, 1], xq[:, 2]), dtype=np.float64)
        self.solver._torch_sync()
        _check(lib().mgb_op_compute_residual(self.solver.h, self.level, _ptr(dst), _ptr(src),
                                             f.ctypes.data_as(_dp) if f is not None else None))

    def el(self, row, col=None):
        """MatrixFreeOperators::Base::el(row, row)"""
        if col is not None and col != row:
            raise MgbError("matrix-free operator: only diagonal entries are available")
        v = C.c_double()
        _check(lib().mgb_op_el(self.solver.h, self.level, self.precision, int(row), C.byref(v)))
        return v.value

    def vmult_with_cg_update(self, alpha, beta, r, q, p, x):
        self._chk(r, q, p, x)
        sums = (C.c_double * 4)()
        self.solver._torch_sync()
        _check(lib().mgb_op_vmult_with_cg_update(self.solver.h, self.level, self.precision, alpha, beta,
                                                 _ptr(r), _ptr(q), _ptr(p), _ptr(x), sums))
        return list(sums)

    def get_matrix_diagonal_inverse(self):
        d = self.initialize_dof_vector()
        self.solver._torch_sync()
        _check(lib().mgb_op_inverse_diagonal(self.solver.h, self.level, self.precision, _ptr(d)))
        self.solver.sync()
        return d

    compute_diagonal = get_matrix_diagonal_inverse


class MultigridSolver:
    """Mixed-precision FMG / CG+V-cycle solver on a refined box mesh.  One process per GPU: with
    ``comm=(rank, size, id)`` (see comm_from_torch_distributed) the mesh is cut into z-slabs and the
    constructor is collective over the ranks."""

    def __init__(self, degree, n_subdivisions, n_refinements, lower=(-0.9,) * 3, upper=(1.0,) * 3,
                 degree_pre=3, degree_post=3, n_cycles=2, problem=PROBLEM_CUBE_SINE,
                 kernel=KERNEL_AUTO, device=0, tile_cells=0, comm=None, geometry=GEOMETRY_BOX, geometry_params=None):
        """geometry: GEOMETRY_BOX (axis-aligned cells on [lower, upper]), GEOMETRY_SHELL_SECTOR (one coarse cell of
        hyper_shell(0, r0, r1, 6) and its refinements, geometry_params = (r0, r1, opening half-angle), default
        (0.5, 1.0, pi/4)), GEOMETRY_AFFINE (x = lower + A ((upper - lower) o xi), geometry_params = A row-major)."""
        import torch
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        o = Options()
        lib().mgb_default_options(C.byref(o))
        o.degree = degree
        o.n_subdivisions[:] = list(n_subdivisions)
        o.n_refinements = n_refinements
        o.lower[:] = list(lower)
        o.upper[:] = list(upper)
        o.degree_pre, o.degree_post, o.n_cycles = degree_pre, degree_post, n_cycles
        o.problem, o.kernel, o.device, o.tile_cells = problem, kernel, device, tile_cells
        o.geometry = geometry
        if geometry_params is not None:
            gp = list(geometry_params)
            o.geometry_params[:len(gp)] = gp
        self.rank, self.n_ranks = 0, 1
        if comm is not None:
            self.rank, self.n_ranks, cid = comm
            o.comm_rank, o.comm_size = self.rank, self.n_ranks
            o.comm_id[:] = list(cid)
        self.h = C.c_void_p()
        self.device = torch.device("cuda", device)
        self.degree = degree
        _check(lib().mgb_solver_create(C.byref(o), C.byref(self.h)))
        self.n_levels = lib().mgb_solver_n_levels(self.h)

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mgb_solver_destroy(self.h)
            self.h = None

    __del__ = close

    def _torch_sync(self):
        import torch
        torch.cuda.synchronize(self.device)

    def sync(self):
        _check(lib().mgb_solver_sync(self.h))

    def _l(self, level):
        return level if level >= 0 else self.n_levels + level

    def n_dofs(self, level=-1):
        return lib().mgb_solver_n_dofs(self.h, self._l(level))

    def partition(self, level=-1):
        """z-slab of this rank on one level (see mgb_solver_partition)"""
        info = (C.c_int64 * 12)()
        _check(lib().mgb_solver_partition(self.h, self._l(level), info))
        return dict(zip(PARTITION_KEYS, info))

    def exchange_count(self):
        return lib().mgb_solver_exchange_count(self.h)

    def dof_grid(self, level=-1):
        nd = (C.c_int * 3)()
        _check(lib().mgb_solver_dof_grid(self.h, self._l(level), nd))
        return tuple(nd)

    def tile_cells(self, level=-1):
        return lib().mgb_solver_tile_cells(self.h, self._l(level))

    def lex_map(self, level=-1):
        m = np.zeros(self.n_dofs(level), dtype=np.uint32)
        _check(lib().mgb_solver_lex_map(self.h, self._l(level), m.ctypes.data_as(C.POINTER(C.c_uint32))))
        return m.astype(np.int64)

    def matrix(self, level=-1):
        return LaplaceOperator(self, self._l(level), F32)

    def matrix_dp(self, level=-1):
        return LaplaceOperator(self, self._l(level), F64)

    # --- reference driver interface (poisson_cube/program.cc:262-380) ---
    def solve(self, do_analyze=False):
        r = C.c_double()
        trace = np.zeros((self.n_levels, 4))
        _check(lib().mgb_solver_solve(self.h, int(do_analyze), C.byref(r), trace.ctypes.data_as(_dp)))
        return (r.value, trace) if do_analyze else r.value

    def solve_cg(self):
        its, rate = C.c_uint(), C.c_double()
        _check(lib().mgb_solver_solve_cg(self.h, C.byref(its), C.byref(rate)))
        return its.value, rate.value

    def solve_cg_host(self, rhs_host=None, solution_host=None):
        """rhs_host / solution_host: pinned torch CPU tensors or numpy arrays (float64)."""
        its, rate = C.c_uint(), C.c_double()
        pr = C.c_void_p(rhs_host.data_ptr() if hasattr(rhs_host, "data_ptr") else rhs_host.ctypes.data) \
            if rhs_host is not None else None
        ps = C.c_void_p(solution_host.data_ptr() if hasattr(solution_host, "data_ptr")
                        else solution_host.ctypes.data) if solution_host is not None else None
        _check(lib().mgb_solver_solve_cg_host(self.h, pr, ps, C.byref(its), C.byref(rate)))
        return its.value, rate.value

    def cg_history(self):
        n = C.c_int(1100)
        hist = np.zeros(1100)
        _check(lib().mgb_solver_cg_history(self.h, hist.ctypes.data_as(_dp), C.byref(n)))
        return hist[:n.value]

    def compute_l2_error(self, level=-1):
        e = C.c_double()
        _check(lib().mgb_solver_l2_error(self.h, self._l(level), C.byref(e)))
        return e.value

    def vmult(self, dst, src):
        """Preconditioner: one float V-cycle (multigrid_solver.h:498-510)."""
        self._torch_sync()
        _check(lib().mgb_solver_vcycle(self.h, _ptr(dst), _ptr(src)))
        self.sync()

    # ---- problem data supplied by the caller (the reference's Function<dim> arguments, multigrid_solver.h:100-106)
    def quadrature_points(self, level=-1):
        """(n_cells (p+1)^3, 3): cells lexicographic (x fastest) on this rank's box, points x fastest inside a cell"""
        l = self._l(level)
        out = np.zeros((lib().mgb_solver_n_quadrature_points(self.h, l), 3))
        _check(lib().mgb_solver_quadrature_points(self.h, l, out.ctypes.data_as(_dp)))
        return out

    def support_points(self, level=-1):
        """(n_dofs, 3) in the library's numbering"""
        out = np.zeros((self.n_dofs(level), 3))
        _check(lib().mgb_solver_support_points(self.h, self._l(level), out.ctypes.data_as(_dp)))
        return out

    def set_problem(self, coefficient=None, right_hand_side=None, boundary_values=None, exact_solution=None):
        """Functions f(x, y, z) -> values on arrays, evaluated here at the library's quadrature / support points of every
        level and handed over as samples (LaplaceOperator::evaluate_coefficient(const Function &), compute_residual(...,
        rhs_function), interpolate_boundary_values)."""
        for l in range(self.n_levels):
            xq = None
            if coefficient is not None or right_hand_side is not None or exact_solution is not None:
                xq = self.quadrature_points(l)
            if coefficient is not None:
                c = np.ascontiguousarray(coefficient(xq[:, 0], xq[:, 1], xq[:, 2]), dtype=np.float64)
                _check(lib().mgb_solver_set_coefficient_values(self.h, l, c.ctypes.data_as(_dp)))
        if coefficient is not None:
            _check(lib().mgb_solver_reinit_smoothers(self.h))
        for l in range(self.n_levels):
            if right_hand_side is not None:
                xq = self.quadrature_points(l)
                f = np.ascontiguousarray(right_hand_side(xq[:, 0], xq[:, 1], xq[:, 2]), dtype=np.float64)
                g = None
                if boundary_values is not None:
                    xs = self.support_points(l)
                    g = np.ascontiguousarray(boundary_values(xs[:, 0], xs[:, 1], xs[:, 2]), dtype=np.float64)
                _check(lib().mgb_solver_set_rhs_values(self.h, l, f.ctypes.data_as(_dp),
                                                       g.ctypes.data_as(_dp) if g is not None else None))
            if exact_solution is not None:
                xq = self.quadrature_points(l)
                u = np.ascontiguousarray(exact_solution(xq[:, 0], xq[:, 1], xq[:, 2]), dtype=np.float64)
                _check(lib().mgb_solver_set_exact_solution_values(self.h, l, u.ctypes.data_as(_dp)))

    # ---- minimal_surface/program.cc: Newton's method with line search on -div(grad u / sqrt(1 + |grad u|^2)) = 0
    def minimal_surface_set_boundary_values(self, boundary_values=None):
        """interpolate_boundary_values (:396-408): a callable g(x, y, z) evaluated at the support points of the finest
        level, or None for the analytic solution of `problem`; the iterate becomes g on the boundary, zero inside"""
        g = None
        if boundary_values is not None:
            xs = self.support_points()
            g = np.ascontiguousarray(boundary_values(xs[:, 0], xs[:, 1], xs[:, 2]), dtype=np.float64)
        _check(lib().mgb_solver_minimal_surface_set_boundary_values(self.h, g.ctypes.data_as(_dp) if g is not None else None))

    def minimal_surface_step(self, first_time):
        """LaplaceProblem::solve(first_time) (:412-572): coefficients from the iterate on every level, smoothers, CG with
        the V-cycle on the linearised operator, line search"""
        info = np.zeros(5)
        _check(lib().mgb_solver_minimal_surface_step(self.h, int(first_time), info.ctypes.data_as(_dp)))
        return dict(initial_residual=info[0], final_residual=info[1], linear_iterations=int(info[2]),
                    line_search_steps=int(info[3]), step_length=info[4])

    def minimal_surface_solve(self, max_steps=100, tolerance=1e-12):
        """the inner loop of LaplaceProblem::run (:652-659) -> the list of the steps' records"""
        steps = []
        for it in range(max_steps):
            steps.append(self.minimal_surface_step(it == 0))
            if steps[-1]["final_residual"] < tolerance:
                break
        return steps

    def minimal_surface_residual(self, dst, u, first_time=False):
        """MinimalSurfaceOperator::compute_residual (:169-198) on double device tensors; returns |dst|_2"""
        import torch
        assert dst.dtype == torch.float64 and u.dtype == torch.float64 and dst.numel() == u.numel() == self.n_dofs()
        self._torch_sync()
        norm = C.c_double()
        _check(lib().mgb_solver_minimal_surface_residual(self.h, _ptr(dst), _ptr(u), int(first_time), C.byref(norm)))
        return norm.value

    def minimal_surface_level_solution(self, level):
        """host copy of coefficient_solutions[level] (:417-456) of the last step, library numbering"""
        p = C.c_void_p()
        _check(lib().mgb_solver_minimal_surface_level_solution(self.h, self._l(level), C.byref(p)))
        out = np.zeros(self.n_dofs(level), dtype=np.float32)
        _check(lib().mgb_copy_to_host(self.h, C.c_void_p(out.ctypes.data), p, out.nbytes))
        return out

    def vmult_with_residual_update(self, residual, update, factor):
        """multigrid_solver.h:516-619: residual += factor update; update = V-cycle(residual); returns
        (<update, residual>, <update, factor update_old>)."""
        self._torch_sync()
        sums = (C.c_double * 2)()
        _check(lib().mgb_solver_vmult_with_residual_update(self.h, _ptr(residual), _ptr(update), float(factor), sums))
        return sums[0], sums[1]

    def do_matvec(self):
        _check(lib().mgb_solver_do_matvec(self.h))

    def do_matvec_smoother(self):
        _check(lib().mgb_solver_do_matvec_smoother(self.h))

    def get_solution(self):
        out = np.zeros(self.n_dofs())
        _check(lib().mgb_solver_get_solution(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def timings(self, reset=True):
        t = np.zeros((self.n_levels, 6))
        _check(lib().mgb_solver_timings(self.h, t.ctypes.data_as(_dp), int(reset)))
        return t

    def profile(self, on=True):
        _check(lib().mgb_solver_profile(self.h, int(on)))

    CLASSES = ("mg_mv", "restrict", "prolongate", "inhomBC", "mg_vec", "smoother", "apply_f64", "apply_f32",
               "apply_f32_chebyshev", "apply_epilogue", "cg_vec", "other")

    def class_times(self, level=-1, reset=True):
        """{class: (milliseconds, count)} of the CUDA-event ranges on one level."""
        ms = (C.c_double * 16)()
        cnt = (C.c_int64 * 16)()
        _check(lib().mgb_solver_class_times(self.h, self._l(level), ms, cnt, int(reset)))
        return {k: (ms[i], cnt[i]) for i, k in enumerate(self.CLASSES)}

    def timer_start(self):
        _check(lib().mgb_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_double()
        _check(lib().mgb_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def vmult_host(self, dst_host, src_host, level=-1, precision=F64):
        ptr = lambda a: C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        _check(lib().mgb_op_vmult_host(self.h, self._l(level), precision, ptr(dst_host), ptr(src_host)))

    def print_wall_times(self):
        t = self.timings()
        print("level  smoother    mg_mv     mg_vec    restrict  prolongate  inhomBC")
        for l in range(1, self.n_levels):
            print("L%-2d    %-12.4g%-10.4g%-10.4g%-10.4g%-12.4g%-10.4g"
                  % (l, t[l][5], t[l][0], t[l][4], t[l][1], t[l][2], t[l][3]))

    def launch_count(self):
        return lib().mgb_solver_launch_count(self.h)

    # --- access to solver-owned vectors as host arrays (library numbering) ---
    _VEC = dict(solution=(0, np.float64), rhs=(1, np.float64), residual=(2, np.float64),
                defect=(3, np.float32), t=(4, np.float32), solution_update=(5, np.float32),
                inverse_diagonal=(6, np.float32))

    def vector_ptr(self, name, level=-1):
        which, dt = self._VEC[name]
        p = C.c_void_p()
        _check(lib().mgb_solver_vector(self.h, which, self._l(level), C.byref(p)))
        return p, dt

    def get_vector(self, name, level=-1):
        p, dt = self.vector_ptr(name, level)
        out = np.zeros(self.n_dofs(level), dtype=dt)
        self.sync()
        _check(lib().mgb_copy_to_host(self.h, C.c_void_p(out.ctypes.data), p, out.nbytes))
        return out

    def set_vector(self, name, values, level=-1):
        p, dt = self.vector_ptr(name, level)
        v = np.ascontiguousarray(values, dtype=dt)
        _check(lib().mgb_copy_to_device(self.h, p, C.c_void_p(v.ctypes.data), v.nbytes))

    # --- smoother / transfer hooks ---
    def smoother_info(self, level):
        info = (C.c_double * 6)()
        _check(lib().mgb_smoother_info(self.h, self._l(level), info))
        return dict(min_eig=info[0], max_eig=info[1], theta=info[2], delta=info[3], degree=int(info[4]),
                    cg_iterations=int(info[5]))

    def smoother_reinit(self, level, smoothing_range, degree, eig_cg_n_iterations):
        _check(lib().mgb_smoother_reinit(self.h, self._l(level), smoothing_range, degree, eig_cg_n_iterations))

    def vcycle_inplace(self):
        _check(lib().mgb_solver_vcycle_inplace(self.h))

    def smoother_vmult(self, level, x, b):
        self._torch_sync()
        _check(lib().mgb_smoother_vmult(self.h, self._l(level), _ptr(x), _ptr(b)))
        self.sync()

    def smoother_step(self, level, x, b):
        self._torch_sync()
        _check(lib().mgb_smoother_step(self.h, self._l(level), _ptr(x), _ptr(b)))
        self.sync()

    def prolongate(self, level, fine, coarse):
        self._torch_sync()
        _check(lib().mgb_transfer_prolongate(self.h, self._l(level), _ptr(fine), _ptr(coarse)))
        self.sync()

    def prolongate_and_add(self, level, fine, coarse):
        self._torch_sync()
        _check(lib().mgb_transfer_prolongate_and_add(self.h, self._l(level), _ptr(fine), _ptr(coarse)))
        self.sync()

    def restrict_and_add(self, level, coarse, fine):
        self._torch_sync()
        _check(lib().mgb_transfer_restrict_and_add(self.h, self._l(level), _ptr(coarse), _ptr(fine)))
        self.sync()

    def prolongate_no_boundary(self, level, fine, coarse):
        self._torch_sync()
        _check(lib().mgb_transfer_prolongate_no_boundary(self.h, self._l(level), _ptr(fine), _ptr(coarse)))
        self.sync()

    def set_kernel(self, kernel):
        _check(lib().mgb_op_set_kernel(self.h, kernel))


# 1-D basis of the DG space: MGB_DG_BASIS_* of include/mgb200.h; dg_basis_of_type() maps the reference's template
# parameter `type` (laplace_operator_dg.h:346-350: 0 Hermite, 1 FE_DGQ, 2 Gauss)
DG_BASIS_GAUSS_LOBATTO, DG_BASIS_HERMITE, DG_BASIS_GAUSS = 0, 1, 2


def dg_basis_of_type(type_):
    return {0: DG_BASIS_HERMITE, 1: DG_BASIS_GAUSS_LOBATTO, 2: DG_BASIS_GAUSS}[type_]


class LaplaceOperatorDG:
    """multigrid::LaplaceOperatorCompactCombine<3,p,Number,type> (common/laplace_operator_dg.h:360-2026): SIP-DG
    Laplacian on the affine box mesh of matvec_dg/program.cc:40-80 with FE_DGQ(p) (default), FE_DGQHermite(p) or
    FE_DGQArbitraryNodes(QGauss(p+1)) (`basis`).  Vectors are torch tensors on the device, cell-wise contiguous
    coefficients of that basis; `precision` picks the float or the double kernels."""

    def __init__(self, degree=4, n_subdivisions=(1, 1, 1), n_refinements=0, lower=None, upper=None, trafo=None,
                 device=0, precision=F64, comm=None, basis=DG_BASIS_GAUSS_LOBATTO):
        """comm = (rank, size, unique_id bytes) for one process per GPU (comm_from_torch_distributed())"""
        import torch
        o = DGOptions()
        lib().mgb_dg_default_options(C.byref(o))
        if comm is not None and comm[1] > 1:
            o.comm_rank, o.comm_size = comm[0], comm[1]
            o.comm_id[:] = list(comm[2])
        o.degree = degree
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        o.n_subdivisions[:] = list(n_subdivisions)
        o.n_refinements = n_refinements
        if lower is not None:
            o.lower[:] = list(lower)
        if upper is not None:
            o.upper[:] = list(upper)
        if trafo is not None:
            o.trafo[:] = [float(v) for v in np.asarray(trafo, dtype=np.float64).ravel()]
        o.device = device
        o.basis = basis
        self.h = C.c_void_p()
        self.device = torch.device("cuda", device)
        self.precision = precision
        self.dtype = torch.float64 if precision == F64 else torch.float32
        self.degree = degree
        _check(lib().mgb_dg_create(C.byref(o), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mgb_dg_destroy(self.h)
            self.h = None

    __del__ = close

    def _torch_sync(self):
        import torch
        torch.cuda.synchronize(self.device)

    def sync(self):
        _check(lib().mgb_dg_sync(self.h))

    def m(self):
        return lib().mgb_dg_n_dofs(self.h)

    n = m

    def n_cells(self):
        nc = (C.c_int * 3)()
        _check(lib().mgb_dg_n_cells(self.h, nc))
        return tuple(nc)

    def partition(self):
        """z-slab of this rank: vectors have `n_local` entries, of which [owned_offset, owned_offset + n_owned) are owned"""
        v = (C.c_int64 * 6)()
        _check(lib().mgb_dg_partition(self.h, v))
        return dict(n_global=v[0], n_local=v[1], owned_offset=v[2], n_owned=v[3], first_layer=v[4], n_layers=v[5])

    def exchange_count(self):
        return lib().mgb_dg_exchange_count(self.h)

    def get_penalty(self):
        s = (C.c_double * 3)()
        _check(lib().mgb_dg_penalty(self.h, s))
        return np.array(s)

    def initialize_dof_vector(self):
        import torch
        return torch.zeros(self.m(), dtype=self.dtype, device=self.device)

    def _chk(self, *ts):
        for t in ts:
            if t.dtype != self.dtype or t.numel() != self.m() or not t.is_cuda or not t.is_contiguous():
                raise MgbError("vector does not match the operator (dtype/size/device)")

    def vmult(self, dst, src):
        self._chk(dst, src)
        self._torch_sync()
        _check(lib().mgb_dg_vmult(self.h, self.precision, _ptr(dst), _ptr(src)))
        self.sync()

    def vmult_residual(self, rhs, lhs, residual):
        self._chk(rhs, lhs, residual)
        self._torch_sync()
        _check(lib().mgb_dg_vmult_residual(self.h, self.precision, _ptr(rhs), _ptr(lhs), _ptr(residual)))
        self.sync()

    def vmult_with_cg_update(self, alpha, beta, r, q, p, x):
        self._chk(r, q, p, x)
        sums = (C.c_double * 4)()
        self._torch_sync()
        _check(lib().mgb_dg_vmult_with_cg_update(self.h, self.precision, alpha, beta, _ptr(r), _ptr(q), _ptr(p),
                                                 _ptr(x), sums))
        return list(sums)

    def get_matrix_diagonal_inverse(self):
        d = self.initialize_dof_vector()
        self._torch_sync()
        _check(lib().mgb_dg_inverse_diagonal(self.h, self.precision, _ptr(d)))
        self.sync()
        return d

    def lumped_mass_inverse(self):
        """the diagonal preconditioner of solver_dg/program.cc:134-148: 1 / (1, phi_i)"""
        d = self.initialize_dof_vector()
        self._torch_sync()
        _check(lib().mgb_dg_lumped_mass_inverse(self.h, self.precision, _ptr(d)))
        self.sync()
        return d

    def solve_cg(self, x, b, diagonal_preconditioner, n_iterations, interleaved=True):
        """SolverCG with IterationNumberControl(n_iterations, 1e-20) and a diagonal preconditioner from x = 0
        (solver_dg/program.cc:180-262); interleaved: the vector updates merged with the operator (LaplaceOperatorInterleave,
        :39-70) instead of separate passes.  -> (iterations, last residual norm)"""
        self._chk(x, b, diagonal_preconditioner)
        stats = (C.c_double * 2)()
        self._torch_sync()
        _check(lib().mgb_dg_solve_cg(self.h, self.precision, int(n_iterations), int(interleaved), _ptr(x), _ptr(b),
                                     _ptr(diagonal_preconditioner), stats))
        return int(stats[0]), stats[1]

    def vmult_with_chebyshev_update(self, rhs, iteration_index, factor1, factor2, solution, solution_old):
        """as the reference (laplace_operator_dg.h:907-955): the new iterate lands in `solution_old`; the caller
        swaps (iteration_index 0 writes `solution` directly)"""
        self._chk(rhs, solution, solution_old)
        self._torch_sync()
        _check(lib().mgb_dg_vmult_with_chebyshev_update(self.h, self.precision, _ptr(rhs), iteration_index, factor1,
                                                        factor2, _ptr(solution), _ptr(solution_old)))
        self.sync()

    POINT_JACOBI, JACOBI_TRANSFORMED = 0, 1

    def prolongate_add_cg_to_dg(self, solver, level, dst_dg, src_cg):
        """laplace_operator_dg.h:1863-1894: dst_dg += (FE_Q vector of `level` of `solver` seen from the DG space)"""
        self._torch_sync()
        solver.sync()
        _check(lib().mgb_dg_prolongate_add_cg_to_dg(self.h, solver.h, solver._l(level), self.precision, _ptr(dst_dg),
                                                    _ptr(src_cg)))
        self.sync()

    def vmult_residual_and_restrict_to_cg(self, solver, level, rhs, lhs, restricted_residual):
        """laplace_operator_dg.h:851-860: restricted_residual = P^T (rhs - A lhs), an FE_Q vector"""
        self._torch_sync()
        solver.sync()
        _check(lib().mgb_dg_vmult_residual_and_restrict_to_cg(self.h, solver.h, solver._l(level), self.precision, _ptr(rhs),
                                                              _ptr(lhs), _ptr(restricted_residual)))
        self.sync()

    def set_preconditioner(self, kind):
        """preconditioner of vmult_with_chebyshev_update: POINT_JACOBI or JACOBI_TRANSFORMED
        (multigrid::JacobiTransformed, laplace_operator_dg.h:2028-2256)"""
        self._torch_sync()
        _check(lib().mgb_dg_set_preconditioner(self.h, kind))

    def preconditioner_vmult(self, dst, src):
        """JacobiTransformed::vmult / point-Jacobi scaling: dst = P^-1 src"""
        self._chk(dst, src)
        self._torch_sync()
        _check(lib().mgb_dg_preconditioner_vmult(self.h, self.precision, _ptr(dst), _ptr(src)))
        self.sync()

    def vmult_async(self, dst, src):
        """no synchronisation (benchmark loops)"""
        _check(lib().mgb_dg_vmult(self.h, self.precision, _ptr(dst), _ptr(src)))

    def chebyshev_update_async(self, rhs, iteration_index, factor1, factor2, solution, solution_old):
        _check(lib().mgb_dg_vmult_with_chebyshev_update(self.h, self.precision, _ptr(rhs), iteration_index, factor1,
                                                        factor2, _ptr(solution), _ptr(solution_old)))

    def vmult_host(self, dst_host, src_host):
        ptr = lambda a: C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        _check(lib().mgb_dg_vmult_host(self.h, self.precision, ptr(dst_host), ptr(src_host)))

    def timer_start(self):
        _check(lib().mgb_dg_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_double()
        _check(lib().mgb_dg_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def launch_count(self):
        return lib().mgb_dg_launch_count(self.h)


class MultigridSolverDG:
    """multigrid::MultigridSolverDG<3,p,float,double> (common/multigrid_solver_dg.h:44-735): CG in double on the
    SIP-DG operator preconditioned by one float cycle (Chebyshev/JacobiTransformed on the DG space, same-degree
    transfer to FE_Q, the FE_Q V-cycle below).  Problem data of poisson_dg/program.cc:70-135."""

    JACOBI, JACOBI_TRANSFORMED = 0, 1

    def __init__(self, degree=4, n_subdivisions=(1, 1, 1), n_refinements=3, lower=(-1.,) * 3, upper=(1.,) * 3,
                 degree_pre=3, degree_post=3, preconditioner=1, device=0, comm=None, basis=DG_BASIS_GAUSS_LOBATTO):
        """basis: DG_BASIS_* (poisson_dg/program.cc:174 uses FE_DGQHermite = DG_BASIS_HERMITE).
        comm = (rank, size, unique_id bytes) for one process per GPU (comm_from_torch_distributed()): the DG
        operator and the finest FE_Q level share the z-slabs; vectors are the local DG vectors (partition())"""
        import torch
        o = DGSolverOptions()
        lib().mgb_dg_solver_default_options(C.byref(o))
        if comm is not None and comm[1] > 1:
            o.comm_rank, o.comm_size = comm[0], comm[1]
            o.comm_id[:] = list(comm[2])
        o.degree = degree
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        o.n_subdivisions[:] = list(n_subdivisions)
        o.n_refinements = n_refinements
        o.lower[:] = list(lower)
        o.upper[:] = list(upper)
        o.degree_pre, o.degree_post, o.preconditioner, o.device = degree_pre, degree_post, preconditioner, device
        o.basis = basis
        self.h = C.c_void_p()
        self.device = torch.device("cuda", device)
        self.degree = degree
        _check(lib().mgb_dg_solver_create(C.byref(o), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mgb_dg_solver_destroy(self.h)
            self.h = None

    __del__ = close

    def n_dofs(self):
        return lib().mgb_dg_solver_n_dofs(self.h)

    def sync(self):
        _check(lib().mgb_dg_solver_sync(self.h))

    def solve_cg(self, tolerance=1e-3):
        """-> (log(tol)/log(rate), rate) as the reference; `last_step` holds the iteration count"""
        est, rate, its = C.c_double(), C.c_double(), C.c_uint()
        _check(lib().mgb_dg_solver_solve_cg(self.h, tolerance, C.byref(est), C.byref(rate), C.byref(its)))
        self.last_step = its.value
        return est.value, rate.value

    def solve_cg_host(self, tolerance, solution_host=None):
        est, rate, its = C.c_double(), C.c_double(), C.c_uint()
        ptr = None if solution_host is None else C.c_void_p(
            solution_host.data_ptr() if hasattr(solution_host, "data_ptr") else solution_host.ctypes.data)
        _check(lib().mgb_dg_solver_solve_cg_host(self.h, tolerance, ptr, C.byref(est), C.byref(rate), C.byref(its)))
        self.last_step = its.value
        return est.value, rate.value

    def residual_history(self):
        hist = np.zeros(128)
        cnt = C.c_int()
        _check(lib().mgb_dg_solver_residual_history(self.h, hist.ctypes.data_as(_dp), 128, C.byref(cnt)))
        return hist[:cnt.value]

    def vmult(self, dst, src):
        """one cycle as preconditioner: double device tensors"""
        import torch
        assert dst.dtype == torch.float64 and src.dtype == torch.float64 and dst.numel() == src.numel() == self.n_dofs()
        torch.cuda.synchronize(self.device)
        _check(lib().mgb_dg_solver_vmult(self.h, C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr())))
        self.sync()

    def vmult_with_residual_update(self, residual, update, factor):
        """the preconditioner call of CG with merged vector operations (double device tensors, both updated in place)
        -> (update_new . residual_new, update_new . (factor * update_old))"""
        import torch
        assert residual.dtype == torch.float64 and update.dtype == torch.float64
        assert residual.numel() == update.numel() == self.n_dofs()
        torch.cuda.synchronize(self.device)
        out = np.zeros(2)
        _check(lib().mgb_dg_solver_vmult_with_residual_update(self.h, C.c_void_p(residual.data_ptr()), C.c_void_p(update.data_ptr()),
                                                   float(factor), out.ctypes.data_as(_dp)))
        return out[0], out[1]

    def compute_l2_error(self):
        e = C.c_double()
        _check(lib().mgb_dg_solver_l2_error(self.h, C.byref(e)))
        return e.value

    def get_solution(self):
        out = np.zeros(self.n_dofs())
        _check(lib().mgb_dg_solver_get_solution(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def get_vector(self, which):
        """host copy of 0 solution, 1 rhs, 2 residual (double), 3 defect_dg, 4 solution_update_dg (float)"""
        p = C.c_void_p()
        self.sync()
        _check(lib().mgb_dg_solver_vector(self.h, which, C.byref(p)))
        out = np.zeros(self.n_dofs(), dtype=np.float64 if which < 3 else np.float32)
        fe, _ = self.parts()
        _check(lib().mgb_copy_to_host(fe, C.c_void_p(out.ctypes.data), p, out.nbytes))
        return out

    def smoother_info(self):
        info = np.zeros(6)
        _check(lib().mgb_dg_solver_smoother_info(self.h, info.ctypes.data_as(_dp)))
        return dict(min_eig=info[0], max_eig=info[1], theta=info[2], delta=info[3], degree=int(info[4]),
                    cg_iterations=int(info[5]))

    def parts(self):
        """(handle of the FE_Q hierarchy, handle of the DG operator), owned by this solver"""
        fe, dg = C.c_void_p(), C.c_void_p()
        _check(lib().mgb_dg_solver_parts(self.h, C.byref(fe), C.byref(dg)))
        return fe, dg

    def partition(self):
        """z-slab of this rank (as LaplaceOperatorDG.partition())"""
        v = (C.c_int64 * 6)()
        _check(lib().mgb_dg_partition(self.parts()[1], v))
        return dict(n_global=v[0], n_local=v[1], owned_offset=v[2], n_owned=v[3], first_layer=v[4], n_layers=v[5])

    def fe_smoother_info(self, level):
        fe, _ = self.parts()
        info = np.zeros(6)
        _check(lib().mgb_smoother_info(fe, level, info.ctypes.data_as(_dp)))
        return dict(min_eig=info[0], max_eig=info[1], theta=info[2], delta=info[3], degree=int(info[4]),
                    cg_iterations=int(info[5]))

    def do_matvec(self):
        _check(lib().mgb_dg_solver_do_matvec(self.h))

    def do_matvec_smoother(self):
        _check(lib().mgb_dg_solver_do_matvec_smoother(self.h))

    def launch_count(self):
        return lib().mgb_dg_solver_launch_count(self.h)


class MultigridSolverDGPlain:
    """multigrid::MultigridSolverDGPlain<3,p,float,double,type> (common/multigrid_solver_dg_plain.h:54-590): CG in
    double on the SIP-DG operator preconditioned by one float V-cycle whose levels are the DG spaces of the refinement
    hierarchy (Chebyshev/JacobiTransformed smoothers, cell-wise embedding as transfer).  Problem data and mesh of
    poisson_dg_plain/program.cc:70-135,356-409.  One GPU."""

    def __init__(self, degree=4, n_subdivisions=(1, 1, 1), n_refinements=3, lower=(-1.,) * 3, upper=(1.,) * 3,
                 degree_pre=3, degree_post=3, device=0, basis=DG_BASIS_HERMITE):
        """basis: DG_BASIS_* = the reference's template argument `type` (poisson_dg_plain/program.cc:420-446 runs all three)"""
        import torch
        o = DGSolverOptions()
        lib().mgb_dg_solver_default_options(C.byref(o))
        o.degree = degree
        if isinstance(n_subdivisions, int):
            n_subdivisions = (n_subdivisions,) * 3
        o.n_subdivisions[:] = list(n_subdivisions)
        o.n_refinements = n_refinements
        o.lower[:] = list(lower)
        o.upper[:] = list(upper)
        o.degree_pre, o.degree_post, o.preconditioner, o.device = degree_pre, degree_post, 1, device
        o.basis = basis
        self.h = C.c_void_p()
        self.device = torch.device("cuda", device)
        self.degree = degree
        _check(lib().mgb_dg_plain_solver_create(C.byref(o), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mgb_dg_plain_solver_destroy(self.h)
            self.h = None

    __del__ = close

    def n_dofs(self):
        return lib().mgb_dg_plain_solver_n_dofs(self.h)

    def n_levels(self):
        return lib().mgb_dg_plain_solver_n_levels(self.h)

    def level_n_dofs(self, level):
        op = C.c_void_p()
        _check(lib().mgb_dg_plain_solver_level_operator(self.h, level, C.byref(op)))
        return lib().mgb_dg_n_dofs(op)

    def sync(self):
        _check(lib().mgb_dg_plain_solver_sync(self.h))

    def solve_cg(self, tolerance=1e-3):
        """-> (log(tol)/log(rate), rate) as the reference; `last_step` holds the iteration count"""
        est, rate, its = C.c_double(), C.c_double(), C.c_uint()
        _check(lib().mgb_dg_plain_solver_solve_cg(self.h, tolerance, C.byref(est), C.byref(rate), C.byref(its)))
        self.last_step = its.value
        return est.value, rate.value

    def solve_cg_host(self, tolerance, solution_host=None):
        est, rate, its = C.c_double(), C.c_double(), C.c_uint()
        ptr = None if solution_host is None else C.c_void_p(
            solution_host.data_ptr() if hasattr(solution_host, "data_ptr") else solution_host.ctypes.data)
        _check(lib().mgb_dg_plain_solver_solve_cg_host(self.h, tolerance, ptr, C.byref(est), C.byref(rate), C.byref(its)))
        self.last_step = its.value
        return est.value, rate.value

    def residual_history(self):
        hist = np.zeros(128)
        cnt = C.c_int()
        _check(lib().mgb_dg_plain_solver_residual_history(self.h, hist.ctypes.data_as(_dp), 128, C.byref(cnt)))
        return hist[:cnt.value]

    def vmult(self, dst, src):
        """one V-cycle as preconditioner: double device tensors"""
        import torch
        assert dst.dtype == torch.float64 and src.dtype == torch.float64 and dst.numel() == src.numel() == self.n_dofs()
        torch.cuda.synchronize(self.device)
        _check(lib().mgb_dg_plain_solver_vmult(self.h, C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr())))
        self.sync()

    def _transfer(self, fn, level, dst, src, n_dst, n_src, add):
        import torch
        assert dst.dtype == torch.float32 and src.dtype == torch.float32
        assert dst.numel() == n_dst and src.numel() == n_src and dst.is_cuda and src.is_cuda
        torch.cuda.synchronize(self.device)
        _check(fn(self.h, level, C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr()), int(add)))
        self.sync()

    def prolongate(self, to_level, dst, src, add=False):
        """dst (level to_level) = or += P src (level to_level - 1), float device tensors"""
        self._transfer(lib().mgb_dg_plain_solver_prolongate, to_level, dst, src, self.level_n_dofs(to_level),
                       self.level_n_dofs(to_level - 1), add)

    def restrict(self, from_level, dst, src, add=False):
        """dst (level from_level - 1) = or += P^T src (level from_level), float device tensors"""
        self._transfer(lib().mgb_dg_plain_solver_restrict, from_level, dst, src, self.level_n_dofs(from_level - 1),
                       self.level_n_dofs(from_level), add)

    def vmult_with_residual_update(self, residual, update, factor):
        """the preconditioner call of CG with merged vector operations (double device tensors, both updated in place)
        -> (update_new . residual_new, update_new . (factor * update_old))"""
        import torch
        assert residual.dtype == torch.float64 and update.dtype == torch.float64
        assert residual.numel() == update.numel() == self.n_dofs()
        torch.cuda.synchronize(self.device)
        out = np.zeros(2)
        _check(lib().mgb_dg_plain_solver_vmult_with_residual_update(self.h, C.c_void_p(residual.data_ptr()), C.c_void_p(update.data_ptr()),
                                                   float(factor), out.ctypes.data_as(_dp)))
        return out[0], out[1]

    def compute_l2_error(self):
        e = C.c_double()
        _check(lib().mgb_dg_plain_solver_l2_error(self.h, C.byref(e)))
        return e.value

    def get_solution(self):
        out = np.zeros(self.n_dofs())
        _check(lib().mgb_dg_plain_solver_get_solution(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def get_vector(self, which, level=None):
        """host copy of 0 solution, 1 rhs, 2 residual (double, finest level), 3 defect[level], 4 solution_update[level],
        5 t[level] (float)"""
        level = self.n_levels() - 1 if level is None else level
        out = np.zeros(self.n_dofs() if which < 3 else self.level_n_dofs(level), dtype=np.float64 if which < 3 else np.float32)
        _check(lib().mgb_dg_plain_solver_get_vector(self.h, which, level, C.c_void_p(out.ctypes.data)))
        return out

    def smoother_info(self, level):
        info = np.zeros(6)
        _check(lib().mgb_dg_plain_solver_smoother_info(self.h, level, info.ctypes.data_as(_dp)))
        return dict(min_eig=info[0], max_eig=info[1], theta=info[2], delta=info[3], degree=int(info[4]),
                    cg_iterations=int(info[5]))

    def do_matvec(self):
        _check(lib().mgb_dg_plain_solver_do_matvec(self.h))

    def do_matvec_smoother(self):
        _check(lib().mgb_dg_plain_solver_do_matvec_smoother(self.h))

    def launch_count(self):
        return lib().mgb_dg_plain_solver_launch_count(self.h)


class LaplaceOperatorAdaptive:
    """FE_Q(p) Laplacian on an adaptively refined box forest with hanging nodes (the operator of
    poisson_l/program.cc:232-260 on the meshes of :465-541).  Mesh and numbering live on the host; vmult and the
    diagonal are kernels.  `root_mask[z][y][x]` selects root cells (hyper_L: 2x2x2 with one root removed)."""

    def __init__(self, degree, n_roots=(1, 1, 1), root_mask=None, lower=(-1.,) * 3, upper=(1.,) * 3, n_global_refinements=0,
                 device=0):
        nr = (C.c_int * 3)(*n_roots)
        lo = (C.c_double * 3)(*lower)
        hi = (C.c_double * 3)(*upper)
        mask = None
        if root_mask is not None:
            m = np.ascontiguousarray(root_mask, dtype=np.uint8).ravel()
            assert m.size == n_roots[0] * n_roots[1] * n_roots[2]
            mask = m.ctypes.data_as(C.POINTER(C.c_ubyte))
        self.h = C.c_void_p()
        self.degree, self.device_index = degree, device
        self.lower, self.upper, self.n_roots = tuple(lower), tuple(upper), tuple(n_roots)
        _check(lib().mgb_adaptive_create(degree, nr, mask, lo, hi, n_global_refinements, device, C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mgb_adaptive_destroy(self.h)
            self.h = None

    __del__ = close

    def n_cells(self):
        return lib().mgb_adaptive_n_cells(self.h)

    def cells(self):
        """(level[n_cells], position[n_cells, 3]) of the active cells, in the library's order"""
        n = self.n_cells()
        level = np.zeros(n, dtype=np.int32)
        pos = np.zeros((n, 3), dtype=np.int32)
        _check(lib().mgb_adaptive_cells(self.h, level.ctypes.data_as(_ip), pos.ctypes.data_as(_ip)))
        return level, pos

    def cell_boxes(self):
        """lower corner and size of every active cell, shape (n_cells, 3) each"""
        level, pos = self.cells()
        H = (np.array(self.upper) - np.array(self.lower)) / np.array(self.n_roots)
        size = H[None, :] / (2.0 ** level)[:, None]
        return np.array(self.lower)[None, :] + pos * size, size

    def refine(self, flags):
        f = np.ascontiguousarray(flags, dtype=np.uint8)
        assert f.size == self.n_cells()
        _check(lib().mgb_adaptive_refine(self.h, f.ctypes.data_as(C.POINTER(C.c_ubyte))))

    def m(self):
        n = lib().mgb_adaptive_n_dofs(self.h)
        if n < 0:
            raise MgbError(lib().mgb_last_error().decode())
        return n

    n_dofs = m

    def info(self):
        v = (C.c_int64 * 6)()
        _check(lib().mgb_adaptive_info(self.h, v))
        return dict(n_dofs=v[0], n_dirichlet=v[1], hanging_local_nodes=v[2], constraint_rows=v[3],
                    constraint_entries=v[4], max_level=v[5])

    def dof_points(self):
        out = np.zeros((self.m(), 3))
        _check(lib().mgb_adaptive_dof_points(self.h, out.ctypes.data_as(_dp)))
        return out

    def dirichlet_mask(self):
        out = np.zeros(self.m(), dtype=np.uint8)
        _check(lib().mgb_adaptive_dirichlet(self.h, out.ctypes.data_as(C.POINTER(C.c_ubyte))))
        return out.astype(bool)

    def tables(self):
        inf = self.info()
        n3 = (self.degree + 1) ** 3
        ci = np.zeros((self.n_cells(), n3), dtype=np.int32)
        rs = np.zeros(inf["constraint_rows"] + 1, dtype=np.int32)
        rd = np.zeros(max(inf["constraint_entries"], 1), dtype=np.int32)
        rw = np.zeros(max(inf["constraint_entries"], 1))
        i32 = C.POINTER(C.c_int32)
        _check(lib().mgb_adaptive_tables(self.h, ci.ctypes.data_as(i32), rs.ctypes.data_as(i32), rd.ctypes.data_as(i32),
                                         rw.ctypes.data_as(_dp)))
        return ci, rs, rd[:inf["constraint_entries"]], rw[:inf["constraint_entries"]]

    def sync(self):
        _check(lib().mgb_adaptive_sync(self.h))

    # ---- local-smoothing multigrid and CG (poisson_l/program.cc:283-420)
    MG_INFO = ("n_dofs", "n_cells", "n_edge_dofs", "n_edge_cells", "n_copy", "n_constrained", "degree", "min_eig", "max_eig",
               "theta", "delta", "cg_iterations")

    def setup_multigrid(self, n_smoother=3):
        """level operators, edge matrices, transfer and Chebyshev smoothers on the current mesh (redo after refine)"""
        _check(lib().mgb_adaptive_mg_setup(self.h, int(n_smoother)))
        self.n_levels = lib().mgb_adaptive_mg_n_levels(self.h)

    def mg_level_info(self, level):
        v = np.zeros(12)
        _check(lib().mgb_adaptive_mg_level_info(self.h, level, v.ctypes.data_as(_dp)))
        return dict(zip(self.MG_INFO, v))

    def mg_level_vmult(self, level, which, dst, src):
        """which 0: matrix[level].vmult, 1: vmult_interface_down, 2: vmult_interface_up, on float level vectors"""
        import torch
        assert dst.dtype == torch.float32 and src.dtype == torch.float32
        n = int(self.mg_level_info(level)["n_dofs"])
        assert dst.numel() == src.numel() == n
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_mg_level_vmult(self.h, int(level), int(which), _ptr(dst), _ptr(src)))

    def mg_vmult(self, dst, src):
        """PreconditionMG::vmult: one local-smoothing V-cycle (float) on a double vector of the active mesh"""
        import torch
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_mg_vmult(self.h, _ptr(dst), _ptr(src)))
        self.sync()

    def solve_cg(self, x, rhs, rel_tol=1e-9):
        """SolverCG (double) preconditioned by the V-cycle, zero start vector; returns (iterations, rate, history)"""
        import torch
        torch.cuda.synchronize()
        its, rate = C.c_uint(), C.c_double()
        hist = np.zeros(1024)
        nh = C.c_int(hist.size)
        _check(lib().mgb_adaptive_solve_cg(self.h, _ptr(x), _ptr(rhs), float(rel_tol), C.byref(its), C.byref(rate),
                                           hist.ctypes.data_as(_dp), C.byref(nh)))
        return its.value, rate.value, hist[:nh.value]

    def vmult(self, dst, src, precision=F64):
        import torch
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_vmult(self.h, precision, _ptr(dst), _ptr(src)))
        self.sync()

    def get_matrix_diagonal_inverse(self, out, precision=F64):
        import torch
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_inverse_diagonal(self.h, precision, _ptr(out)))
        self.sync()

    # ---- the adaptive loop around the solver (poisson_l/program.cc:497-602)
    FUNCTION_ZERO, FUNCTION_L_SINGULARITY = 0, 1
    NORM_L2, NORM_H1_SEMINORM = 0, 1

    def refine_and_coarsen(self, refine_flags, coarsen_flags=None):
        """execute_coarsening_and_refinement with explicit flags (one per active cell)"""
        r = np.ascontiguousarray(refine_flags, dtype=np.uint8)
        c = None if coarsen_flags is None else np.ascontiguousarray(coarsen_flags, dtype=np.uint8)
        assert r.size == self.n_cells() and (c is None or c.size == r.size)
        ub = C.POINTER(C.c_ubyte)
        _check(lib().mgb_adaptive_refine_and_coarsen(self.h, r.ctypes.data_as(ub), None if c is None else c.ctypes.data_as(ub)))

    def refine_and_coarsen_fixed_number(self, indicators, top_fraction, bottom_fraction):
        """GridRefinement::refine_and_coarsen_fixed_number + execute_coarsening_and_refinement (:532-541)"""
        e = np.ascontiguousarray(indicators, dtype=np.float32)
        assert e.size == self.n_cells()
        _check(lib().mgb_adaptive_refine_and_coarsen_fixed_number(self.h, e.ctypes.data_as(C.POINTER(C.c_float)),
                                                                  float(top_fraction), float(bottom_fraction)))

    def face_neighbours(self):
        t = np.zeros((self.n_cells(), 6), dtype=np.int8)
        ids = np.zeros((self.n_cells(), 6, 4), dtype=np.int32)
        _check(lib().mgb_adaptive_face_neighbours(self.h, t.ctypes.data_as(C.POINTER(C.c_byte)),
                                                  ids.ctypes.data_as(C.POINTER(C.c_int32))))
        return t, ids

    def plain_table(self):
        """cell -> DoF table that keeps the Dirichlet DoFs (rows refer to the CSR list of tables())"""
        ci = np.zeros((self.n_cells(), (self.degree + 1) ** 3), dtype=np.int32)
        _check(lib().mgb_adaptive_plain_table(self.h, ci.ctypes.data_as(C.POINTER(C.c_int32))))
        return ci

    def interpolate_boundary_values(self, u, function=1):
        import torch
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_interpolate_boundary_values(self.h, int(function), _ptr(u)))

    def compute_residual(self, dst, src, f=0.):
        """dst = (f, phi) - K src on the rows without Dirichlet condition; src carries the boundary values"""
        import torch
        torch.cuda.synchronize()
        _check(lib().mgb_adaptive_compute_residual(self.h, _ptr(dst), _ptr(src), float(f)))
        self.sync()

    def estimate_kelly(self, u):
        """KellyErrorEstimator::estimate (:524-531) -> float32 indicator per active cell"""
        import torch
        torch.cuda.synchronize()
        out = np.zeros(self.n_cells(), dtype=np.float32)
        _check(lib().mgb_adaptive_estimate_kelly(self.h, _ptr(u), out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def integrate_difference(self, u, function=1, norm=0, n_q_points=None, per_cell=False):
        """VectorTools::integrate_difference (:556-575) -> norm (and the cells' squared contributions)"""
        import torch
        torch.cuda.synchronize()
        nq = self.degree + (2 if norm == 0 else 1) if n_q_points is None else n_q_points
        res = C.c_double()
        sq = np.zeros(self.n_cells()) if per_cell else None
        _check(lib().mgb_adaptive_integrate_difference(self.h, _ptr(u), int(function), int(norm), int(nq), C.byref(res),
                                                       sq.ctypes.data_as(_dp) if per_cell else None))
        return (res.value, sq) if per_cell else res.value

    def solve_cycle(self, function=1, f=0., n_smoother=3):
        """one cycle of LaplaceProblem::run on the current mesh (:545-575) -> dict(iterations, rate, l2_error, h1_error),
        the solution as a device tensor view owned by the operator (valid until the mesh changes)"""
        import torch
        stats = np.zeros(4)
        ptr = C.c_void_p()
        _check(lib().mgb_adaptive_solve_cycle(self.h, int(function), float(f), int(n_smoother), stats.ctypes.data_as(_dp),
                                              C.byref(ptr)))
        self.sync()
        out = dict(iterations=int(stats[0]), rate=stats[1], l2_error=stats[2], h1_error=stats[3], n_cells=self.n_cells(),
                   n_dofs=self.m())
        self._solution_ptr = ptr.value
        return out

    def solution(self):
        """host copy of the solution of the last solve_cycle"""
        out = np.zeros(self.m())
        _check(lib().mgb_copy_to_host(None, C.c_void_p(out.ctypes.data), C.c_void_p(self._solution_ptr), out.nbytes))
        return out

    def run_adaptive_loop(self, n_cycles, function=1, f=0., n_smoother=3, top_fraction=0.3, bottom_fraction=0.03,
                          max_size=None):
        """LaplaceProblem::run (:497-602) from the current mesh: solve, estimate, refine / coarsen, n_cycles times"""
        table = []
        for cycle in range(n_cycles):
            if cycle > 0:
                import torch
                u = torch.tensor(self.solution(), device="cuda")
                eta = self.estimate_kelly(u)
                self.refine_and_coarsen_fixed_number(eta, top_fraction, bottom_fraction)
            table.append(self.solve_cycle(function, f, n_smoother))
            if max_size is not None and table[-1]["n_dofs"] > max_size:
                break
        return table

    def vmult_host(self, dst_host, src_host, precision=F64):
        _check(lib().mgb_adaptive_vmult_host(self.h, precision, C.c_void_p(dst_host.ctypes.data),
                                             C.c_void_p(src_host.ctypes.data)))

    def launch_count(self):
        return lib().mgb_adaptive_launch_count(self.h)
