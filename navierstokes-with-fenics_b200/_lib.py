"""ctypes binding of libnsb200.so (include/nsb200.h).  There is no CPU fallback: a missing
library or a missing GPU raises, loudly."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnsb200.so")

# constants of include/nsb200.h
VEC, MAT, PICARD, BDF, SYMBC = 1, 2, 4, 8, 16
PAT_P2P2, PAT_P1P1, PAT_WW = 0, 1, 2
CSR, BSR = 0, 1
BC_LINEAR, BC_NEWTON = 0, 1
# entries per TMA tile of the bulk-copy SpMV kernels, keyed by (matrix kind, bs) (csrc/spmv_tma.cuh)
TILE_ENTRIES = {(BSR, 3): 384, (BSR, 2): 768}
SCHED_MAX_BLOCKS = 160     # entries of a row group in shared memory (csrc/assemble_jac.cu JacCfg::MAXB)
TIMER_CLASSES = ("asm_jacobian", "asm_residual", "asm_constant", "asm_rhs", "spmv_bsr", "spmv_csr",
                 "krylov_vec", "dirichlet", "vector", "comm", "spmv_bsr_f32", "cg_persistent")
HALO_NONE, HALO_NODES, HALO_VERTS = 0, 1, 2

c_i32p = C.POINTER(C.c_int32)
c_u8p = C.POINTER(C.c_uint8)
c_f64p = C.POINTER(C.c_double)
c_i64p = C.POINTER(C.c_int64)
vp = C.c_void_p


class Sell(C.Structure):
    """struct nsb_sell"""
    _fields_ = [("nslices", C.c_int32), ("reserved", C.c_int32), ("slice_ptr", vp), ("col", vp), ("vals", vp),
                ("vals32", vp), ("col16", vp)]


class Matrix(C.Structure):
    """struct nsb_matrix"""
    _fields_ = [("kind", C.c_int32), ("bs", C.c_int32), ("nrows", C.c_int32), ("nnz", C.c_int32),
                ("rowptr", vp), ("col", vp), ("vals", vp), ("bc_mask", vp), ("tiles", vp),
                ("ntiles", C.c_int32), ("halo", C.c_int32), ("vals32", vp), ("diag32", vp), ("sell", C.POINTER(Sell))]


class Saddle(C.Structure):
    """struct nsb_saddle"""
    _fields_ = [("n_u", C.c_int32), ("n_p", C.c_int32), ("Mp", Matrix), ("Lp", Matrix), ("Mp_inv_diag", vp),
                ("Lp_inv_diag", vp), ("p_identity", vp), ("c_mass", C.c_double), ("c_lap", C.c_double),
                ("lp_nullspace", C.c_int32), ("vel_degree", C.c_int32), ("vel_lambda_min", C.c_double),
                ("vel_lambda_max", C.c_double), ("lp_degree", C.c_int32), ("lp_lambda_min", C.c_double),
                ("lp_lambda_max", C.c_double), ("inner_rtol", C.c_double), ("inner_maxit", C.c_int32),
                ("reserved", C.c_int32), ("work", vp), ("inner_iterations", C.c_int64)]


class Precond(C.Structure):
    """struct nsb_precond"""
    _fields_ = [("kind", C.c_int32), ("degree", C.c_int32), ("inv_diag", vp), ("lambda_min", C.c_double),
                ("lambda_max", C.c_double), ("lowp", C.c_int32), ("reserved", C.c_int32), ("saddle", C.POINTER(Saddle))]


PC_NONE, PC_JACOBI, PC_CHEBYSHEV, PC_SADDLE = 0, 1, 2, 3

# name -> (restype, argtypes); every symbol include/nsb200.h declares
PROTOTYPES = {
    "nsb_version": (C.c_int, []),
    "nsb_create": (C.c_int, [C.POINTER(vp), C.c_int]),
    "nsb_destroy": (C.c_int, [vp]),
    "nsb_last_error": (C.c_char_p, [vp]),
    "nsb_pattern_build": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, vp, C.c_int32, vp, C.POINTER(vp)]),
    "nsb_pattern_sizes": (C.c_int, [vp, c_i64p, c_i64p, c_i32p]),
    "nsb_pattern_fill": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "nsb_pattern_free": (C.c_int, [vp]),
    "nsb_tiles_build": (C.c_int, [C.c_int32, vp, C.c_int32, c_i32p, vp]),
    "nsb_set_mesh": (C.c_int, [vp, C.c_int32, C.c_int32, vp, C.c_int32, vp, C.c_int32, vp]),
    "nsb_set_pattern": (C.c_int, [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "nsb_sched_build": (C.c_int, [C.c_int32, vp, vp, vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, vp, vp]),
    "nsb_set_schedule": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "nsb_set_scratch": (C.c_int, [vp, vp, C.c_int64]),
    "nsb_scratch_doubles_per_cell": (C.c_int, [C.c_int32]),
    "nsb_set_mixed_dofmap": (C.c_int, [vp, C.c_int32, vp]),
    "nsb_set_coefficients": (C.c_int, [vp, c_f64p]),
    "nsb_set_forms": (C.c_int, [vp, C.c_int32, C.c_int32, c_f64p, C.c_int32]),
    "nsb_assemble_ipcs_momentum": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, vp, vp, vp]),
    "nsb_assemble_poisson": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, vp]),
    "nsb_assemble_correction": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "nsb_assemble_p1_mass": (C.c_int, [vp, C.c_int32, vp, vp, vp]),
    "nsb_assemble_mixed": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "nsb_bc_apply_matrix": (C.c_int, [vp, C.c_int32, C.c_int32, vp, vp, vp, C.c_int32, vp, vp]),
    "nsb_bc_apply_vector": (C.c_int, [vp, C.c_int32, vp, vp, C.c_int32, vp, vp, vp]),
    "nsb_set_mixed_bc": (C.c_int, [vp, vp, vp]),
    "nsb_spmv": (C.c_int, [vp, C.POINTER(Matrix), vp, vp, vp]),
    "nsb_sell_fill": (C.c_int, [vp, C.POINTER(Matrix), C.c_int32, vp, vp, vp, vp, vp, c_i32p, vp]),
    "nsb_inv_diagonal": (C.c_int, [vp, C.POINTER(Matrix), vp, vp]),
    "nsb_cg_workspace_doubles": (C.c_int64, [C.c_int64]),
    "nsb_gmres_workspace_doubles": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32]),
    "nsb_lambda_max": (C.c_int, [vp, C.POINTER(Matrix), vp, C.c_int32, C.c_int32, vp, c_f64p, vp]),
    "nsb_cell_vorticity": (C.c_int, [vp, vp, vp, vp]),
    "nsb_cell_pressure_gradient": (C.c_int, [vp, vp, vp, vp]),
    "nsb_rhs_guess": (C.c_int, [vp, C.c_int64, C.c_int32, vp, vp, vp, vp]),
    "nsb_rhs_update": (C.c_int, [vp, C.POINTER(Matrix), C.POINTER(C.c_int32), C.c_int32, vp, vp, vp, vp, vp]),
    "nsb_lambda_max32": (C.c_int, [vp, C.POINTER(Matrix), vp, C.c_int32, C.c_int32, vp, c_f64p, vp]),
    "nsb_saddle_lambda_max": (C.c_int, [vp, C.POINTER(Matrix), vp, C.c_int32, C.c_int32, vp, c_f64p, vp]),
    "nsb_cg": (C.c_int, [vp, C.POINTER(Matrix), C.POINTER(Precond), vp, vp, C.c_double, C.c_double, C.c_int32, C.c_int32,
                         vp, c_i32p, c_f64p, vp]),
    "nsb_gmres": (C.c_int, [vp, C.POINTER(Matrix), C.POINTER(Precond), vp, vp, C.c_double, C.c_double, C.c_int32, C.c_int32,
                            C.c_int32, vp, c_i32p, c_f64p, vp]),
    "nsb_assemble_cell_rhs": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp, vp]),
    "nsb_cell_error_norm2": (C.c_int, [vp, C.c_int32, C.c_int32, vp, C.c_int32, vp, C.c_int32, vp, vp, vp, vp, c_f64p, vp]),
    "nsb_cfl_max": (C.c_int, [vp, vp, C.c_double, C.c_int32, vp, vp, c_f64p, vp]),
    "nsb_vec_shift": (C.c_int, [vp, C.c_int64, C.c_double, vp, vp]),
    "nsb_vec_copy": (C.c_int, [vp, C.c_int64, vp, vp, vp]),
    "nsb_vec_axpy": (C.c_int, [vp, C.c_int64, C.c_double, vp, vp, vp]),
    "nsb_vec_scale": (C.c_int, [vp, C.c_int64, C.c_double, vp, vp]),
    "nsb_vec_to_f32": (C.c_int, [vp, C.c_int64, vp, vp, vp]),
    "nsb_vec_dot": (C.c_int, [vp, C.c_int64, vp, vp, c_f64p, vp]),
    "nsb_vec_sum": (C.c_int, [vp, C.c_int64, vp, c_f64p, vp]),
    "nsb_vec_scatter": (C.c_int, [vp, C.c_int64, vp, vp, vp, vp]),
    "nsb_vec_gather": (C.c_int, [vp, C.c_int64, vp, vp, vp, vp]),
    "nsb_ipcs_advance": (C.c_int, [vp, C.c_int64, vp, vp, vp, C.c_int64, vp, vp, vp]),
    "nsb_fp64_peak": (C.c_int, [vp, c_f64p]),
    "nsb_boundary_force": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, C.c_double, c_f64p, vp]),
    "nsb_direct_solve": (C.c_int, [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, C.c_int32, c_f64p, vp]),
    "nsb_direct_seconds": (C.c_int, [vp, c_f64p, c_i64p, C.c_int32]),
    "nsb_comm_unique_id": (C.c_int, [vp]),
    "nsb_comm_init": (C.c_int, [vp, C.c_int32, C.c_int32, vp]),
    "nsb_comm_finalize": (C.c_int, [vp]),
    "nsb_set_halo": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "nsb_static_scratch_doubles": (C.c_int64, []),
    "nsb_assemble_ipcs_momentum_residual_noalloc": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "nsb_bsr_block_scalar_f32": (C.c_int, [vp, C.c_int64, C.c_int32, vp, vp, vp]),
    "nsb_set_block_scalar_out": (C.c_int, [vp, vp, C.c_int64]),
    "nsb_bsr_block_scalar_rows_f32": (C.c_int, [vp, C.c_int32, vp, vp, C.c_int32, vp, vp, vp]),
    "nsb_p2p_export": (C.c_int, [vp, C.c_int32, vp]),
    "nsb_p2p_connect": (C.c_int, [vp, C.c_int32, C.c_int32, vp, vp, vp, vp]),
    "nsb_p2p_status": (C.c_int, [vp, c_i32p]),
    "nsb_p2p_disconnect": (C.c_int, [vp]),
    "nsb_allreduce": (C.c_int, [vp, vp, C.c_int64, vp]),
    "nsb_halo_exchange": (C.c_int, [vp, C.c_int32, C.c_int32, vp, vp, vp]),
    "nsb_timer_enable": (C.c_int, [vp, C.c_int32]),
    "nsb_timer_report": (C.c_int, [vp, c_f64p, c_i64p]),
    "nsb_launch_count": (C.c_int, [vp, c_i64p]),
}

_lib = None


def load():
    """dlopen libnsb200.so and attach the prototypes (raises if the library is not built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libnsb200.so is not built (%s): run __graft_entry__.build(); there is no CPU fallback"
                % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)     # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error(ctx=None):
    msg = load().nsb_last_error(ctx)
    return msg.decode() if msg else ""


def check(status, ctx=None):
    """0 -> ok; otherwise RuntimeError (what dolfin raises on solver failure,
    source/ns_problem.py:470 catches it)."""
    if status != 0:
        raise RuntimeError("nsb200: " + last_error(ctx))
