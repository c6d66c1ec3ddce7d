"""ctypes binding of libsu2b200.so (include/su2b200.h).

This is the host-side glue used by the tests, bench.py and the Python mirror of the solver
interface.  It never computes anything itself and has no CPU fallback: if the CUDA library
cannot be loaded, or no GPU is present, every compute entry point raises.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SU2B200_LIB selects another build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("SU2B200_LIB") or os.path.join(_HERE, "libsu2b200.so")

c_ulong_p = C.POINTER(C.c_ulong)
c_double_p = C.POINTER(C.c_double)


class GeometryDesc(C.Structure):
    _fields_ = [
        ("nDim", C.c_int),
        ("nPoint", C.c_ulong), ("nPointDomain", C.c_ulong), ("nEdge", C.c_ulong),
        ("edge_node0", c_ulong_p), ("edge_node1", c_ulong_p),
        ("edge_normal", c_double_p), ("coord", c_double_p), ("volume", c_double_p),
        ("point_ptr", c_ulong_p), ("point_nbr", c_ulong_p), ("point_edge", c_ulong_p),
        ("row_ptr", c_ulong_p), ("col_ind", c_ulong_p), ("dia_ptr", c_ulong_p), ("edge_ptr", c_ulong_p),
        ("nNeighbor", C.POINTER(C.c_ushort)), ("physical_boundary", C.POINTER(C.c_ubyte)),
        ("wall_distance", c_double_p),
        ("color_ptr", c_ulong_p), ("color_edges", c_ulong_p), ("nColor", C.c_ulong),
    ]


class FlowConfig(C.Structure):
    _fields_ = [
        ("gamma", C.c_double), ("gas_constant", C.c_double),
        ("viscous", C.c_int), ("rans_sa", C.c_int), ("sa_negative", C.c_int),
        ("mu_ref", C.c_double), ("mu_temperature_ref", C.c_double), ("mu_sutherland", C.c_double),
        ("mu_constant", C.c_double),
        ("prandtl_lam", C.c_double), ("prandtl_turb", C.c_double),
        ("density_inf", C.c_double), ("pressure_inf", C.c_double), ("velocity_inf", C.c_double * 3),
        ("energy_inf", C.c_double), ("temperature_inf", C.c_double),
        ("nu_tilde_inf", C.c_double), ("eddy_visc_inf", C.c_double),
        ("conv_scheme", C.c_int), ("muscl", C.c_int), ("muscl_kappa", C.c_double),
        ("limiter", C.c_int), ("venkat_coeff", C.c_double), ("ref_elem_length", C.c_double),
        ("grad_method", C.c_int), ("grad_method_recon", C.c_int),
        ("roe_kappa", C.c_double), ("entropy_fix", C.c_double),
        ("jst_kappa2", C.c_double), ("jst_kappa4", C.c_double), ("cent_jac_fix_factor", C.c_double),
        ("cfl", C.c_double), ("cfl_turb_ratio", C.c_double), ("max_delta_time", C.c_double),
        ("max_update_flow", C.c_double), ("max_update_sa", C.c_double),
        ("linear_solver", C.c_int), ("linear_prec", C.c_int), ("linear_tol", C.c_double),
        ("linear_iter", C.c_int), ("linear_restart", C.c_int),
        ("turb_linear_solver", C.c_int), ("turb_linear_prec", C.c_int), ("turb_linear_iter", C.c_int),
        ("turb_linear_tol", C.c_double),
        ("aoa_deg", C.c_double), ("aos_deg", C.c_double), ("ref_area", C.c_double), ("ref_length", C.c_double),
        ("ref_origin", C.c_double * 3),
        ("assembly", C.c_int), ("prec_partitions", C.c_int), ("turb_reducer", C.c_int),
        ("cfl_adapt", C.c_int), ("cfl_adapt_param", C.c_double * 6), ("n_inner_iter", C.c_ulong),
        ("mg_level", C.c_int), ("lax_kappa1", C.c_double),
        ("rans_sst", C.c_int), ("kine_inf", C.c_double), ("omega_inf", C.c_double),
        ("sst_lower", C.c_double * 2), ("sst_upper", C.c_double * 2), ("max_update_sst", C.c_double),
        ("mixed_precision", C.c_int),
    ]


class MgConfig(C.Structure):
    _fields_ = [
        ("nLevels", C.c_int), ("cycle", C.c_int), ("pre_smooth", C.c_int * 8), ("post_smooth", C.c_int * 8),
        ("correction_smooth", C.c_int * 8), ("damp_restriction", C.c_double), ("damp_prolongation", C.c_double),
    ]


class IterReport(C.Structure):
    _fields_ = [
        ("rms_flow", C.c_double * 5), ("rms_turb", C.c_double), ("CL", C.c_double), ("CD", C.c_double),
        ("linsol_iter_flow", C.c_ulong), ("linsol_iter_turb", C.c_ulong),
        ("linsol_res_flow", C.c_double), ("linsol_res_turb", C.c_double),
        ("nonphysical_points", C.c_ulong), ("nonphysical_edges", C.c_ulong), ("rms_turb2", C.c_double),
    ]


# enums (include/su2b200.h)
EULER_WALL, FAR_FIELD, SYMMETRY_PLANE, HEAT_FLUX, SEND_RECEIVE = 1, 2, 3, 26, 99
CONV_ROE, CONV_JST = 0, 1
GRAD_GREEN_GAUSS, GRAD_LEAST_SQUARES, GRAD_WEIGHTED_LEAST_SQUARES = 0, 1, 2
LIMITER_NONE, LIMITER_VENKATAKRISHNAN, LIMITER_BARTH_JESPERSEN = 0, 1, 2
SOLVER_FGMRES, SOLVER_BCGSTAB, SOLVER_RESTARTED_FGMRES = 0, 1, 2
PREC_JACOBI, PREC_LU_SGS, PREC_ILU = 0, 1, 2
SYS_FLOW, SYS_TURB = 0, 1
ASSEMBLY_GATHER, ASSEMBLY_COLORED = 0, 1
ARR = dict(
    SOLUTION=0, SOLUTION_OLD=1, PRIMITIVE=2, GRAD_RECON=3, GRAD_PRIM=4, LIMITER=5, DELTA_TIME=6, LINSYSRES=7,
    LINSYSSOL=8, JACOBIAN=9, LAMBDA=10, SENSOR=11, UND_LAPL=12, MAX_LAMBDA_INV=13, MAX_LAMBDA_VISC=14, ILU=15,
    INVM=16, UNDER_RELAX=17, RES_TRUNC=18, LOCAL_CFL=19, TURB_SOLUTION=20, TURB_GRADIENT=21, TURB_LINSYSRES=22, TURB_LINSYSSOL=23,
    TURB_JACOBIAN=24, EDDY_VISC=25, VORTICITY=26, STRAIN_MAG=27, WALL_DIST=28, TURB_DELTA_TIME=29,
    TURB_SOLUTION_OLD=30, SST_F1=31, SST_F2=32, SST_CDKW=33)

# every symbol include/su2b200.h declares (tests check that the library exports all of them)
EXPORTED_SYMBOLS = [
    "su2b200_create", "su2b200_destroy", "su2b200_last_error", "su2b200_version", "su2b200_synchronize",
    "su2b200_geometry_upload", "su2b200_marker_upload", "su2b200_marker_charac_download", "su2b200_p2p_upload", "su2b200_flow_configure",
    "su2b200_flow_allocate", "su2b200_array_size", "su2b200_array_download", "su2b200_array_upload",
    "su2b200_edge_counter_download", "su2b200_flow_preprocessing", "su2b200_flow_set_old_solution",
    "su2b200_flow_set_time_step", "su2b200_flow_edge_residual", "su2b200_flow_bc_far_field",
    "su2b200_flow_bc_sym_plane", "su2b200_flow_bc_heatflux_wall", "su2b200_flow_space_integration",
    "su2b200_flow_prepare_implicit", "su2b200_flow_linear_solve", "su2b200_flow_complete_implicit",
    "su2b200_flow_forces", "su2b200_flow_adapt_cfl", "su2b200_turb_preprocessing", "su2b200_turb_set_time_step",
    "su2b200_turb_space_integration", "su2b200_turb_prepare_implicit", "su2b200_turb_linear_solve",
    "su2b200_turb_complete_implicit", "su2b200_turb_postprocessing", "su2b200_iterate", "su2b200_linsys_create",
    "su2b200_linsys_set_partitions", "su2b200_linsys_set_schedule", "su2b200_linsys_set_wide_threshold",
    "su2b200_linsys_set_partition_rows", "su2b200_linsys_set_matrix", "su2b200_linsys_spmv",
    "su2b200_linsys_build_preconditioner", "su2b200_linsys_apply_preconditioner", "su2b200_linsys_precondition_and_multiply", "su2b200_linsys_solve",
    "su2b200_vector_dot", "su2b200_vector_dot_noalloc", "su2b200_mg_link", "su2b200_mg_set_restricted_solution", "su2b200_mg_set_residual_term",
    "su2b200_mg_set_forcing_term", "su2b200_mg_get_prolongated_correction", "su2b200_mg_smooth_prolongated_correction",
    "su2b200_mg_set_prolongated_correction", "su2b200_mg_iterate", "su2b200_restart_write", "su2b200_restart_read",
    "su2b200_restart_write_arrays", "su2b200_restart_read_arrays", "su2b200_comm_unique_id", "su2b200_comm_init", "su2b200_comm_mailbox_create", "su2b200_comm_mailbox_connect", "su2b200_comm_mailbox_enable", "su2b200_halo_exchange",
    "su2b200_timer_start", "su2b200_timer_stop", "su2b200_launch_count", "su2b200_bench_edge_kernel",
    "su2b200_bench_spmv", "su2b200_bench_prec", "su2b200_struct_sizes",
    "su2b200_mesh_create_box", "su2b200_mesh_read_su2", "su2b200_mesh_write_su2", "su2b200_mesh_stretch", "su2b200_mesh_partition", "su2b200_mesh_destroy", "su2b200_mesh_set_marker",
    "su2b200_mesh_preprocess", "su2b200_mesh_sizes", "su2b200_mesh_array", "su2b200_mesh_marker_info",
    "su2b200_mesh_upload", "su2b200_profile_enable", "su2b200_profile_ncat", "su2b200_profile_name",
    "su2b200_profile_report",
]

NOALLOC_DOT_MAX = 1 << 20  # SU2B200_NOALLOC_DOT_MAX (include/su2b200.h)

_lib = None


class Su2B200Error(RuntimeError):
    pass


def load():
    """Load libsu2b200.so (built in-tree by __graft_entry__.build()). Fails loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Su2B200Error(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    lib.su2b200_last_error.restype = C.c_char_p
    lib.su2b200_array_size.restype = C.c_long
    lib.su2b200_launch_count.restype = C.c_ulong
    lib.su2b200_profile_name.restype = C.c_char_p
    g = C.c_int()
    f = C.c_int()
    r = C.c_int()
    lib.su2b200_struct_sizes(C.byref(g), C.byref(f), C.byref(r))
    if (g.value, f.value, r.value) != (C.sizeof(GeometryDesc), C.sizeof(FlowConfig), C.sizeof(IterReport)):
        raise Su2B200Error("ctypes structures out of sync with include/su2b200.h: "
                           f"C {(g.value, f.value, r.value)} vs Python "
                           f"{(C.sizeof(GeometryDesc), C.sizeof(FlowConfig), C.sizeof(IterReport))}")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise Su2B200Error(load().su2b200_last_error().decode())


def as_ulong(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def ptr_ul(a):
    return a.ctypes.data_as(c_ulong_p) if a is not None else None


def ptr_d(a):
    return a.ctypes.data_as(c_double_p) if a is not None else None
