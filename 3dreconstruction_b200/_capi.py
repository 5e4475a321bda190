"""ctypes mirror of include/ba_engine.h and include/ba_synth.h (field order and types must match the headers)."""
import ctypes as C

ABI_VERSION = 1

ROBUST_NONE, ROBUST_HUBER, ROBUST_CAUCHY = 0, 1, 2
MODEL_PINHOLE, MODEL_SNAVELY = 0, 1
LINEAR_EXACT, LINEAR_PCG = 0, 1
PRECOND_JACOBI, PRECOND_SCHUR_JACOBI = 0, 1
TERM_CONVERGENCE, TERM_NO_CONVERGENCE, TERM_FAILURE, TERM_TARGET_REACHED = 0, 1, 2, 3

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_uint8_p = C.POINTER(C.c_uint8)


class BAOptions(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("model", C.c_int32),
        ("robust_kind", C.c_int32),
        ("robust_a", C.c_double),
        ("intrinsics_free", C.c_int32),
        ("taylor_rule", C.c_int32),
        ("max_num_iterations", C.c_int32),
        ("initial_trust_region_radius", C.c_double),
        ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double),
        ("min_relative_decrease", C.c_double),
        ("min_lm_diagonal", C.c_double),
        ("max_lm_diagonal", C.c_double),
        ("max_num_consecutive_invalid_steps", C.c_int32),
        ("function_tolerance", C.c_double),
        ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double),
        ("jacobi_scaling", C.c_int32),
        ("linear_solver", C.c_int32),
        ("preconditioner", C.c_int32),
        ("pcg_max_iterations", C.c_int32),
        ("pcg_rel_tolerance", C.c_double),
        ("pcg_forcing", C.c_int32),
        ("eta", C.c_double),
        ("target_chi2", C.c_double),
        ("device", C.c_int32),
        ("num_threads", C.c_int32),
        ("rank", C.c_int32),
        ("world_size", C.c_int32),
        ("nccl_unique_id", C.c_void_p),
        ("stream", C.c_void_p),
        ("scatter_mode", C.c_int32),
        ("p2p_allreduce", C.c_int32),
        ("lm_policy", C.c_int32),
        ("warm_start_radius", C.c_int32),
        ("reserved", C.c_int32 * 4),
    ]


class BAProblem(C.Structure):
    _fields_ = [
        ("num_cameras", C.c_int32),
        ("num_groups", C.c_int32),
        ("num_points", C.c_int32),
        ("num_observations", C.c_int32),
        ("pose", c_double_p),
        ("intr", c_double_p),
        ("cam_group", c_int32_p),
        ("cam_fixed", c_uint8_p),
        ("pt", c_double_p),
        ("obs_cam", c_int32_p),
        ("obs_pt", c_int32_p),
        ("obs_uv", c_double_p),
    ]


class BADelta(C.Structure):
    _fields_ = [
        ("num_new_cameras", C.c_int32), ("pose", c_double_p), ("cam_group", c_int32_p), ("cam_fixed", c_uint8_p),
        ("num_new_groups", C.c_int32), ("intr", c_double_p),
        ("num_new_points", C.c_int32), ("pt", c_double_p),
        ("num_new_observations", C.c_int32), ("obs_cam", c_int32_p), ("obs_pt", c_int32_p), ("obs_uv", c_double_p),
    ]


class BASummary(C.Structure):
    _fields_ = [
        ("termination", C.c_int32),
        ("usable", C.c_int32),
        ("num_iterations", C.c_int32),
        ("num_successful_steps", C.c_int32),
        ("num_unsuccessful_steps", C.c_int32),
        ("num_residuals", C.c_int32),
        ("total_pcg_iterations", C.c_int64),
        ("initial_cost", C.c_double),
        ("final_cost", C.c_double),
        ("solve_seconds", C.c_double),
        ("kernel_launches", C.c_int64),
        ("message", C.c_char * 160),
    ]


class BAIterRecord(C.Structure):
    _fields_ = [
        ("iteration", C.c_int32),
        ("step_is_valid", C.c_int32),
        ("step_is_successful", C.c_int32),
        ("pcg_iterations", C.c_int32),
        ("cost", C.c_double),
        ("candidate_cost", C.c_double),
        ("model_cost_change", C.c_double),
        ("relative_decrease", C.c_double),
        ("radius_used", C.c_double),
        ("radius", C.c_double),
        ("step_norm", C.c_double),
        ("gradient_max_norm", C.c_double),
        ("pcg_rel_residual", C.c_double),
    ]


class BAKernelStat(C.Structure):
    _fields_ = [
        ("name", C.c_char * 32),
        ("launches", C.c_int64),
        ("total_ms", C.c_double),
        ("bytes_per_launch", C.c_double),
    ]


class SynthSpec(C.Structure):
    _fields_ = [
        ("num_cameras", C.c_int32),
        ("num_points", C.c_int32),
        ("num_observations", C.c_int32),
        ("seed", C.c_uint64),
        ("max_track", C.c_int32),
        ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
        ("noise_px", C.c_double),
        ("outlier_fraction", C.c_double),
        ("outlier_px", C.c_double),
        ("sigma_aa", C.c_double),
        ("sigma_t_rel", C.c_double),
        ("sigma_pt_rel", C.c_double),
        ("round_obs_to_float", C.c_int32),
        ("first_pose_zero", C.c_int32),
        ("fixed_camera", C.c_int32),
        ("arc", C.c_int32),
    ]


class SynthProblem(C.Structure):
    _fields_ = [
        ("num_cameras", C.c_int32), ("num_groups", C.c_int32), ("num_points", C.c_int32), ("num_observations", C.c_int32),
        ("pose", c_double_p),
        ("intr", c_double_p),
        ("cam_group", c_int32_p),
        ("cam_fixed", c_uint8_p),
        ("pt", c_double_p),
        ("obs_cam", c_int32_p),
        ("obs_pt", c_int32_p),
        ("obs_uv", c_double_p),
        ("pose_true", c_double_p),
        ("pt_true", c_double_p),
        ("extent", C.c_double),
    ]


# Every symbol include/ba_engine.h declares: (name, restype, argtypes)
ENGINE_SYMBOLS = [
    ("ba_options_default", C.c_int, [C.POINTER(BAOptions)]),
    ("ba_create", C.c_int, [C.POINTER(BAOptions), C.POINTER(C.c_void_p)]),
    ("ba_set_problem", C.c_int, [C.c_void_p, C.POINTER(BAProblem)]),
    ("ba_reset_state", C.c_int, [C.c_void_p]),
    ("ba_append", C.c_int, [C.c_void_p, C.POINTER(BADelta)]),
    ("ba_update_parameters", C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p]),
    ("ba_get_observations", C.c_int, [C.c_void_p, c_int32_p, c_int32_p, c_double_p, C.c_int32, c_int32_p]),
    ("ba_solve", C.c_int, [C.c_void_p, C.POINTER(BASummary)]),
    ("ba_get_state", C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    ("ba_get_trajectory", C.c_int, [C.c_void_p, C.POINTER(BAIterRecord), C.c_int32, c_int32_p]),
    ("ba_get_structure", C.c_int, [C.c_void_p, c_int32_p, c_int32_p, c_int32_p, c_int32_p, C.c_int64, C.POINTER(C.c_int64)]),
    ("ba_evaluate", C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    ("ba_set_profiling", C.c_int, [C.c_void_p, C.c_int32]),
    ("ba_set_profiling_stride", C.c_int, [C.c_void_p, C.c_int32]),
    ("ba_linear_solver_in_use", C.c_int, [C.c_void_p, c_int32_p]),
    ("ba_get_kernel_stats", C.c_int, [C.c_void_p, C.POINTER(BAKernelStat), C.c_int32, c_int32_p]),
    ("ba_get_execution_info", C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.c_int32, c_int32_p]),
    ("ba_cost_noalloc", C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 9),
    ("ba_get_io_bytes", C.c_int, [C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("ba_shard_points", C.c_int, [C.c_int32, C.c_int32, c_int32_p, C.c_int32, c_int32_p]),
    ("ba_nccl_unique_id", C.c_int, [C.c_void_p]),
    ("ba_last_error", C.c_char_p, [C.c_void_p]),
    ("ba_backend", C.c_char_p, []),
    ("ba_destroy", None, [C.c_void_p]),
]

SYNTH_SYMBOLS = [
    ("ba_synth_spec_default", C.c_int, [C.POINTER(SynthSpec)]),
    ("ba_synth_named", C.c_int, [C.c_char_p, C.POINTER(SynthSpec)]),
    ("ba_synth_generate", C.c_int, [C.POINTER(SynthSpec), C.POINTER(C.POINTER(SynthProblem))]),
    ("ba_synth_free", None, [C.POINTER(SynthProblem)]),
    ("ba_bal_read", C.c_int, [C.c_char_p, C.POINTER(C.POINTER(SynthProblem))]),
    ("ba_bal_write", C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p, c_int32_p, c_double_p, c_int32_p, c_int32_p, c_double_p]),
    ("ba_write_ply", C.c_int, [C.c_char_p, C.c_int32, c_double_p, C.c_int32, c_double_p]),
]


def bind(lib, symbols):
    for name, restype, argtypes in symbols:
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = restype
        fn.argtypes = argtypes
    return lib
