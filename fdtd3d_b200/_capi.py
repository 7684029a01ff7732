"""ctypes declarations of the C ABI in include/fdtd3d_b200.h (struct layouts, enums, prototypes).

Pure declarations: loading the shared library happens in fdtd3d_b200/__init__.py.
"""
import ctypes as C

ABI_VERSION = 1

# fdtd_status
OK, ERROR, INVALID_ARGUMENT, NO_DEVICE, COMM_ERROR = 0, 1, 2, 3, 4

# fdtd_field
EX, EY, EZ, HX, HY, HZ, DX, DY, DZ, BX, BY, BZ, EPS, MU, EINC, HINC = range(16)
FIELD_NAMES = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz", "Dx", "Dy", "Dz", "Bx", "By", "Bz", "Eps", "Mu", "EInc", "HInc"]
FIELD_IDS = {n: i for i, n in enumerate(FIELD_NAMES)}

# fdtd_order
ORDER_REFERENCE, ORDER_NATIVE = 0, 1


class Config(C.Structure):
    """struct fdtd_config"""
    _fields_ = [
        ("abi_version", C.c_int32),
        ("dims", C.c_int32),
        ("value_type", C.c_int32),
        ("complex_values", C.c_int32),
        ("use_pml", C.c_int32),
        ("use_tfsf", C.c_int32),
        ("exact_mode", C.c_int32),
        ("device", C.c_int32),
        ("size", C.c_int64 * 3),
        ("pml_size", C.c_int64 * 3),
        ("tfsf_size", C.c_int64 * 3),
        ("source_pos", C.c_int64 * 3),
        ("dx", C.c_double),
        ("dt", C.c_double),
        ("source_frequency", C.c_double),
        ("rel_phase_velocity", C.c_double),
        ("angle_teta", C.c_double),
        ("angle_phi", C.c_double),
        ("angle_psi", C.c_double),
        ("rank", C.c_int32),
        ("nranks", C.c_int32),
        ("time_step_start", C.c_int64),
    ]


class HostArrays(C.Structure):
    """struct fdtd_host_arrays"""
    _fields_ = [
        ("field", C.c_void_p * 12),
        ("eps", C.c_void_p),
        ("mu", C.c_void_p),
        ("sigma", C.c_void_p * 3),
        ("einc", C.c_void_p),
        ("hinc", C.c_void_p),
    ]


# name -> (restype, argtypes); every symbol include/fdtd3d_b200.h declares
PROTOTYPES = {
    "fdtd_last_error": (C.c_char_p, []),
    "fdtd_device_info": (C.c_int, [C.c_int, C.c_char_p, C.c_int64]),
    "fdtd_init_scheme": (C.c_int, [C.POINTER(Config), C.c_double, C.c_double, C.c_double, C.c_double]),
    "fdtd_pml_sigma_profile": (C.c_int, [C.POINTER(Config), C.c_int, C.c_void_p, C.c_int64]),
    "fdtd_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "fdtd_destroy": (C.c_int, [C.c_void_p]),
    "fdtd_set_sigma": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]),
    "fdtd_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_int64]),
    "fdtd_download": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_int64]),
    "fdtd_fill": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double]),
    "fdtd_init_eps_sphere": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double]),
    "fdtd_run": (C.c_int, [C.c_void_p, C.c_int64]),
    "fdtd_subtract_incident": (C.c_int, [C.c_void_p]),
    "fdtd_run_async": (C.c_int, [C.c_void_p, C.c_int64]),
    "fdtd_sync": (C.c_int, [C.c_void_p]),
    "fdtd_tfsf_face_masks": (C.c_int, [C.POINTER(Config), C.c_int, C.c_void_p, C.c_int64]),
    "fdtd_get_fused_box": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "fdtd_fused_time": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "fdtd_set_tuning": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "fdtd_step_breakdown": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "fdtd_busy": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "fdtd_last_run_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "fdtd_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "fdtd_phase_times": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "fdtd_launch_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "fdtd_get_time_step": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "fdtd_get_core_box": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "fdtd_plan_core_box": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "fdtd_field_sumsq": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double)]),
    "fdtd_sumsq_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int64]),
    "fdtd_comm_unique_id": (C.c_int, [C.c_void_p]),
    "fdtd_comm_init": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fdtd_execute_steps": (C.c_int, [C.POINTER(C.c_int), C.POINTER(Config), C.POINTER(HostArrays), C.c_int64]),
}
