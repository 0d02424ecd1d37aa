"""ctypes declarations for libsunsky_b200.so (include/sunsky_b200.h). Loading fails loudly when the
library has not been built: there is no Python or CPU fallback for the compute entry points."""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_PKG)                      # mitsuba3-sunsky_b200/
LIB_PATH = os.environ.get("SSK_LIB") or os.path.join(ROOT, "libsunsky_b200.so")   # SSK_LIB: tuning builds only
DATASET_DIR = os.path.join(ROOT, "data", "resources", "sunsky", "datasets")

SSK_OK, SSK_ERR_INVALID, SSK_ERR_IO, SSK_ERR_CUDA, SSK_ERR_UNSUPPORTED = 0, -1, -2, -3, -4
MODE_RGB, MODE_SPECTRAL = 0, 1
FLAG_SUN_DIR_NORMALIZED = 1
FLAG_TO_LOCAL_GIVEN = 2
ARITH_EXACT, ARITH_FAST = 0, 1


class Params(C.Structure):
    _fields_ = [
        ("turbidity", C.c_float), ("n_albedo", C.c_int32), ("albedo", C.c_float * 11),
        ("use_record", C.c_int32), ("sun_direction", C.c_float * 3),
        ("latitude", C.c_float), ("longitude", C.c_float), ("timezone", C.c_float),
        ("year", C.c_int32), ("month", C.c_int32), ("day", C.c_int32),
        ("hour", C.c_float), ("minute", C.c_float), ("second", C.c_float),
        ("sun_scale", C.c_float), ("sky_scale", C.c_float), ("sun_aperture_deg", C.c_float),
        ("to_world", C.c_float * 16), ("flags", C.c_uint32), ("arith", C.c_int32),
        ("to_local", C.c_float * 16),
    ]


class State(C.Structure):
    _fields_ = [
        ("n_channels", C.c_int32), ("sky_params", C.c_float * 99), ("sky_radiance", C.c_float * 11),
        ("n_sun_radiance", C.c_int32), ("sun_radiance", C.c_float * 3240), ("sun_ld", C.c_float * 66),
        ("gaussians", C.c_float * 100), ("pmf", C.c_float * 20), ("cdf", C.c_float * 20),
        ("cdf_sum", C.c_float), ("cdf_normalization", C.c_float), ("sky_sampling_w", C.c_float),
        ("sun_angles", C.c_float * 2), ("frame_s", C.c_float * 3), ("frame_t", C.c_float * 3),
        ("frame_n", C.c_float * 3), ("sun_half_aperture", C.c_float), ("sun_dir_world", C.c_float * 3),
        ("bsphere_center", C.c_float * 3), ("bsphere_radius", C.c_float),
        ("spectral_pdf", C.c_float * 10), ("spectral_cdf", C.c_float * 10),
        ("spectral_integral", C.c_float), ("spectral_normalization", C.c_float),
        ("area_ratio", C.c_float), ("sun_below_horizon", C.c_int32), ("to_local", C.c_float * 9),
    ]


FP = C.POINTER(C.c_float)
FPP = C.POINTER(FP)
CTX = C.c_void_p

# every symbol include/sunsky_b200.h declares, with its signature
SYMBOLS = {
    "ssk_last_error": (C.c_char_p, []),
    "ssk_version": (C.c_char_p, []),
    "ssk_params_default": (None, [C.POINTER(Params)]),
    "ssk_create": (C.c_int, [C.POINTER(Params), C.c_char_p, C.c_int, C.c_int, C.POINTER(CTX)]),
    "ssk_update": (C.c_int, [CTX, C.POINTER(Params)]),
    "ssk_update_batch": (C.c_int, [C.POINTER(CTX), C.POINTER(Params), C.c_int]),
    "ssk_update_batch_noalloc": (C.c_int, [C.POINTER(CTX), C.POINTER(Params), C.c_int]),
    "ssk_set_scene": (C.c_int, [CTX, FP, FP]),
    "ssk_destroy": (None, [CTX]),
    "ssk_get_state": (C.c_int, [CTX, C.POINTER(State)]),
    "ssk_channels": (C.c_int, [CTX]),
    "ssk_device": (C.c_int, [CTX]),
    "ssk_to_string": (C.c_int, [CTX, C.c_char_p, C.c_size_t]),
    "ssk_eval_device": (C.c_int, [CTX, FPP, FPP, C.c_size_t, FPP, C.c_void_p]),
    "ssk_eval_direction_device": (C.c_int, [CTX, FPP, FPP, C.c_size_t, FPP, C.c_void_p]),
    "ssk_pdf_direction_device": (C.c_int, [CTX, FPP, C.c_size_t, FP, C.c_void_p]),
    "ssk_sample_direction_device": (C.c_int, [CTX, FPP, FPP, FPP, C.c_size_t, FPP, FP, FPP, FPP, FP, C.c_void_p]),
    "ssk_eval_jvp_device": (C.c_int, [CTX, C.c_float, FP, C.c_int, FP, FPP, FPP, C.c_size_t, FPP, C.c_void_p]),
    "ssk_sample_wavelengths_device": (C.c_int, [CTX, FPP, FP, C.c_size_t, FPP, FPP, C.c_void_p]),
    "ssk_sample_ray_device": (C.c_int, [CTX, FP, FPP, FPP, C.c_size_t, FPP, FPP, FPP, FPP, C.c_void_p]),
    "ssk_sample_wavelengths_host": (C.c_int, [CTX, FPP, FP, C.c_size_t, FPP, FPP]),
    "ssk_sample_ray_host": (C.c_int, [CTX, FP, FPP, FPP, C.c_size_t, FPP, FPP, FPP, FPP]),
    "ssk_envmap_device": (C.c_int, [C.POINTER(CTX), C.c_int, C.c_int, C.c_int, FP, FP, C.c_void_p]),
    "ssk_eval_host": (C.c_int, [CTX, FPP, FPP, C.c_size_t, FPP]),
    "ssk_eval_direction_host": (C.c_int, [CTX, FPP, FPP, C.c_size_t, FPP]),
    "ssk_pdf_direction_host": (C.c_int, [CTX, FPP, C.c_size_t, FP]),
    "ssk_sample_direction_host": (C.c_int, [CTX, FPP, FPP, FPP, C.c_size_t, FPP, FP, FPP, FPP, FP]),
    "ssk_launch_count": (C.c_uint64, []),
    "ssk_host_copy_ceiling": (C.c_int, [C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.POINTER(C.c_double),
                                      C.POINTER(C.c_double)]),
    "ssk_debug_fdiv_by_device": (C.c_int, [FP, FP, FP, C.c_size_t, C.c_void_p]),
    "ssk_debug_math_device": (C.c_int, [C.c_int, FP, FP, FP, C.c_size_t, C.c_void_p]),
    "ssk_debug_set_exact_limits": (C.c_int, [CTX, C.c_float, C.c_float]),
    "ssk_sun_coordinates": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                      C.c_float, C.c_float, FP]),
    "ssk_normalize_direction": (C.c_int, [FP, FP]),
    "ssk_hosek_sun_rad": (C.c_int, [C.c_char_p, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]),
    "ssk_array_from_file": (C.c_int, [C.c_char_p, C.c_int, FP, C.c_size_t, C.POINTER(C.c_size_t),
                                      C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "ssk_array_to_file": (C.c_int, [C.c_char_p, FP, C.POINTER(C.c_size_t), C.c_size_t, C.c_int]),
    # ---- multi-GPU
    "ssk_multi_last_error": (C.c_char_p, []),
    "ssk_shard_ray_range": (C.c_int, [C.c_size_t, C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "ssk_shard_frames": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ssk_peer_alloc": (C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "ssk_peer_open": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    "ssk_peer_close": (C.c_int, [C.c_int, C.c_void_p]),
    "ssk_peer_free": (C.c_int, [C.c_int, C.c_void_p]),
    "ssk_sweep_render": (C.c_int, [C.POINTER(Params), C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, FP, FPP,
                                   C.c_int, C.c_void_p]),
    "ssk_sweep_release": (None, []),
    "ssk_comm_unique_id": (C.c_int, [C.c_char_p]),
    "ssk_comm_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "ssk_comm_destroy": (None, [C.c_void_p]),
    "ssk_comm_version": (C.c_int, []),
    "ssk_comm_gather_frames": (C.c_int, [C.c_void_p, FP, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, FP, C.c_int, C.c_void_p]),
    "ssk_sweep_render_gather_nccl": (C.c_int, [C.c_void_p, C.POINTER(Params), C.c_int, C.c_int, C.c_char_p, C.c_int, C.c_int,
                                               C.c_int, FP, FP, FP, C.c_int, C.c_int, C.c_void_p]),
    "ssk_group_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(Params), C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ssk_group_destroy": (None, [C.c_void_p]),
    "ssk_group_size": (C.c_int, [C.c_void_p]),
    "ssk_group_context": (C.c_void_p, [C.c_void_p, C.c_int]),
    "ssk_group_update": (C.c_int, [C.c_void_p, C.POINTER(Params)]),
    "ssk_group_eval_host": (C.c_int, [C.c_void_p, FPP, FPP, C.c_size_t, FPP]),
    "ssk_group_sample_direction_host": (C.c_int, [C.c_void_p, FPP, FPP, C.c_size_t, FPP, FP, FPP]),
    "ssk_group_sweep": (C.c_int, [C.c_void_p, C.POINTER(Params), C.c_int, C.c_int, C.c_int, C.c_int, FP, FPP, C.c_int]),
}
SHARD_BLOCKED, SHARD_ROUND_ROBIN = 0, 1
PEER_HANDLE_BYTES, COMM_ID_BYTES = 64, 128
# not in the public header: test hook for the pixel -> direction convention of ssk_envmap_device
EXTRA_SYMBOLS = {
    "ssk_envmap_directions_device": (C.c_int, [CTX, C.c_int, C.c_int, FPP, C.c_void_p]),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.join(ROOT, 'csrc')}` "
                "(or __graft_entry__.build()). The sunsky kernels have no Python/CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in {**SYMBOLS, **EXTRA_SYMBOLS}.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
