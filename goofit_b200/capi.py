"""ctypes view of the C ABI (include/goofit_b200.h) — used by bench.py / tests that talk to the library
directly (resident device columns, NCCL communicator attachment) instead of through the host classes."""
import ctypes as C
import os

_here = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_here, "lib", "libgoofit_b200.so")

# every symbol include/goofit_b200.h declares
SYMBOLS = [
    "gf_last_error", "gf_version", "gf_create", "gf_destroy", "gf_device_info", "gf_set_program", "gf_validate_program", "gf_reset_aux",
    "gf_set_params", "gf_set_norms", "gf_upload_columns", "gf_adopt_device_columns", "gf_eval_sum",
    "gf_eval_sum_async", "gf_partial_device_ptr", "gf_stream", "gf_sync", "gf_eval_events", "gf_integrate_grid",
    "gf_conv_setup", "gf_conv_fill", "gf_conv_sizes", "gf_conv_get_workspace", "gf_amp3_alloc", "gf_amp3_cache_waves",
    "gf_amp3_integrals", "gf_amp3_get_wave", "gf_amp3_wave_gram", "gf_amp4_alloc", "gf_amp4_set_mc", "gf_amp4_cache", "gf_amp4_norm",
    "gf_amp4_norm_async", "gf_amp4_gram", "gf_amp4_get_amp", "gf_last_kernel_ms", "gf_last_kernel_name", "gf_set_option", "gf_amp4_cache_gram", "gf_amp4_cache_data", "gf_genbod", "gf_genbod_static", "gf_table_set", "gf_launch_count", "gf_comm_unique_id", "gf_comm_init",
    "gf_comm_destroy", "gf_allgather_events", "gf_comm_attached", "gf_comm_path", "gf_profile_begin", "gf_profile_end", "gf_timer_start", "gf_timer_stop", "gf_fp64_peak", "gf_set_params_async",
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(PATH):
            raise OSError("goofit_b200: %s is missing; build it with `make cuda` (there is no CPU fallback)" % PATH)
        _lib = C.CDLL(PATH)
        _lib.gf_last_error.restype = C.c_char_p
        _lib.gf_create.restype = C.c_void_p
        _lib.gf_create.argtypes = [C.c_int]
        _lib.gf_destroy.argtypes = [C.c_void_p]
        _lib.gf_partial_device_ptr.restype = C.c_void_p
        _lib.gf_partial_device_ptr.argtypes = [C.c_void_p]
        _lib.gf_stream.restype = C.c_void_p
        _lib.gf_stream.argtypes = [C.c_void_p]
        _lib.gf_last_kernel_ms.restype = C.c_float
        _lib.gf_last_kernel_ms.argtypes = [C.c_void_p]
        _lib.gf_last_kernel_name.restype = C.c_char_p
        _lib.gf_last_kernel_name.argtypes = [C.c_void_p]
        _lib.gf_launch_count.restype = C.c_uint64
        _lib.gf_launch_count.argtypes = [C.c_void_p]
    return _lib


def comm_path(ctx):
    f = lib().gf_comm_path
    f.restype = C.c_char_p
    f.argtypes = [C.c_void_p]
    return f(ctx).decode()


def last_error():
    return lib().gf_last_error().decode()
