"""ctypes binding of libdau_conv_b200.so (C ABI: include/dau_conv_b200.h).

There is no CPU or PyTorch fallback: if the library is missing this module raises at import
time, and every entry point raises RuntimeError with dau_last_error() on a non-zero status.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdau_conv_b200.so")

# exported symbols, in the order of include/dau_conv_b200.h (checked by tests/test_capi.py)
SYMBOLS = [
    "dau_last_error", "dau_abi_version", "dau_settings_default", "dau_create", "dau_destroy",
    "dau_reshape", "dau_get_output_size", "dau_set_sigma_hint", "dau_set_offset_hint", "dau_forward", "dau_backward", "dau_set_activation", "dau_backward_act", "dau_take_nan_flag",
    "dau_get_kernels",
    "dau_forward_host", "dau_backward_host", "dau_train_step_host", "dau_plan_mosaic", "dau_plan_strips", "dau_mosaic_to_real", "dau_bias_grad_static", "dau_edge_mask_rule", "dau_debug_unit_tables", "dau_measure_fma_peak", "dau_enable_timing",
    "dau_get_timing", "dau_timing_name", "dau_launch_count",
]

GRAD_INPUT, GRAD_W, GRAD_MU1, GRAD_MU2, GRAD_SIGMA, GRAD_BIAS, GRAD_ALL = 1, 2, 4, 8, 16, 32, 31
ACT_NONE, ACT_RELU = 0, 1


class Settings(C.Structure):
    """dau_settings_t (mirrors DAUConvSettings, base_dau_conv_layer.hpp:109-130)."""
    _fields_ = [
        ("num_output", C.c_int32), ("number_units_x", C.c_int32), ("number_units_y", C.c_int32),
        ("number_units_ignore", C.c_int32), ("kernel_size", C.c_int32), ("pad", C.c_int32),
        ("stride", C.c_int32), ("bias_term", C.c_int32), ("unit_normalization", C.c_int32),
        ("square_unit_normalization", C.c_int32), ("use_interpolation", C.c_int32),
        ("single_dimension_kernel", C.c_int32), ("forbid_positive_dim1", C.c_int32),
        ("ignore_edge_gradients", C.c_int32), ("unit_bounds_guard", C.c_int32),
        ("offsets_already_centered", C.c_int32), ("learn_sigma", C.c_int32),
        ("component_border_bound", C.c_float), ("sigma_lower_bound", C.c_float),
        ("mu_learning_rate_factor", C.c_float),
        ("mean_iteration_step", C.c_int32), ("sigma_iteration_step", C.c_int32),
        ("merge_iteration_step", C.c_int32), ("merge_threshold", C.c_float),
    ]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "libdau_conv_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
        "or `make -C dau-convnet_b200/csrc`. There is no CPU fallback for the DAU layer." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_fp = C.POINTER(C.c_float)
_vp = C.c_void_p
lib.dau_last_error.restype = C.c_char_p
lib.dau_abi_version.restype = C.c_int
lib.dau_settings_default.argtypes = [C.POINTER(Settings)]
lib.dau_settings_default.restype = None
lib.dau_create.argtypes = [C.POINTER(Settings), C.c_int, C.POINTER(_vp)]
lib.dau_destroy.argtypes = [_vp]
lib.dau_destroy.restype = None
lib.dau_reshape.argtypes = [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
lib.dau_get_output_size.argtypes = [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
lib.dau_set_sigma_hint.argtypes = [_vp, C.c_float]
lib.dau_set_offset_hint.argtypes = [_vp, C.c_float]
lib.dau_forward.argtypes = [_vp] + [_vp] * 7 + [_vp, C.c_size_t, _vp]
lib.dau_backward.argtypes = [_vp] + [_vp] * 12 + [C.c_uint, _vp, C.c_size_t, _vp]
lib.dau_take_nan_flag.argtypes = [_vp, C.POINTER(C.c_int), _vp]
lib.dau_set_activation.argtypes = [_vp, C.c_int]
lib.dau_backward_act.argtypes = [_vp] + [_vp] * 13 + [C.c_uint, _vp, C.c_size_t, _vp]
lib.dau_get_kernels.argtypes = [C.POINTER(Settings), C.c_float, C.POINTER(C.c_int), _fp, C.c_size_t]
lib.dau_forward_host.argtypes = [_vp, C.c_int, C.c_int, C.c_int, C.c_int] + [_vp] * 7
lib.dau_backward_host.argtypes = [_vp, C.c_int, C.c_int, C.c_int, C.c_int] + [_vp] * 13 + [C.c_uint]
lib.dau_train_step_host.argtypes = [_vp] + [C.c_int] * 5 + [C.c_float] + [_vp] * 12 + [_vp]
lib.dau_plan_mosaic.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_int)]
lib.dau_plan_strips.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.c_longlong]
lib.dau_mosaic_to_real.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 3
lib.dau_mosaic_to_real.restype = C.c_int
lib.dau_bias_grad_static.argtypes = [C.c_int, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp]
lib.dau_edge_mask_rule.argtypes = [C.c_int]
lib.dau_edge_mask_rule.restype = C.c_int
lib.dau_debug_unit_tables.argtypes = [_vp] * 7 + [C.POINTER(C.c_int), _vp]
lib.dau_measure_fma_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
lib.dau_enable_timing.argtypes = [_vp, C.c_int]
lib.dau_get_timing.argtypes = [_vp, C.c_int, _fp, C.c_int, C.POINTER(C.c_int)]
lib.dau_timing_name.argtypes = [C.c_int, C.c_int]
lib.dau_timing_name.restype = C.c_char_p
lib.dau_launch_count.argtypes = [_vp]
lib.dau_launch_count.restype = C.c_longlong


class DAUError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("dau_conv_b200 error %d: %s" % (status, message))
        self.status = status


def check(status):
    if status != 0:
        raise DAUError(status, lib.dau_last_error().decode("utf-8", "replace"))


def default_settings(**kw):
    s = Settings()
    lib.dau_settings_default(C.byref(s))
    for k, v in kw.items():
        if not hasattr(s, k):
            raise TypeError("unknown DAU setting %r" % k)
        setattr(s, k, v)
    return s


def get_kernels(settings, sigma):
    """Six Kb x Kb numpy arrays: blur, d_w, d_mu1, d_mu2, d_sigma, d_error."""
    import numpy as np
    kb = C.c_int(0)
    buf = np.zeros(6 * 17 * 17, dtype=np.float32)
    check(lib.dau_get_kernels(C.byref(settings), float(sigma), C.byref(kb),
                              buf.ctypes.data_as(_fp), buf.size))
    k = kb.value
    return [buf[i * k * k:(i + 1) * k * k].reshape(k, k).copy() for i in range(6)]
