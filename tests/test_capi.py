"""CPU-only checks of the C-ABI library: it loads, exports every symbol include/dau_conv_b200.h
declares, its host-side DAUKernelCompute matches the oracle, and argument errors are reported
the reference's way (status + message instead of DAUException)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import dau_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dau_conv_b200.h")


@pytest.fixture(scope="module")
def dau_conv():
    import dau_conv as m
    return m


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dau_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(dau_conv):
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(dau_conv._lib.lib, s), "missing export: " + s
    assert sorted(dau_conv._lib.SYMBOLS) == syms
    assert dau_conv._lib.lib.dau_abi_version() == 1


def test_settings_struct_layout_matches_header(dau_conv):
    src = open(HEADER).read()
    body = src[src.index("typedef struct dau_settings_t {"):src.index("} dau_settings_t;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in re.findall(r"(?:int32_t|float)\s+([^;]+);", body):
        names += [n.strip() for n in decl.split(",")]
    assert names == [f[0] for f in dau_conv._lib.Settings._fields_]
    assert C.sizeof(dau_conv._lib.Settings) == 4 * len(names)


def test_default_settings_are_the_tf_op_defaults(dau_conv):
    s = dau_conv.default_settings()
    assert (s.num_output, s.number_units_x, s.number_units_y, s.kernel_size, s.pad, s.stride) == (64, 2, 2, 9, 4, 1)
    assert s.unit_normalization == 1 and s.use_interpolation == 1 and s.mu_learning_rate_factor == 1.0


@pytest.mark.parametrize("sigma", [0.5, 0.8, 1.2, 1.6])
@pytest.mark.parametrize("opts", [
    {}, {"single_dimension_kernel": 1}, {"forbid_positive_dim1": 1},
    {"square_unit_normalization": 1}, {"unit_normalization": 0},
])
def test_get_kernels_matches_oracle(dau_conv, sigma, opts):
    s = dau_conv.default_settings(**opts)
    got = dau_conv.get_kernels(s, sigma)
    want = O.get_filters(sigma, None, bool(s.unit_normalization), bool(s.square_unit_normalization),
                         bool(s.single_dimension_kernel), bool(s.forbid_positive_dim1))
    assert got[0].shape == (O.blur_kernel_size(sigma),) * 2
    for g, w in zip(got, want):
        np.testing.assert_allclose(g, w, rtol=1e-6, atol=1e-7)


def test_get_kernels_rejects_large_sigma(dau_conv):
    with pytest.raises(dau_conv.DAUError) as e:
        dau_conv.get_kernels(dau_conv.default_settings(), 2.0)     # Kb = 21 > 17 (convolve.cu:40)
    assert e.value.status == 4


def test_null_arguments_are_reported(dau_conv):
    lib = dau_conv._lib.lib
    assert lib.dau_create(None, 0, None) == 1
    assert b"NULL" in lib.dau_last_error()
    assert lib.dau_reshape(None, 1, 1, 1, 1, None, None) == 1
    lib.dau_destroy(None)
    assert lib.dau_launch_count(None) == 0
    assert lib.dau_timing_name(0, 3) == b"gather_forward"
    assert lib.dau_timing_name(1, 2) == b"grad_units"
    assert lib.dau_timing_name(2, 0) is None


def test_create_without_gpu_fails_loudly(dau_conv):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    st = dau_conv._lib.lib.dau_create(C.byref(dau_conv.default_settings()), 0, C.byref(h))
    assert st == 2 and not h.value          # CUDA error, no silent CPU path
    with pytest.raises(RuntimeError):
        import torch as t
        dau_conv.dau_conv_op(*[t.zeros(1, 1, 1, 1)] * 5)


def test_invalid_settings_rejected_before_touching_the_device(dau_conv):
    h = C.c_void_p()
    lib = dau_conv._lib.lib
    for bad in ({"stride": 2}, {"kernel_size": 8}, {"pad": 5}, {"pad": -1}, {"num_output": 0}, {"kernel_size": 67, "pad": 33}):
        st = lib.dau_create(C.byref(dau_conv.default_settings(**bad)), 0, C.byref(h))
        assert st in (1, 4), bad
