"""dau_bias_grad_static: the allocation-free bias gradient (include/dau_conv_b200.h; no counterpart in the
reference).  Its per-slice partial sums live in a file-scope
`static __device__` array of the library instead of the caller's workspace.  Held here, through the C ABI,
against the existing entry points (dau_backward / dau_backward_act with DAU_GRAD_BIAS) on two calls in a row
whose sizes differ -- the second call meets the first call's leftovers in the shared array -- and against the
plain sum.  Tolerance: the tests' own (tests/_tolerance.py, 1e-4 rel / 1e-5 abs scaled by max|ref| for sums over
N*H*W terms); the two entry points run the same kernels in the same order, so they also agree bit for bit.

Named test_zz_* so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

from _tolerance import close

pytestmark = pytest.mark.gpu


def _existing_entry_point(dev, err, y_act, F, ks=9):
    """dbias through dau_backward (y_act None) or dau_backward_act (ReLU mask from the forward output)."""
    import torch
    import dau_conv
    from dau_conv import _lib
    N, _, H, W = err.shape
    S, G = 2, 2
    st = _lib.default_settings(num_output=F, number_units_x=1, number_units_y=G, kernel_size=ks, pad=ks // 2, bias_term=1)
    h = dau_conv.DAULayerHandle(st, dev)
    h.set_sigma_hint(0.5)
    x = torch.zeros((N, S, H, W), device=dev)
    p = [torch.zeros((1, S, G, F), device=dev) for _ in range(3)] + [torch.full((1, S, G, F), 0.5, device=dev)]
    dbias = torch.full((F,), 0.25, device=dev)                    # accumulated into
    h.backward(x, err, *p, flags=_lib.GRAD_BIAS, dbias=dbias, top_data=y_act,
               activation=_lib.ACT_RELU if y_act is not None else _lib.ACT_NONE)
    h.set_activation(_lib.ACT_NONE)
    torch.cuda.synchronize(dev)
    return dbias.cpu().numpy()


def _static_entry_point(dev, err, y_act, F):
    import torch
    from dau_conv import _lib
    N, _, H, W = err.shape
    import gc
    dbias = torch.full((F,), 0.25, device=dev)
    gc.collect()
    # torch's allocator counts every allocation it serves (monotonic; frees by a late garbage collection of other
    # tests' tensors do not touch it): the call must not add one
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    _lib.check(_lib.lib.dau_bias_grad_static(dev.index, err.data_ptr(), y_act.data_ptr() if y_act is not None else None,
                                             dbias.data_ptr(), N, F, H, W, torch.cuda.current_stream(dev).cuda_stream))
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
    torch.cuda.synchronize(dev)
    return dbias.cpu().numpy()


def test_static_bias_gradient_matches_the_existing_entry_point_twice_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev); gen.manual_seed(7)
    # first call: 20 images (all 16 slices of the scratch in use), F = 48; second: 3 images, F = 7, other extents
    for n, F, H, W, masked in ((20, 48, 13, 17, False), (3, 7, 9, 30, True), (20, 48, 13, 17, True)):
        err = torch.randn((n, F, H, W), device=dev, generator=gen)
        y_act = torch.relu(torch.randn((n, F, H, W), device=dev, generator=gen)) if masked else None
        got = _static_entry_point(dev, err, y_act, F)
        ref = _existing_entry_point(dev, err, y_act, F)
        close(got, ref, "dbias static vs dau_backward", scale_atol=True)
        assert np.array_equal(got, ref)                            # same kernels, same order of additions
        e = err.double() if y_act is None else torch.where(y_act > 0, err, torch.zeros_like(err)).double()
        close(got, 0.25 + e.sum((0, 2, 3)).cpu().numpy(), "dbias static vs sum", scale_atol=True)


def test_static_bias_gradient_refuses_sizes_beyond_its_scratch():
    import torch
    from dau_conv import _lib
    dev = torch.device("cuda", 0)
    t = torch.zeros(16, device=dev)
    rc = _lib.lib.dau_bias_grad_static(0, t.data_ptr(), None, t.data_ptr(), 1, 4097, 1, 1, None)
    assert rc == 4 and "4096" in _lib.lib.dau_last_error().decode()
    rc = _lib.lib.dau_bias_grad_static(0, None, None, t.data_ptr(), 1, 4, 1, 1, None)
    assert rc == 1
    torch.cuda.synchronize(dev)
    assert torch.cuda.current_device() == 0
