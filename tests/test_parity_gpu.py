"""Parity of the CUDA path (through the C ABI) against the oracle and the golden fixtures.

Tolerance (BASELINE.json north_star): 1e-4 relative / 1e-5 absolute in fp32 for the output
and all five gradients.  For gradients, which are fp32 sums of up to N*H*W terms, the absolute
term is scaled by max(1, max|ref|) of the tensor (documented in DESIGN.md).
"""
import ast
import glob
import os

import numpy as np
import pytest

from oracle import dau_oracle as O

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(p for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))
                if not os.path.basename(p).startswith("filters_"))
from _tolerance import ATOL, RTOL, close  # noqa: E402,F401  (the rule and its bookkeeping)


def run_layer(x, err, w, mu1, mu2, sigma, ks, **attrs):
    import torch
    import dau_conv
    dev = torch.device("cuda:0")
    S, G, F = w.shape[1:]
    sig = np.full(w.shape, sigma, np.float32)
    t = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev) for a in (x, w, mu1, mu2, sig)]
    a = dict(num_output=F, number_units_x=1, number_units_y=G, kernel_size=ks, pad=(ks - 1) // 2)
    a.update(attrs)
    y = dau_conv.dau_conv_op(*t, **a)
    g = dau_conv.dau_conv_grad_op(torch.from_numpy(np.ascontiguousarray(err, dtype=np.float32)).to(dev), *t, **a)
    torch.cuda.synchronize()
    return y.cpu().numpy(), [v.cpu().numpy() for v in g]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_fixture(path):
    """Outputs of the reference's own DAUConvPython (tests/golden, oracle/make_golden.py)."""
    d = np.load(path)
    kw = ast.literal_eval(str(d["kwargs"]))
    m = max(float(np.abs(d["mu1"]).max()), float(np.abs(d["mu2"]).max()))
    ks = 9 if m < 4 else 17
    attrs = dict(use_interpolation=kw.get("use_interpolation", True),
                 number_units_ignore=kw.get("num_dau_units_ignore", 0),
                 unit_testing=kw.get("unit_testing", True),
                 single_dim_kernel=kw.get("single_dim_kernel", False),
                 forbid_positive_dim1=kw.get("aggr_forbid_positive", False))
    y, g = run_layer(d["x"], d["err"], d["w"], d["mu1"], d["mu2"], float(d["sigma"]), ks, **attrs)
    close(y, d["y"], "y")
    for got, name in zip(g, ("dx", "dw", "dmu1", "dmu2", "dsigma")):
        close(got, d[name], name, scale_atol=name != "dx")


# (N, W, H, S, F, units, max_kernel, mu_range): the reference's Quick matrix
# (dau_conv_test.py:418-437) plus shapes that exercise partial tiles and odd channel counts
SHAPES = [
    (2, 65, 8, 33, 32, (1, 2), 9, 3), (1, 65, 8, 32, 32, (1, 2), 9, 3),
    (1, 8, 8, 32, 32, (1, 2), 9, 3), (4, 8, 8, 32, 32, (1, 2), 9, 3),
    (16, 32, 32, 32, 32, (2, 2), 9, 3), (16, 32, 32, 32, 32, (2, 2), 17, 6),
    (16, 32, 32, 32, 32, (2, 2), 17, 3), (16, 32, 32, 3, 32, (2, 2), 17, 3),
    (4, 64, 64, 3, 32, (2, 2), 33, 10),
    (3, 27, 27, 8, 16, (1, 2), 9, 3), (5, 13, 13, 16, 8, (1, 2), 9, 3),
    (2, 30, 21, 5, 7, (1, 3), 9, 3.9), (2, 6, 6, 16, 24, (2, 1), 17, 8),
    (1, 40, 40, 2, 3, (1, 1), 65, 20), (2, 17, 70, 4, 9, (3, 3), 9, 2),
    (2, 20, 20, 2, 2, (3, 6), 9, 3),
    # narrow maps packed into mosaics (several images per 64-px tile, cells beyond the batch)
    (16, 13, 13, 8, 8, (1, 2), 9, 3.9), (9, 13, 13, 3, 5, (1, 2), 9, 3.9), (12, 27, 27, 4, 6, (1, 2), 9, 3.9),
    (8, 28, 28, 4, 4, (1, 2), 17, 8), (33, 6, 5, 2, 3, (2, 2), 9, 3.9),
    # the gathers' cell-row mode (cells of 13..16 rows stacked in the mosaic: strip height = image height) at every such
    # height, with several 64-column tiles and an odd batch; the gradient kernel's tail strips of 1, 2 and 3 rows
    (20, 40, 14, 3, 4, (1, 2), 9, 3.9), (30, 64, 15, 2, 3, (1, 2), 9, 3.9), (40, 16, 16, 4, 5, (1, 2), 9, 3.9),
    (23, 50, 13, 2, 9, (2, 1), 9, 3.9),
]


def _case(shape, seed):
    N, W, H, S, F, units, ks, m = shape
    G = units[0] * units[1]
    rng = np.random.default_rng(seed)
    x = rng.random((N, S, H, W)).astype(np.float32)
    err = rng.standard_normal((N, F, H, W)).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    lim = ks // 2 - 0.01
    mu1 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    mu2 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    return x, err, w, mu1, mu2, ks


@pytest.mark.parametrize("shape", SHAPES, ids=["x".join(str(v) for v in s[:5]) + "_k%d" % s[6] for s in SHAPES])
@pytest.mark.parametrize("edge", [True, False], ids=["edgemask", "exact"])
def test_oracle_parity(shape, edge):
    x, err, w, mu1, mu2, ks = _case(shape, 1234 + len(str(shape)))
    y, g = run_layer(x, err, w, mu1, mu2, 0.5, ks, unit_testing=edge)
    close(y, O.forward(x, w, mu1, mu2, [0.5]), "y")
    ref = O.backward(x, err, w, mu1, mu2, [0.5], ignore_edge_gradients=edge)
    for got, r, name in zip(g, ref, ("dx", "dw", "dmu1", "dmu2", "dsigma")):
        close(got, r, name, scale_atol=name != "dx")


@pytest.mark.parametrize("sigma", [0.8, 1.2, 1.6])
def test_wide_blur_in_mosaic(sigma):
    """Blur half-width (4, 6, 8) >= the displacement bound: the mosaic gap must follow sigma."""
    x, err, w, mu1, mu2, ks = _case((10, 13, 11, 3, 4, (1, 2), 9, 3.9), 77)
    y, g = run_layer(x, err, w, mu1, mu2, sigma, ks, unit_testing=False)
    close(y, O.forward(x, w, mu1, mu2, [sigma]), "y")
    ref = O.backward(x, err, w, mu1, mu2, [sigma], ignore_edge_gradients=False)
    for got, r, name in zip(g, ref, ("dx", "dw", "dmu1", "dmu2", "dsigma")):
        close(got, r, name, scale_atol=name != "dx")


def test_integer_offsets_are_pure_shifts_and_exact_at_borders():
    """mu integer => bilinear weights are exactly {1,0,0,0}: output = shifted blur, zero fill."""
    rng = np.random.default_rng(3)
    x = rng.random((2, 2, 19, 23)).astype(np.float32)
    w = np.ones((1, 2, 1, 3), np.float32)
    mu1 = np.zeros_like(w); mu2 = np.zeros_like(w)
    mu1[0, :, 0, 0], mu2[0, :, 0, 0] = 3, -2
    mu1[0, :, 0, 1], mu2[0, :, 0, 1] = -4, 3
    err = np.zeros((2, 3, 19, 23), np.float32)
    y, _ = run_layer(x, err, w, mu1, mu2, 0.5, 9)
    blur = O.correlate_same(x, O.get_filters(0.5)[0]).sum(1)
    exp = np.zeros_like(y, dtype=np.float64)
    exp[:, 0, 2:, :20] = blur[:, :17, 3:]
    exp[:, 1, :16, 4:] = blur[:, 3:, :19]
    exp[:, 2] = blur
    close(y, exp, "shifted")
    assert np.all(y[:, 0, :2, :] == 0) and np.all(y[:, 0, :, 20:] == 0)      # bit-exact zero fill
    assert np.all(y[:, 1, 16:, :] == 0) and np.all(y[:, 1, :, :4] == 0)


def test_mu_learning_rate_factor_and_accumulation():
    import torch
    import dau_conv
    x, err, w, mu1, mu2, ks = _case((2, 16, 16, 4, 8, (1, 2), 9, 3), 99)
    _, g1 = run_layer(x, err, w, mu1, mu2, 0.5, ks)
    _, g2 = run_layer(x, err, w, mu1, mu2, 0.5, ks, mu_learning_rate_factor=500.0)
    np.testing.assert_allclose(g2[2], 500.0 * g1[2], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(g2[3], 500.0 * g1[3], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(g2[1], g1[1], rtol=0, atol=0)
    # backward ACCUMULATES into the parameter-gradient buffers (base_dau_conv_layer.cu:340-343)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    t = [torch.from_numpy(a).to(dev) for a in (x, err, w, mu1, mu2, sig)]
    h = dau_conv.get_handle(dev, num_output=8, number_units_x=1, number_units_y=2)
    grads = [torch.full_like(t[2], 2.0) for _ in range(4)]
    h.backward(t[0], t[1], t[2], t[3], t[4], t[5], grads=grads)
    torch.cuda.synchronize()
    np.testing.assert_allclose(grads[0].cpu().numpy(), g1[1] + 2.0, rtol=1e-6, atol=1e-6)


def test_bias_term_bias_gradient_and_unit_bounds_guard():
    """The layer-level options the TF op leaves off: fused bias (base_dau_conv_layer.cu:119-123), bias
    gradient (:188-196) and the in-place parameter guard (:33-49)."""
    import torch
    import dau_conv
    from dau_conv import _lib
    x, err, w, mu1, mu2, ks = _case((3, 21, 17, 4, 5, (1, 2), 9, 3.9), 11)
    mu1[0, 0, 0, 0], mu2[0, 1, 1, 2] = 6.5, -9.0                    # outside +-(4 - 1.0): the guard clips them
    dev = torch.device("cuda:0")
    st = _lib.default_settings(num_output=5, number_units_x=1, number_units_y=2, kernel_size=ks, pad=4,
                               bias_term=1, unit_bounds_guard=1, component_border_bound=1.0, sigma_lower_bound=0.3)
    h = dau_conv.DAULayerHandle(st, dev)
    sig = np.full(w.shape, 0.25, np.float32)                        # below the lower bound: raised to 0.3
    bias = np.linspace(-1, 1, 5).astype(np.float32)
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in
         dict(x=x, err=err, w=w, mu1=mu1, mu2=mu2, sig=sig, bias=bias).items()}
    y = h.forward(t["x"], t["w"], t["mu1"], t["mu2"], t["sig"], bias=t["bias"])
    torch.cuda.synchronize()
    m1c, m2c = np.clip(mu1, -3.0, 3.0), np.clip(mu2, -3.0, 3.0)
    assert np.array_equal(t["mu1"].cpu().numpy(), m1c) and np.array_equal(t["mu2"].cpu().numpy(), m2c)
    assert float(t["sig"].min()) == pytest.approx(0.3)
    ref = O.forward(x, w, m1c, m2c, [0.3]) + bias.reshape(1, -1, 1, 1)
    close(y.cpu().numpy(), ref, "y+bias")
    dbias = torch.full((5,), 1.5, device=dev)
    h.set_sigma_hint(0.3)
    out = h.backward(t["x"], t["err"], t["w"], t["mu1"], t["mu2"], t["sig"], flags=_lib.GRAD_ALL | _lib.GRAD_BIAS,
                     dbias=dbias)
    torch.cuda.synchronize()
    close(dbias.cpu().numpy(), 1.5 + err.sum((0, 2, 3)), "dbias", scale_atol=True)     # accumulated into
    g_ref = O.backward(x, err, w, m1c, m2c, [0.3], ignore_edge_gradients=False)
    for got, r, name in zip(out, g_ref, ("dx", "dw", "dmu1", "dmu2", "dsigma")):
        close(got.cpu().numpy(), r, name, scale_atol=name != "dx")


def test_host_training_step_matches_device_path():
    """dau_train_step_host (chunked, copy/compute overlapped) == dau_forward + dau_backward."""
    import torch
    import dau_conv
    x, err, w, mu1, mu2, ks = _case((11, 20, 14, 5, 6, (1, 2), 9, 3.9), 5)
    y_ref, g_ref = run_layer(x, err, w, mu1, mu2, 0.5, ks)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    wd, m1, m2, sg = [torch.from_numpy(a).to(dev) for a in (w, mu1, mu2, sig)]
    h = dau_conv.get_handle(dev, num_output=6, number_units_x=1, number_units_y=2, kernel_size=ks, pad=4)
    xp, ep = torch.from_numpy(x).pin_memory(), torch.from_numpy(err).pin_memory()
    for chunk in (4, 11, 64):                       # ragged tail, one chunk, chunk > batch
        yp = torch.zeros(y_ref.shape).pin_memory()
        dxp = torch.zeros(x.shape).pin_memory()
        grads = h.train_step_host(xp, ep, wd, m1, m2, sg, 0.5, y_host=yp, dx_host=dxp, chunk_images=chunk)
        torch.cuda.synchronize()
        assert np.array_equal(yp.numpy(), y_ref), chunk          # per-image work: bit-identical
        assert np.array_equal(dxp.numpy(), g_ref[0]), chunk
        for got, r, name in zip(grads, g_ref[1:], ("dw", "dmu1", "dmu2", "dsigma")):
            close(got.cpu().numpy(), r, name, scale_atol=True)   # chunking reorders the fp32 sums


FULL_SIZE = {   # BASELINE.json configs at full size: (N, S, F, units, H, W, kernel_size, mu range)
    "c2": (128, 128, 128, (2, 2), 64, 64, 9, 4.0),
    "c4": (256, 512, 512, (1, 2), 28, 28, 17, 8.0),
    "alexnet_l2": (256, 96, 256, (1, 2), 27, 27, 9, 3.0),
    "alexnet_l4": (256, 384, 384, (1, 2), 13, 13, 9, 3.0),
}


@pytest.mark.parametrize("cfg", sorted(FULL_SIZE))
def test_full_size_properties(cfg):
    """BASELINE configs at FULL size (the oracle cannot run them): size-independent identities.
    (a) sum(E*Y) == sum_units w*dL/dw  (Y is linear in w; a checksum over every unit-pixel)
    (b) w=1, mu=0  =>  Y[:, f] == G * blur(sum_s X_s) for every f.
    (<E, Y(x)> == <dL/dx(E), x> is NOT an identity of the reference: it blurs the error first and
    gathers second (base_dau_conv_layer.cu:278-325), which equals the adjoint only away from the
    image border; dL/dx is pinned by the oracle tests instead.)"""
    import torch
    import dau_conv
    dev = torch.device("cuda:0")
    N, S, F, units, H, W, ks, m = FULL_SIZE[cfg]
    G = units[0] * units[1]
    lim = ks // 2 - 0.01
    gen = torch.Generator(device=dev); gen.manual_seed(5)
    x = torch.rand((N, S, H, W), device=dev, generator=gen)
    e = torch.randn((N, F, H, W), device=dev, generator=gen)
    w = 0.1 * torch.randn((1, S, G, F), device=dev, generator=gen)
    mu1 = (torch.rand((1, S, G, F), device=dev, generator=gen) * 2 * m - m).clamp(-lim, lim)
    mu2 = (torch.rand((1, S, G, F), device=dev, generator=gen) * 2 * m - m).clamp(-lim, lim)
    sg = torch.full((1, S, G, F), 0.5, device=dev)
    a = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=ks // 2)
    y = dau_conv.dau_conv_op(x, w, mu1, mu2, sg, **a)
    dx, dw, dm1, dm2, ds = dau_conv.dau_conv_grad_op(e, x, w, mu1, mu2, sg, **a)
    lhs = float((y.double() * e.double()).sum())
    rhs = float((w.double() * dw.double()).sum())
    scale = float((w.double() * dw.double()).abs().sum())
    assert abs(lhs - rhs) <= 1e-4 * abs(rhs) + 1e-6 * scale, (lhs, rhs, scale)
    for t in (dx, dm1, dm2, ds):
        assert bool(torch.isfinite(t).all())
    del dx, dw, dm1, dm2, ds, y
    # (b)
    ones = torch.ones_like(w); zeros = torch.zeros_like(w)
    y1 = dau_conv.dau_conv_op(x, ones, zeros, zeros, sg, **a)
    k = torch.from_numpy(O.get_filters(0.5)[0].astype(np.float32)).to(dev)[None, None]
    ref = G * torch.nn.functional.conv2d(x.sum(1, keepdim=True), k, padding=3)
    assert float((y1[:, :1] - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
    # every feature saw the same units; features of the gather's split tail wave add their input
    # channels in split_k partial sums, so equality holds to fp32 rounding, not bit for bit
    assert float((y1 - y1[:, :1]).abs().max()) <= 1e-5 * float(ref.abs().max())


# ---- row f3: bias + ReLU fused into the forward store, ReLU mask fused into every error load ------
FUSED_SHAPES = [
    (4, 64, 64, 8, 16, (1, 2), 9, 3),      # identity mosaic, TMA blur path
    (3, 64, 40, 4, 6, (2, 2), 9, 3),       # odd batch (second image of the last pair is not there)
    (2, 30, 21, 5, 7, (1, 3), 9, 3.9),     # odd width: element-wise staging
    (9, 13, 13, 3, 5, (1, 2), 9, 3.9),     # mosaic
    (5, 27, 27, 8, 16, (1, 2), 17, 6),     # mosaic, larger kernel
]


@pytest.mark.parametrize("shape", FUSED_SHAPES, ids=["x".join(str(v) for v in s[:5]) for s in FUSED_SHAPES])
@pytest.mark.parametrize("which", ["all", "input_only", "params_only"])
@pytest.mark.parametrize("edge", [False, True], ids=["exact", "edgemask"])
def test_fused_bias_relu_matches_unfused(shape, which, edge):
    """act(DAUConv(x) + b) and its gradients through dau_conv_fused (one pass of the operator) against
    the un-fused composition dau_conv -> torch add -> torch relu, whose DAU part is oracle-checked
    above.  Forward is the same arithmetic followed by max(0, .): compared tightly."""
    import torch
    import dau_conv
    x, err, w, mu1, mu2, ks = _case(shape, 77)
    N, W, H, S, F, units, _, _ = shape
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(5)
    b = (0.2 * rng.standard_normal(F)).astype(np.float32)
    sig = np.full(w.shape, 0.5, np.float32)
    attrs = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=(ks - 1) // 2,
                 unit_testing=edge)

    def leaves():
        t = [torch.from_numpy(a.copy()).to(dev) for a in (x, w, mu1, mu2, sig, b)]
        t[0].requires_grad_(which != "params_only")
        for v in t[1:]:
            v.requires_grad_(which != "input_only")
        return t

    e = torch.from_numpy(err).to(dev)
    ref = leaves()
    y_ref = torch.relu(dau_conv.dau_conv(*ref[:5], sigma_value=0.5, **attrs) + ref[5].reshape(1, -1, 1, 1))
    y_ref.backward(e)
    got = leaves()
    y = dau_conv.dau_conv_fused(*got[:5], bias=got[5], activation="relu", sigma_value=0.5, **attrs)
    y.backward(e)
    torch.cuda.synchronize()
    assert float((y.detach() < 0).sum()) == 0
    assert 0.05 < float((y.detach() == 0).float().mean()) < 0.95        # the mask really cuts
    close(y.detach().cpu().numpy(), y_ref.detach().cpu().numpy(), "relu(y + b)")
    names = ["dx", "dw", "dmu1", "dmu2", "dsigma", "dbias"]
    for name, g, r in zip(names, got, ref):
        if r.grad is None:
            assert g.grad is None, name
            continue
        close(g.grad.cpu().numpy(), r.grad.cpu().numpy(), name, scale_atol=name != "dx")


def test_fused_activation_needs_top_data_and_none_equals_plain_backward():
    import torch
    import dau_conv
    from dau_conv import _lib
    x, err, w, mu1, mu2, ks = _case((2, 32, 32, 4, 4, (1, 2), 9, 3), 3)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    t = [torch.from_numpy(a).to(dev) for a in (x, w, mu1, mu2, sig)]
    e = torch.from_numpy(err).to(dev)
    h = dau_conv.get_handle(dev, num_output=4, number_units_x=1, number_units_y=2, kernel_size=ks, pad=4)
    h.set_sigma_hint(0.5)
    y = h.forward(*t)
    plain = h.backward(t[0], e, *t[1:])
    same = h.backward(t[0], e, *t[1:], top_data=y, activation=_lib.ACT_NONE)
    for a, b_ in zip(plain, same):
        assert torch.equal(a, b_)
    h.set_activation(_lib.ACT_RELU)
    ws = h.workspace(h.bwd_bytes)
    rc = _lib.lib.dau_backward(h._h, t[0].data_ptr(), e.data_ptr(), *[v.data_ptr() for v in t[1:]],
                               plain[0].data_ptr(), *[g.data_ptr() for g in plain[1:]], None, _lib.GRAD_ALL,
                               ws.data_ptr(), ws.numel(), None)
    assert rc != 0 and "top_data" in _lib.lib.dau_last_error().decode()
    with pytest.raises(dau_conv.DAUError):
        h.set_activation(7)
    h.set_activation(_lib.ACT_NONE)


# ---- row f4: pad < (kernel_size-1)/2, output smaller than the input --------------------------------
CROP_SHAPES = [
    # (N, W, H, S, F, units, kernel, pad, mu_range)
    (4, 64, 64, 4, 8, (1, 2), 9, 2, 3),        # identity mosaic; crop 2 -> 60 x 60
    (3, 40, 24, 3, 5, (2, 2), 9, 0, 3.9),      # "valid" output 32 x 16, odd batch
    (6, 13, 13, 4, 6, (1, 2), 9, 3, 3),        # mosaic; 11 x 11
    (2, 48, 48, 2, 4, (1, 2), 17, 4, 6),       # crop 4 -> 40 x 40: the edge rule fires on the OUTPUT size
]


@pytest.mark.parametrize("shape", CROP_SHAPES, ids=["x".join(str(v) for v in s[:5]) + "_k%d_p%d" % (s[6], s[7])
                                                    for s in CROP_SHAPES])
@pytest.mark.parametrize("edge", [True, False], ids=["edgemask", "exact"])
def test_output_smaller_than_input(shape, edge):
    N, W, H, S, F, units, ks, pad, m = shape
    x, _, w, mu1, mu2, _ = _case((N, W, H, S, F, units, ks, m), 41)
    by, ho = O.output_border(H, ks, pad)
    bx, wo = O.output_border(W, ks, pad)
    assert by == (ks - 1) // 2 - pad and ho == H - 2 * by and wo == W - 2 * bx
    err = np.random.default_rng(8).standard_normal((N, F, ho, wo)).astype(np.float32)
    y, g = run_layer(x, err, w, mu1, mu2, 0.5, ks, pad=pad, number_units_x=units[0], number_units_y=units[1],
                     unit_testing=edge)
    assert y.shape == (N, F, ho, wo) and g[0].shape == x.shape
    close(y, O.forward_cropped(x, w, mu1, mu2, [0.5], ks, pad), "y")
    ref = O.backward_cropped(x, err, w, mu1, mu2, [0.5], ks, pad, ignore_edge_gradients=edge)
    for name, a, b in zip(["dx", "dw", "dmu1", "dmu2", "dsigma"], g, ref):
        close(a, b, name, scale_atol=name != "dx")


def test_pad_larger_than_same_is_refused():
    import dau_conv
    with pytest.raises(dau_conv.DAUError):
        dau_conv.get_handle("cuda:0", num_output=4, kernel_size=9, pad=5)


# ---- seeded random sweep over shapes and flags (the reference's own test draws its cases the same
#      way, dau_conv_test.py:418-437; this widens it to every attribute the op accepts) -------------
def _fuzz_cases(count, seed):
    rng = np.random.default_rng(seed)
    cases = []
    for i in range(count):
        ks = int(rng.choice([3, 5, 9, 9, 17, 33]))
        pad = (ks - 1) // 2 if rng.random() < 0.7 else int(rng.integers(0, (ks - 1) // 2 + 1))
        crop = (ks - 1) // 2 - pad
        H = int(rng.integers(2 * crop + 1, 2 * crop + 40)); W = int(rng.integers(2 * crop + 1, 2 * crop + 80))
        cases.append(dict(
            N=int(rng.integers(1, 8)), H=H, W=W, S=int(rng.integers(1, 10)), F=int(rng.integers(1, 12)),
            ux=int(rng.integers(1, 3)), uy=int(rng.integers(1, 4)), ks=ks, pad=pad,
            sigma=float(rng.choice([0.4, 0.5, 0.5, 0.65, 0.8, 1.0])),
            attrs=dict(unit_testing=bool(rng.integers(0, 2)), use_interpolation=bool(rng.random() < 0.8),
                       single_dim_kernel=bool(rng.random() < 0.15), forbid_positive_dim1=bool(rng.random() < 0.15),
                       square_unit_normalization=bool(rng.random() < 0.2), unit_normalization=bool(rng.random() < 0.85),
                       number_units_ignore=int(rng.integers(0, 2))),
            seed=1000 + i))
    return cases


@pytest.mark.parametrize("case", _fuzz_cases(100, 2024), ids=lambda c: "N%dx%dx%dx%d_F%d_G%d_k%d_p%d_s%g" % (
    c["N"], c["S"], c["H"], c["W"], c["F"], c["ux"] * c["uy"], c["ks"], c["pad"], c["sigma"]))
def test_random_shapes_and_flags(case):
    c = case
    G = c["ux"] * c["uy"]
    a = dict(c["attrs"])
    a["number_units_ignore"] = min(a["number_units_ignore"], G - 1)
    rng = np.random.default_rng(c["seed"])
    x = rng.random((c["N"], c["S"], c["H"], c["W"])).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, c["S"], G, c["F"]))).astype(np.float32)
    lim = c["ks"] // 2 - 0.01
    mu1 = rng.uniform(-lim, lim, w.shape).astype(np.float32)
    mu2 = rng.uniform(-lim, lim, w.shape).astype(np.float32)
    if a["single_dim_kernel"]:
        mu2[:] = 0                                      # DAUConv1d pins mu2 (dau_conv.py:560-571)
    by, ho = O.output_border(c["H"], c["ks"], c["pad"])
    bx, wo = O.output_border(c["W"], c["ks"], c["pad"])
    err = rng.standard_normal((c["N"], c["F"], ho, wo)).astype(np.float32)
    y, g = run_layer(x, err, w, mu1, mu2, c["sigma"], c["ks"], pad=c["pad"], number_units_x=c["ux"],
                     number_units_y=c["uy"], **a)
    kw = dict(num_dau_units_ignore=a["number_units_ignore"], single_dim_kernel=a["single_dim_kernel"],
              forbid_positive_dim1=a["forbid_positive_dim1"], use_interpolation=a["use_interpolation"],
              unit_normalization=a["unit_normalization"], square_unit_normalization=a["square_unit_normalization"])
    close(y, O.forward_cropped(x, w, mu1, mu2, [c["sigma"]], c["ks"], c["pad"], **kw), "y")
    ref = O.backward_cropped(x, err, w, mu1, mu2, [c["sigma"]], c["ks"], c["pad"],
                             ignore_edge_gradients=a["unit_testing"], **kw)
    for name, got, want in zip(["dx", "dw", "dmu1", "dmu2", "dsigma"], g, ref):
        close(got, want, name, scale_atol=name != "dx")


def test_corner_based_offsets_equal_centred_ones():
    """offsets_already_centered = 0 (the Caffe layer's convention): mu counts from the kernel's corner and
    the layer subtracts kernel_size/2 (base_dau_conv_layer.cpp:574-575, :817-818; rot(mu) = ks-1-mu for
    the back-propagation, base_dau_conv_layer.cu:302-314); the guard then clips to [bound, ks-1-bound]
    (:38-42).  Same numbers as the centred layer on mu - ks/2, bit for bit."""
    import torch
    import dau_conv
    from dau_conv import _lib
    x, err, w, mu1, mu2, ks = _case((4, 24, 19, 3, 5, (1, 2), 9, 3.9), 21)
    mu1, mu2 = np.round(mu1 * 1024) / 1024, np.round(mu2 * 1024) / 1024     # mu + ks/2 is then exact in fp32
    mu1, mu2 = mu1.astype(np.float32), mu2.astype(np.float32)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    res = []
    for centred in (1, 0):
        st = _lib.default_settings(num_output=5, number_units_x=1, number_units_y=2, kernel_size=ks, pad=4,
                                   offsets_already_centered=centred, unit_bounds_guard=1 - centred,
                                   component_border_bound=0.01, sigma_lower_bound=0.3)
        h = dau_conv.DAULayerHandle(st, dev)
        h.set_sigma_hint(0.5)
        shift = 0.0 if centred else float(ks // 2)
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x, w, mu1 + shift, mu2 + shift, sig)]
        y = h.forward(*t)
        g = h.backward(t[0], torch.from_numpy(err).to(dev), *t[1:])
        torch.cuda.synchronize()
        if not centred:          # all offsets were inside [0.01, ks-1-0.01]: the guard left them alone
            assert np.array_equal(t[2].cpu().numpy(), (mu1 + shift).astype(np.float32))
        res.append([y.cpu().numpy()] + [v.cpu().numpy() for v in g])
    for a, b in zip(*res):
        assert np.array_equal(a, b)
    close(res[1][0], O.forward(x, w, mu1, mu2, [0.5]), "y")


def test_nan_offsets_are_flagged_and_contribute_nothing():
    """A NaN mu makes the TF op fail the call (dau_conv_op.cpp:258-262).  Here the unit is dropped
    (w = 0) and a sticky flag is raised that dau_take_nan_flag reads and clears; dau_conv_op(check_nan=True)
    turns it into the op's error."""
    import torch
    import dau_conv
    x, err, w, mu1, mu2, ks = _case((2, 20, 12, 3, 4, (1, 2), 9, 3), 9)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    bad = mu1.copy(); bad[0, 1, 0, 2] = np.nan
    w0 = w.copy(); w0[0, 1, 0, 2] = 0.0
    a = dict(num_output=4, number_units_x=1, number_units_y=2, kernel_size=ks, pad=4)
    t = lambda *arrs: [torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in arrs]
    h = dau_conv.get_handle(dev, **a)
    h.take_nan_flag()
    y = dau_conv.dau_conv_op(*t(x, w, bad, mu2, sig), sigma_value=0.5, **a)
    assert h.take_nan_flag() and not h.take_nan_flag()           # raised once, cleared by the read
    close(y.cpu().numpy(), O.forward(x, w0, mu1, mu2, [0.5]), "y without the NaN unit")
    dau_conv.dau_conv_op(*t(x, w, mu1, mu2, sig), sigma_value=0.5, **a)
    assert not h.take_nan_flag()
    with pytest.raises(RuntimeError):
        dau_conv.dau_conv_op(*t(x, w, bad, mu2, sig), sigma_value=0.5, check_nan=True, **a)


def test_offset_hint_selects_the_displacement_bucket():
    """dau_set_offset_hint: with max_kernel_size 33 and |mu| < 3.9 the layer plans the '< 4' bucket the
    reference's op would pick from max|mu| (dau_conv_op.cpp:222-256): same results as the full-kernel
    plan within the tolerance (the tiling differs), much smaller gaps and halos; offsets beyond a
    too-small bound are clamped into its bucket, not undefined."""
    import torch
    import dau_conv
    x, err, w, mu1, mu2, _ = _case((6, 27, 27, 5, 6, (1, 2), 9, 3.9), 31)
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    t = [torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in (x, w, mu1, mu2, sig)]
    e = torch.from_numpy(err).to(dev)
    a = dict(num_output=6, number_units_x=1, number_units_y=2, kernel_size=33, pad=16, unit_testing=True)
    h = dau_conv.get_handle(dev, **a)
    y_ref = O.forward(x, w, mu1, mu2, [0.5])
    g_ref = O.backward(x, err, w, mu1, mu2, [0.5], ignore_edge_gradients=True)
    sizes = []
    for bound in (None, 3.95):
        y = dau_conv.dau_conv_op(*t, sigma_value=0.5, max_offset=bound, **a)
        grads = dau_conv.dau_conv_grad_op(e, *t, sigma_value=0.5, max_offset=bound, **a)
        torch.cuda.synchronize()
        close(y.cpu().numpy(), y_ref, "y bound=%s" % bound)
        for name, got, want in zip(["dx", "dw", "dmu1", "dmu2", "dsigma"], grads, g_ref):
            close(got.cpu().numpy(), want, name, scale_atol=name != "dx")
        sizes.append((h.fwd_bytes, h.bwd_bytes))
    assert sizes[0] != sizes[1]                    # a different plan (gap 4 instead of 16) really ran
    # a bound that is too small clamps the integer offsets into its bucket [-4, 3]
    big1, big2 = (3 * mu1).astype(np.float32), (3 * mu2).astype(np.float32)
    tb = [torch.from_numpy(v).to(dev) for v in (big1, big2)]
    y = dau_conv.dau_conv_op(t[0], t[1], tb[0], tb[1], t[4], sigma_value=0.5, max_offset=3.0, **a)

    def clamp(m):        # what split_offset does: floor clamped to [-4, 3], fraction 0 / 1 at the ends
        return np.clip(m, -4.0, 4.0).astype(np.float32)
    close(y.cpu().numpy(), O.forward(x, w, clamp(big1), clamp(big2), [0.5]), "y with clamped offsets")
    h.set_offset_hint(None)


def test_graphed_step_replays_the_eager_step_bit_for_bit():
    """dau_conv.GraphedStep: forward + backward captured once in a CUDA graph (latency-bound layers, the
    reference's speed-test shape dau_conv_test.py:504-516).  Replays read the CURRENT contents of the captured
    tensors: new data written into them gives the new result, identical to the eager calls."""
    import torch
    import dau_conv
    shape = (4, 16, 16, 8, 6, (2, 1), 9, 3)
    dev = torch.device("cuda:0")
    a = dict(num_output=6, number_units_x=2, number_units_y=1, kernel_size=9, pad=4)
    h = dau_conv.get_handle(dev, **a)
    h.set_sigma_hint(0.5)
    x, err, w, mu1, mu2, ks = _case(shape, 17)
    sig = np.full(w.shape, 0.5, np.float32)
    t = [torch.from_numpy(v).to(dev) for v in (x, err, w, mu1, mu2, sig)]
    gs = dau_conv.GraphedStep(h, *t)
    for seed in (17, 18):
        x2, err2, _, _, _, _ = _case(shape, seed)
        t[0].copy_(torch.from_numpy(x2)); t[1].copy_(torch.from_numpy(err2))
        got = [v.clone() for v in gs.replay()]
        y = h.forward(t[0], t[2], t[3], t[4], t[5])
        ref = (y,) + tuple(h.backward(t[0], t[1], t[2], t[3], t[4], t[5]))
        torch.cuda.synchronize()
        for g_, r_ in zip(got, ref):
            assert torch.equal(g_, r_)
    close(got[0].cpu().numpy(), O.forward(x2, w, mu1, mu2, [0.5]), "y graphed")
