"""DAUConv2d / DAUConv1d / dau_conv2d against the reference's Python
layer semantics (plugins/tensorflow/dau_conv/dau_conv.py) and the oracle."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


def test_grid_mean_values_follow_the_reference_formula():
    from dau_conv.layers import dau_grid_mean
    # dau_conv.py:50 with 2 units, max_value 3: centres of two equal cells of [-3.5, 3.5]
    mu1 = dau_grid_mean((1, 3, 4, 2), (2, 2), 3, 2)
    assert mu1.shape == (1, 3, 4, 2)
    np.testing.assert_allclose(mu1[0, 0, :, 0].numpy(), [-1.75, 1.75, -1.75, 1.75])
    # the reference's tf.tile quirk (dau_conv.py:61-62) interleaves mu2 as well; exact_grid undoes it
    mu2 = dau_grid_mean((1, 3, 4, 2), (2, 2), 3, 1)
    np.testing.assert_allclose(mu2[0, 0, :, 0].numpy(), [-1.75, 1.75, -1.75, 1.75])
    mu2x = dau_grid_mean((1, 3, 4, 2), (2, 2), 3, 1, exact_grid=True)
    np.testing.assert_allclose(mu2x[0, 0, :, 0].numpy(), [-1.75, -1.75, 1.75, 1.75])
    assert float(mu1.std()) > 0 and bool((mu1[0, 1] == mu1[0, 0]).all())      # same for every channel


def test_odd_unit_count_gets_a_zero_weight_dummy_unit():
    from dau_conv import DAUConv2d
    m = DAUConv2d(5, (1, 1), 9, in_channels=3)                      # dau_conv.py:311-327
    assert m.num_dau_units_ignore == 1 and m.num_dau_units_all == 2 and m.dau_units == (1, 2)
    assert m.weights.shape == (1, 3, 2, 5)
    assert bool((m.weights[:, :, -1] == 0).all()) and bool((m.weights[:, :, 0] != 0).any())
    assert m.sigma.shape == (1,) and not m.sigma.requires_grad and float(m.sigma) == 0.5
    assert m._op_attrs()["number_units_ignore"] == 1
    m3 = DAUConv2d(5, (1, 3), 9, in_channels=3)
    assert m3.dau_units[0] * m3.dau_units[1] == m3.num_dau_units_all == 4


def test_lazy_build_and_1d_variant_on_cpu_objects():
    from dau_conv import DAUConv1d, DAUConv2d
    m = DAUConv2d(4, (2, 2), 9)
    assert m.weights is None
    m.build(6)
    assert m.weights.shape == (1, 6, 4, 4) and [n for n, _ in m.named_parameters()] == ["weights", "mu1", "mu2", "sigma", "bias"]
    m1 = DAUConv1d(4, (1, 2), 9, in_channels=2)
    assert m1.dau_unit_single_dim and bool((m1.mu2 == 0).all())
    with pytest.raises(ValueError):
        DAUConv2d(4, (2, 2), 9, data_format="channels_last")


@pytest.mark.gpu
def test_module_forward_backward_matches_oracle():
    from oracle import dau_oracle as O
    from dau_conv import DAUConv2d
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    m = DAUConv2d(6, (2, 2), 9, strides=2, activation=torch.relu, mu_learning_rate_factor=1, in_channels=5).to(dev)
    with torch.no_grad():
        m.mu1.add_(torch.randn_like(m.mu1)); m.mu2.add_(torch.randn_like(m.mu2)); m.bias.uniform_(-0.5, 0.5)
        m.mu1[0, 0, 0, 0] = 7.0                                    # clipped to 4 - 0.01 by the layer
    x = torch.rand((3, 5, 18, 22), device=dev, requires_grad=True)
    y = m(x)
    assert y.shape == (3, 6, 9, 11)
    lim = 4 - 0.01
    w, mu1, mu2 = [t.detach().cpu().numpy() for t in (m.weights, m.mu1.clamp(-lim, lim), m.mu2.clamp(-lim, lim))]
    full = O.forward(x.detach().cpu().numpy(), w, mu1, mu2, [0.5]) + m.bias.detach().cpu().numpy().reshape(1, -1, 1, 1)
    ref = np.maximum(full, 0)[:, :, ::2, ::2]
    np.testing.assert_allclose(y.detach().cpu().numpy(), ref, rtol=1e-4, atol=1e-5)
    g = torch.randn_like(y)
    y.backward(g)
    # error reaching the op: upsampled by the stride slicing, masked by the ReLU
    err = np.zeros_like(full, dtype=np.float32)
    err[:, :, ::2, ::2] = g.cpu().numpy() * (ref > 0)
    dx, dw, dm1, dm2, dsg = O.backward(x.detach().cpu().numpy(), err, w, mu1, mu2, [0.5], ignore_edge_gradients=False)
    np.testing.assert_allclose(x.grad.cpu().numpy(), dx, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(m.weights.grad.cpu().numpy(), dw, rtol=1e-4, atol=1e-4)
    inside = (np.abs(m.mu1.detach().cpu().numpy()) < lim)          # clamp passes no gradient outside
    np.testing.assert_allclose(m.mu1.grad.cpu().numpy(), dm1 * inside, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(m.bias.grad.cpu().numpy(), err.sum((0, 2, 3)), rtol=1e-4, atol=1e-4)
    assert m.sigma.grad is None


@pytest.mark.gpu
def test_functional_wrapper_normaliser_replaces_bias():
    from dau_conv import dau_conv2d
    dev = torch.device("cuda:0")
    x = torch.rand((2, 3, 12, 12), device=dev)
    out, layer = dau_conv2d(x, 4, (2, 2), 9, normalizer_fn=lambda t, scale: t * scale, normalizer_params=dict(scale=2.0))
    assert layer.bias is None and out.shape == (2, 4, 12, 12) and float(out.detach().min()) >= 0.0
    out2, _ = dau_conv2d(x, 4, (2, 2), 9, layer=layer, activation_fn=None, normalizer_fn=None)
    np.testing.assert_allclose(torch.relu(2.0 * out2).detach().cpu().numpy(), out.detach().cpu().numpy(), rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
def test_dynamic_kernel_size_matches_the_static_plan():
    """DAUConv2d(dynamic_kernel_size=True) reads max|mu| and plans its bucket (dau_conv_op.cpp:222-256)."""
    import torch
    import dau_conv
    torch.manual_seed(3)
    dev = torch.device("cuda:0")
    x = torch.rand(4, 6, 24, 24, device=dev)
    a = dau_conv.DAUConv2d(8, (1, 2), 33, in_channels=6, activation=torch.relu).to(dev)
    b = dau_conv.DAUConv2d(8, (1, 2), 33, in_channels=6, activation=torch.relu, dynamic_kernel_size=True).to(dev)
    b.load_state_dict(a.state_dict())
    with torch.no_grad():
        a.mu1.uniform_(-3, 3); a.mu2.uniform_(-3, 3)
        b.mu1.copy_(a.mu1); b.mu2.copy_(a.mu2)
    ya, yb = a(x), b(x)
    ya.square().sum().backward(); yb.square().sum().backward()
    assert torch.allclose(ya, yb, rtol=1e-4, atol=1e-5)
    for pa, pb in zip(a.parameters(), b.parameters()):
        if pa.grad is not None:
            assert torch.allclose(pa.grad, pb.grad, rtol=1e-3, atol=1e-4 * float(pa.grad.abs().max()))
