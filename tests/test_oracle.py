"""The oracle (oracle/dau_oracle.py) against the reference's own numbers.

tests/golden/*.npz were produced by executing the reference's DAUConvPython class
verbatim (oracle/make_golden.py; dau_conv_test.py:13-295).  The reference model
hard-codes a 9x9 filter support (dau_conv_test.py:178-180), so the oracle is called
with kb=9 here; a separate test shows Kb=2*ceil(5*sigma)+1 (what the CUDA layer
uses, base_dau_conv_layer.cpp:146-147) changes nothing measurable at sigma=0.5.
"""
import ast
import glob
import os

import numpy as np
import pytest

from oracle import dau_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(p for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))
               if not os.path.basename(p).startswith("filters_"))


def _kwargs(d):
    kw = ast.literal_eval(str(d["kwargs"]))
    out = {}
    out["use_interpolation"] = kw.get("use_interpolation", True)
    out["num_dau_units_ignore"] = kw.get("num_dau_units_ignore", 0)
    out["single_dim_kernel"] = kw.get("single_dim_kernel", False)
    out["forbid_positive_dim1"] = kw.get("aggr_forbid_positive", False)
    return out, kw.get("unit_testing", True)


def test_golden_present():
    assert len(CASES) >= 10


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[:-4] for p in CASES])
def test_oracle_matches_reference_model(path):
    d = np.load(path)
    kw, unit_testing = _kwargs(d)
    args = [d[k].astype(np.float64) for k in ("w", "mu1", "mu2")]
    sigma = [float(d["sigma"])]
    y = O.forward(d["x"], *args, sigma, kb=9, **kw)
    # the reference accumulates the forward in float32 (dau_conv_test.py:24)
    np.testing.assert_allclose(y, d["y"], rtol=2e-5, atol=2e-6)
    dx, dw, dmu1, dmu2, dsig = O.backward(d["x"], d["err"], *args, sigma, kb=9,
                                           ignore_edge_gradients=unit_testing, **kw)
    np.testing.assert_allclose(dx, d["dx"], rtol=2e-5, atol=2e-6)
    for got, name in ((dw, "dw"), (dmu1, "dmu1"), (dmu2, "dmu2"), (dsig, "dsigma")):
        np.testing.assert_allclose(got, d[name], rtol=2e-5, atol=2e-5, err_msg=name)


@pytest.mark.parametrize("sigma", [0.5, 0.8])
def test_filters_match_reference(sigma, golden_dir):
    d = np.load(os.path.join(golden_dir, "filters_sigma%02d.npz" % int(sigma * 10)))
    f = O.get_filters(sigma, kb=9)
    for got, name in zip(f, ("blur", "d_w", "d_mu1", "d_mu2", "d_sigma", "d_error")):
        np.testing.assert_allclose(got, d[name], rtol=1e-12, atol=1e-15, err_msg=name)


def test_layer_kernel_size_rule_is_equivalent_at_sigma_half():
    assert O.blur_kernel_size(0.5) == 7 and O.blur_kernel_size(0.8) == 9
    f7 = O.get_filters(0.5)
    f9 = O.get_filters(0.5, kb=9)
    for a, b in zip(f7, f9):
        assert np.abs(a - b[1:-1, 1:-1]).max() < 1e-11
        assert np.abs(b[0]).max() < 1e-11 and np.abs(b[:, 0]).max() < 1e-11


def test_correlate_matches_scipy():
    scipy_ndimage = pytest.importorskip("scipy.ndimage")
    rng = np.random.default_rng(0)
    x = rng.random((2, 3, 9, 11))
    k = rng.random((7, 7))
    ref = np.stack([[scipy_ndimage.correlate(x[n, s], k, mode="constant") for s in range(3)]
                    for n in range(2)])
    np.testing.assert_allclose(O.correlate_same(x, k), ref, rtol=1e-12, atol=1e-14)


def test_hand_checkable_anchors():
    """w=1, mu=0, G=2  =>  Y = 2 * sum_s blur(X_s);   integer mu => pure shift."""
    rng = np.random.default_rng(1)
    x = rng.random((1, 3, 8, 9))
    w = np.ones((1, 3, 2, 2)); z = np.zeros_like(w)
    y = O.forward(x, w, z, z, [0.5])
    blur = O.correlate_same(x, O.get_filters(0.5)[0])
    np.testing.assert_allclose(y[0, 0], 2 * blur[0].sum(0), rtol=1e-12)
    mu1 = np.full_like(w, 2.0); mu2 = np.full_like(w, -1.0)
    y2 = O.forward(x, w, mu1, mu2, [0.5])
    exp = np.zeros((8, 9))
    exp[1:, :7] = 2 * blur[0].sum(0)[:-1, 2:]
    np.testing.assert_allclose(y2[0, 1], exp, rtol=1e-12, atol=1e-14)


def test_edge_mask_rule():
    assert O.edge_mask_flags(8, 65) == (True, False)
    assert O.edge_mask_flags(64, 64) == (True, True)
    assert O.edge_mask_flags(27, 13) == (False, False)
    assert O.edge_mask_flags(6, 6) == (False, False)
    assert O.edge_mask_flags(96, 48) == (False, False)   # 96 % 64 != 0, 48 % 32 != 0


def test_backward_is_gradient_of_forward_in_w():
    """dL/dw from the oracle equals the analytic derivative of the oracle forward (linear in w)."""
    rng = np.random.default_rng(2)
    N, S, F, G, H, W = 1, 2, 2, 2, 7, 6
    x = rng.random((N, S, H, W)); e = rng.standard_normal((N, F, H, W))
    w = rng.standard_normal((1, S, G, F)); mu1 = rng.uniform(-2, 2, w.shape); mu2 = rng.uniform(-2, 2, w.shape)
    _, dw, *_ = O.backward(x, e, w, mu1, mu2, [0.5], ignore_edge_gradients=False)
    for idx in [(0, 0, 0, 0), (0, 1, 1, 1), (0, 1, 0, 1)]:
        w1 = np.zeros_like(w); w1[idx] = 1.0
        y = O.forward(x, w1, mu1, mu2, [0.5])
        assert abs((y * e).sum() - dw[idx]) < 1e-9
