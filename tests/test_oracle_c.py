"""oracle/dau_cpu.c (the timed CPU baseline) against the NumPy oracle and the golden fixtures."""
import ast
import glob
import os

import numpy as np
import pytest

from oracle import dau_cpu, dau_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(p for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))
               if not os.path.basename(p).startswith("filters_"))


@pytest.mark.parametrize("sigma", [0.5, 0.8, 1.3])
def test_filters(sigma):
    sigma = float(np.float32(sigma))       # the C side takes sigma as fp32, like the reference
    for a, b in zip(dau_cpu.filters(sigma), O.get_filters(sigma)):
        np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-15)
    for a, b in zip(dau_cpu.filters(sigma, True, True, True, True), O.get_filters(sigma, None, True, True, True, True)):
        np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-15)


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[:-4] for p in CASES])
def test_c_restatement_matches_golden(path):
    d = np.load(path)
    kw = ast.literal_eval(str(d["kwargs"]))
    common = dict(units_ignore=kw.get("num_dau_units_ignore", 0), interp=kw.get("use_interpolation", True),
                  single_dim=kw.get("single_dim_kernel", False), forbid_pos=kw.get("aggr_forbid_positive", False))
    H, W = d["x"].shape[2:]
    row, col = O.edge_mask_flags(H, W) if kw.get("unit_testing", True) else (False, False)
    y = dau_cpu.forward(d["x"], d["w"], d["mu1"], d["mu2"], float(d["sigma"]), **common)
    np.testing.assert_allclose(y, d["y"], rtol=1e-4, atol=1e-5)
    g = dau_cpu.backward(d["x"], d["err"], d["w"], d["mu1"], d["mu2"], float(d["sigma"]),
                         mask_last_row=row, mask_last_col=col, **common)
    for got, name in zip(g, ("dx", "dw", "dmu1", "dmu2", "dsigma")):
        np.testing.assert_allclose(got, d[name], rtol=1e-4, atol=2e-5, err_msg=name)


def test_c_restatement_matches_numpy_oracle_on_quick_shape():
    rng = np.random.default_rng(11)
    N, S, F, G, H, W = 2, 5, 6, 3, 13, 27
    x = rng.random((N, S, H, W)).astype(np.float32)
    e = rng.standard_normal((N, F, H, W)).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    mu1 = rng.uniform(-3.9, 3.9, w.shape).astype(np.float32)
    mu2 = rng.uniform(-3.9, 3.9, w.shape).astype(np.float32)
    np.testing.assert_allclose(dau_cpu.forward(x, w, mu1, mu2, 0.5), O.forward(x, w, mu1, mu2, [0.5]),
                               rtol=1e-4, atol=1e-5)
    for a, b in zip(dau_cpu.backward(x, e, w, mu1, mu2, 0.5),
                    O.backward(x, e, w, mu1, mu2, [0.5], ignore_edge_gradients=False)):
        np.testing.assert_allclose(a, b, rtol=1e-4, atol=2e-5)
