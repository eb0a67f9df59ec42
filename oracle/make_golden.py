#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the REFERENCE's own NumPy model.

Runs only in the authoring container (needs /root/reference).  It lifts the class
``DAUConvPython`` verbatim from
``/root/reference/plugins/tensorflow/tests/dau_conv_test.py:13-295`` with ``exec``
(the module itself imports TensorFlow/pylab, which are absent, so only the class
body is executed) and stores seeded inputs together with the outputs that class
produces.  Nothing from the reference is copied into this repository; only the
numbers it computes are.

Two textual patches are applied to the lifted source, both Python-2 -> Python-3
index fixes in branches the reference's own tests reach only under Python 2
(dau_conv_test.py:189,196: ``shape[0]/2`` used as an index):
    ``valid_filter.shape[0]/2``  ->  ``valid_filter.shape[0]//2``
    ``valid_filter.shape[1]/2``  ->  ``valid_filter.shape[1]//2``

Usage:  python oracle/make_golden.py     (rewrites tests/golden/)
"""
import os
import sys
import warnings

import numpy as np

REF = "/root/reference/plugins/tensorflow/tests/dau_conv_test.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_reference_class():
    with open(REF) as fh:
        lines = fh.readlines()
    src = "".join(lines[12:295])  # file lines 13..295: class DAUConvPython
    src = src.replace("valid_filter.shape[0]/2", "valid_filter.shape[0]//2")
    src = src.replace("valid_filter.shape[1]/2", "valid_filter.shape[1]//2")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from scipy.ndimage import correlate
    ns = {"np": np, "correlate": correlate}
    exec(compile(src, REF, "exec"), ns)
    return ns["DAUConvPython"]


# name: (N, S, F, G, H, W, mu_range, sigma, kwargs)
CASES = {
    "basic_8x10":      (2, 3, 4, 2, 8, 10, 3.0, 0.5, {}),
    "nointerp_8x10":   (2, 3, 4, 2, 8, 10, 3.0, 0.5, {"use_interpolation": False}),
    "ignore1_9x7":     (1, 2, 3, 2, 9, 7, 2.5, 0.5, {"num_dau_units_ignore": 1}),
    "edge16_g4":       (2, 2, 2, 4, 16, 16, 3.9, 0.5, {}),
    "noedge16_g4":     (2, 2, 2, 4, 16, 16, 3.9, 0.5, {"unit_testing": False}),
    "bigmu_12x13":     (1, 2, 2, 2, 12, 13, 7.9, 0.5, {}),
    "sigma08_10x9":    (1, 2, 2, 2, 10, 9, 3.0, 0.8, {}),
    "singledim_6x12":  (2, 2, 3, 2, 6, 12, 3.0, 0.5, {"single_dim_kernel": True}),
    "forbidpos_6x12":  (2, 2, 3, 2, 6, 12, 3.0, 0.5, {"aggr_forbid_positive": True}),
    "quick_1x8x8":     (1, 4, 4, 2, 8, 8, 3.0, 0.5, {}),
    "wide_65x8":       (2, 3, 2, 2, 8, 65, 3.0, 0.5, {}),
}


def main():
    if not os.path.exists(REF):
        sys.exit("reference not mounted; golden vectors can only be generated in the authoring container")
    cls = load_reference_class()
    os.makedirs(OUT, exist_ok=True)
    for idx, (name, (N, S, F, G, H, W, m, sigma, kw)) in enumerate(sorted(CASES.items())):
        rng = np.random.default_rng(4242 + idx)
        # inputs are exactly representable in fp32 so the CUDA path sees the same numbers
        x = rng.random((N, S, H, W)).astype(np.float32)
        err = rng.standard_normal((N, F, H, W)).astype(np.float32)
        w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
        mu1 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
        mu2 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
        if kw.get("single_dim_kernel"):
            mu2[:] = 0.0  # DAUConv1d keeps mu2 == 0 (dau_conv.py DAUConv1d)
        fkw = {k: v for k, v in kw.items() if k != "unit_testing"}
        ref = cls()
        x64, e64 = x.astype(np.float64), err.astype(np.float64)
        w64, m1, m2 = w.astype(np.float64), mu1.astype(np.float64), mu2.astype(np.float64)
        y = ref.forward_cpu(x64, w64, m1, m2, [sigma], **fkw)
        # backward_cpu mutates `error` in place (edge mask) -> hand it a copy
        dx, dw, dmu1, dmu2, dsig = ref.backward_cpu(x64, e64.copy(), w64, m1, m2, [sigma], **kw)
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            x=x, err=err, w=w, mu1=mu1, mu2=mu2, sigma=np.float32(sigma),
            y=np.asarray(y, np.float64), dx=np.asarray(dx, np.float64), dw=np.asarray(dw, np.float64),
            dmu1=np.asarray(dmu1, np.float64), dmu2=np.asarray(dmu2, np.float64),
            dsigma=np.asarray(dsig, np.float64),
            kwargs=np.array(repr(kw)))
        print("wrote", name, "y", np.abs(y).max(), "dw", np.abs(dw).max())
    # the six 9x9 filters of the reference at the sigmas used above
    for sigma in (0.5, 0.8):
        f = cls()._get_filters(sigma)
        np.savez_compressed(os.path.join(OUT, "filters_sigma%02d.npz" % int(sigma * 10)),
                            **{n: np.asarray(a, np.float64) for n, a in
                               zip(("blur", "d_w", "d_mu1", "d_mu2", "d_sigma", "d_error"), f)})


if __name__ == "__main__":
    main()
