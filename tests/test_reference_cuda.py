"""Parity of this repo's CUDA path against the REFERENCE'S OWN CUDA kernels, both run on the same GPU
on the same inputs.  oracle/_ref/libdau_ref_cuda.so is the reference's BaseDAUConvLayer<float> built
for sm_100a from the sources under /root/reference (oracle/Makefile.refcuda).  Round 2 compiles all 11
forward and 17 gradient kernel-template shards (oracle/_ref/shards.txt lists what a given build holds):
CASES stay inside the 15 shards round 1 had, EXTRA_CASES dispatch to the other 13 (F not a multiple of
16 / odd S at offsets above 4 px, 1x1 and 8xN / Nx8 patches).  A reference call that ends with "shard
not built" FAILS the test when shards.txt says the build is complete; a reference process that dies
inside its own kernels is recorded as a reference fault (xfail) -- this repo is held to the oracle on
every shape before the reference is asked.

Two bars, both written here:
 * the reference's own acceptance rule for its CUDA op against its NumPy model
   (`_assertMatrix`, plugins/tensorflow/tests/dau_conv_test.py:300-333, with the last output / dL/dx
   column cropped as :398-404 does) -- applied to (ours vs reference CUDA);
 * the north-star tolerance 1e-4 rel / 1e-5 abs (gradients: abs scaled by max(1, max|ref|)) away from
   the right / bottom edge, where the reference's kernels are documented to be inexact
   (include/dau_conv/base_dau_conv_layer.hpp "does not compute accuretely on bottom/right border (1px)").
The oracle (pinned to the reference's NumPy model) arbitrates: the test also reports how far each side
is from it.
"""
import numpy as np
import pytest

import _tolerance
import ref_cuda
from oracle import dau_oracle as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref/libdau_ref_cuda.so not built")]


def reference_assert_matrix(mat, gt, name, rel_tolerance=0.01):
    """dau_conv_test.py:300-333 restated."""
    diff_abs = np.abs(mat - gt)
    diff_rel = np.nan_to_num(diff_abs / np.abs(gt + 1e-9))
    invalid = np.logical_and(diff_rel > 1e-4, diff_abs > 1e-7)
    rate = invalid.sum() / float(diff_rel.size)
    avg = float(diff_rel[invalid].mean()) if invalid.any() else 0.0
    assert avg <= rel_tolerance or rate <= 1e-2, "%s: avg rel diff %.4g over %.3g of the elements" % (name, avg, rate)


def strict(got, ref, name, scale_atol=False):
    return _tolerance.close(got, ref, "refcuda:" + name, scale_atol=scale_atol)


# shapes on which the reference build (as compiled here for sm_100a) itself faults: illegal memory access inside its
# kernels, process abort.  They stay in the list: this repo is still held to the oracle on them; the reference's
# output is not used on them even when a run happens to survive.
REFERENCE_FAULTS = {
    (2, 15, 32, (1, 2), 32, 32, 9),        # odd S
    (2, 16, 32, (1, 2), 64, 64, 33),       # the 16-px offset bucket (8 px and 32 px run)
}


def _all_shards_built():
    import os
    p = os.path.join(os.path.dirname(ref_cuda.SO), "shards.txt")
    return os.path.exists(p) and len(open(p).read().split()) >= 28


# shapes that dispatch to the 13 shards round 1 had not compiled (src/dau_conv/dau_conv_impl/dau_conv_forward.cpp:
# single_feature = F % 16 != 0, single_subfeature = S odd; dau_conv_backward.cpp:15-36: patch 8 / 16 / 32 / 64 per
# axis from the map extent, 1x1 for 1x1 maps).  The reference's authors never ran most of them: a reference
# process that dies is an expected outcome here, a "shard not built" answer is not.
EXTRA_CASES = [
    (2, 15, 32, (1, 2), 32, 32, 17),       # forward off8 s0_f1 ; dL/dx off8 s1_f0
    (2, 16, 40, (1, 2), 32, 32, 17),       # forward off8 s1_f0
    (2, 15, 40, (1, 2), 32, 32, 17),       # forward off8 s1_f1
    (2, 15, 40, (1, 2), 64, 64, 33),       # forward off16 s1_f1
    (4, 16, 32, (1, 2), 8, 16, 9), (4, 16, 32, (1, 2), 16, 8, 9),          # 8x16 / 16x8 patches
    (4, 16, 32, (1, 2), 8, 32, 9), (4, 16, 32, (1, 2), 32, 8, 9),          # 8x32 / 32x8
    (2, 16, 32, (1, 2), 8, 64, 9), (2, 16, 32, (1, 2), 64, 8, 9),          # 8x64 / 64x8
    (2, 16, 32, (1, 2), 16, 64, 9), (2, 16, 32, (1, 2), 64, 16, 9),        # 16x64 / 64x16
    (8, 16, 32, (1, 2), 1, 1, 9),                                          # 1x1 maps
]

CASES = [
    # N, S, F, (Gx, Gy), H, W, kernel_size   (the reference's gradient kernel needs F >= 32)
    (2, 16, 32, (1, 2), 32, 32, 9),
    (3, 16, 32, (2, 2), 32, 32, 9),
    (2, 32, 32, (2, 1), 64, 64, 9),
    (2, 16, 32, (2, 2), 64, 64, 9),
    (2, 32, 32, (1, 2), 27, 27, 9),        # AlexNet-DAU L2 map: partial 32x32 patches
    (4, 16, 32, (1, 2), 13, 13, 9),        # AlexNet-DAU L3-L5 map: 16x16 patches
    (2, 16, 32, (1, 2), 32, 32, 17),       # offsets up to 8 px (the op's kernel-17 bucket)
    (2, 32, 64, (1, 2), 28, 28, 17),       # the C4 map
    (2, 15, 32, (1, 2), 32, 32, 9),        # odd S: the reference's single-subfeature / single-feature kernel variants
    (2, 16, 32, (1, 2), 64, 32, 9),        # rectangular maps: 32x64 and 64x32 patches
    (2, 16, 32, (2, 1), 32, 64, 9),
    (2, 16, 32, (1, 2), 16, 32, 9),
    (4, 16, 32, (1, 2), 8, 8, 9),          # 8x8 patches
    (2, 16, 32, (1, 2), 64, 64, 33),       # offsets up to 16 px (kernel 33)
    (2, 16, 32, (1, 2), 64, 64, 65),       # offsets up to 32 px (kernel 65)
]


@pytest.mark.parametrize("case", CASES + EXTRA_CASES,
                         ids=lambda c: "N%dS%dF%dG%dx%d_%dx%d_k%d" % (c[0], c[1], c[2], c[3][0], c[3][1], c[4], c[5], c[6]))
def test_matches_reference_cuda(case, tmp_path):
    import torch
    import dau_conv
    N, S, F, units, H, W, ks = case
    m = ks // 2 - 0.01
    G = units[0] * units[1]
    rng = np.random.default_rng(99 + N + S + F + H)
    x = rng.random((N, S, H, W)).astype(np.float32)
    err = rng.standard_normal((N, F, H, W)).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    mu1 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
    mu2 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
    sigma = np.full((1, S, G, F), 0.5, np.float32)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(a).to(dev) for a in (x, w, mu1, mu2, sigma)]
    te = torch.from_numpy(err).to(dev)

    attrs = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=(ks - 1) // 2,
                 unit_testing=True, mu_learning_rate_factor=1.0)
    y = dau_conv.dau_conv_op(*t, **attrs).cpu().numpy()
    ours = [y] + [g.cpu().numpy() for g in dau_conv.dau_conv_grad_op(te, *t, **attrs)]

    # (3) the oracle arbitrates on the whole maps: this repo's path must be inside tolerance everywhere
    y_o = O.forward(x, w, mu1, mu2, [0.5])
    g_o = O.backward(x, err, w, mu1, mu2, [0.5], ignore_edge_gradients=True)
    strict(ours[0], y_o, "top vs oracle")
    strict(ours[1], g_o[0], "bottom_error vs oracle", scale_atol=True)
    for k, name in enumerate(["dw", "dmu1", "dmu2", "dsigma"]):
        strict(ours[2 + k], g_o[1 + k], name + " vs oracle", scale_atol=True)

    try:
        r = ref_cuda.run_isolated(x, err, w, mu1, mu2, sigma, units, tmp_path, kernel_size=ks,
                                  ignore_edge_gradients=True)
    except ref_cuda.ReferenceFault as ex:
        if "not built" in str(ex):
            assert not _all_shards_built(), "complete reference build still answers 'shard not built': %s" % ex
            pytest.skip("this oracle/_ref build lacks the shard this shape dispatches to (%s)" % ex)
        if case in REFERENCE_FAULTS or case in EXTRA_CASES:
            pytest.xfail("reference fault: the reference's own CUDA build dies on this shape (%s); this repo "
                         "matches the oracle" % ex)
        raise
    if case in REFERENCE_FAULTS:
        # an out-of-bounds access does not fault on every box: when the reference survives, its numbers on such a
        # shape are not evidence either way, and this repo has already been held to the oracle above
        pytest.xfail("the reference's CUDA build ran a shape it usually faults on; its output is not trusted")
    ref = list(r[:6])

    names = ["top", "bottom_error", "dw", "dmu1", "dmu2", "dsigma"]
    # (1) the reference's own rule, its crops
    for k, (a, b) in enumerate(zip(ours, ref)):
        if k < 2:
            a, b = a[..., :-1], b[..., :-1]
        reference_assert_matrix(a, b, names[k])
    # (2) the north-star tolerance where the reference's kernels are exact: everything but the last
    # column and row of the two maps; the parameter gradients in full (edge mask on, as in the reference's tests)
    strict(ours[0][..., :-1, :-1], ref[0][..., :-1, :-1], "top (interior)")
    strict(ours[1][..., :-1, :-1], ref[1][..., :-1, :-1], "bottom_error (interior)", scale_atol=True)
    for k in range(2, 6):
        strict(ours[k], ref[k], names[k], scale_atol=True)


def test_full_size_c2_matches_reference_cuda(tmp_path):
    """BASELINE config C2 at its full size (N=128, S=F=128, 64x64, G=4): every output of one forward + backward
    pass against the reference's CUDA kernels on the same tensors.  Sums run over up to 524 288 terms in different
    orders on the two sides, so the absolute term of the tolerance is scaled by max(1, max|ref|) for all six."""
    import torch
    import bench
    import dau_conv
    N, S, F, units, H, W, ks, m = bench.CONFIGS["c2"]
    G = units[0] * units[1]
    dev = torch.device("cuda:0")
    host = bench.make_inputs_numpy("c2", N, 1234)
    x, e, w, mu1, mu2 = [torch.from_numpy(a).to(dev) for a in host]
    sigma = torch.full((1, S, G, F), 0.5, device=dev)
    attrs = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=(ks - 1) // 2,
                 unit_testing=True, mu_learning_rate_factor=1.0)
    y = dau_conv.dau_conv_op(x, w, mu1, mu2, sigma, **attrs)
    g = dau_conv.dau_conv_grad_op(e, x, w, mu1, mu2, sigma, **attrs)
    y, g = y.cpu().numpy(), [v.cpu().numpy() for v in g]
    del x, e
    torch.cuda.empty_cache()
    r = ref_cuda.run_isolated(*host, np.full((1, S, G, F), 0.5, np.float32), units, tmp_path, kernel_size=ks,
                              ignore_edge_gradients=True, timeout=600)
    strict(y[..., :-1, :-1], r[0][..., :-1, :-1], "top (interior)", scale_atol=True)
    strict(g[0][..., :-1, :-1], r[1][..., :-1, :-1], "bottom_error (interior)", scale_atol=True)
    for k, name in enumerate(["dw", "dmu1", "dmu2", "dsigma"]):
        strict(g[1 + k], r[2 + k], name, scale_atol=True)


@pytest.mark.parametrize("sigma", [0.5, 0.8, 1.2, 1.6])
@pytest.mark.parametrize("flags,opts", [
    (0, {}), (1, {"single_dimension_kernel": 1}), (2, {"forbid_positive_dim1": 1}),
    (4, {"unit_normalization": 0}), (8, {"square_unit_normalization": 1}),
], ids=["default", "single_dim", "forbid_positive_dim1", "no_unit_norm", "square_unit_norm"])
def test_kernel_compute_matches_reference_cuda(sigma, flags, opts, tmp_path):
    """Row a1: dau_get_kernels (host, double) against the reference's DAUKernelCompute::get_kernels (its CUDA
    kernels + cub segmented sums, fp32) for the same sigma and options: blur, the four derivative filters and the
    rotated error filter, Kb = 7 ... 17.  fp32 sums of exp() on the reference side: 1e-4 relative, 1e-6 absolute."""
    import dau_conv
    ref = ref_cuda.kernels_isolated(sigma, flags, tmp_path)
    got = dau_conv.get_kernels(dau_conv.default_settings(**opts), sigma)
    assert ref.shape == (6,) + got[0].shape
    for name, g, r in zip(["blur", "d_w", "d_mu1", "d_mu2", "d_sigma", "d_error"], got, ref):
        np.testing.assert_allclose(g, r, rtol=1e-4, atol=1e-6, err_msg=name)


OPTION_CASES = [
    # id, op attrs (TF names), reference driver arguments
    # use_interpolation = false is not here: the reference's kernels refuse it unless built with
    # -DDAU_ALLOW_INTERPOLATION_OFF ("support for non-interpolation disabled"); the oracle covers it (test_parity_gpu.py)
    ("single_dim_kernel", dict(single_dim_kernel=True), dict(opt_flags=2)),
    ("forbid_positive_dim1", dict(forbid_positive_dim1=True), dict(opt_flags=4)),
    ("one_unit_ignored", dict(number_units_ignore=1), dict(units_ignore=1)),
    ("mu_learning_rate_factor_500", dict(mu_learning_rate_factor=500.0), dict(mu_learning_rate_factor=500.0)),
]


@pytest.mark.parametrize("name,attrs,ref_kw", OPTION_CASES, ids=[c[0] for c in OPTION_CASES])
def test_op_attributes_match_reference_cuda(name, attrs, ref_kw, tmp_path):
    """The op attributes beyond the default path (row f4), each against the reference's CUDA layer driven the way
    its TF ops drive it (dau_conv_op.cpp:283-290, dau_conv_grad_op.cpp:270-303)."""
    import torch
    import dau_conv
    N, S, F, units, H, W, ks = 2, 16, 32, (1, 2), 32, 32, 9
    G = 2
    rng = np.random.default_rng(4711)
    x = rng.random((N, S, H, W)).astype(np.float32)
    err = rng.standard_normal((N, F, H, W)).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    mu1 = rng.uniform(-3.99, 3.99, (1, S, G, F)).astype(np.float32)
    mu2 = rng.uniform(-3.99, 3.99, (1, S, G, F)).astype(np.float32)
    if name == "single_dim_kernel":
        mu2[:] = 0                      # DAUConv1d keeps mu2 at zero (dau_conv.py:716-724)
    if name == "one_unit_ignored":
        # The reference's NumPy model (and this repo) leave ignored units out of the forward sum
        # (dau_conv_test.py:34), its CUDA layer only zeroes their gradients (base_dau_conv_layer.cu:346-350);
        # the two agree because the layer initialises the dummy unit's weight to zero and never updates it
        # (ZeroNLast, dau_conv.py:77-110) -- the state this case reproduces.
        w[:, :, -1, :] = 0
    sigma = np.full((1, S, G, F), 0.5, np.float32)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(a).to(dev) for a in (x, w, mu1, mu2, sigma)]
    a = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=4,
             unit_testing=True, mu_learning_rate_factor=1.0)
    a.update(attrs)
    ours = [dau_conv.dau_conv_op(*t, **a).cpu().numpy()]
    ours += [g.cpu().numpy() for g in dau_conv.dau_conv_grad_op(torch.from_numpy(err).to(dev), *t, **a)]
    ref = ref_cuda.run_isolated(x, err, w, mu1, mu2, sigma, units, tmp_path, kernel_size=ks,
                                ignore_edge_gradients=True, **ref_kw)[:6]
    names = ["top", "bottom_error", "dw", "dmu1", "dmu2", "dsigma"]
    # the 1-D layer's crop is two columns (dau_conv_test.py:744-751)
    crop = 2 if name == "single_dim_kernel" else 1
    for k, (o, r) in enumerate(zip(ours, ref)):
        if k < 2:
            o, r = o[..., :-crop], r[..., :-crop]
        reference_assert_matrix(o, r, names[k])
    strict(ours[0][..., :-1, :-crop], ref[0][..., :-1, :-crop], "top (interior)")
    strict(ours[1][..., :-1, :-crop], ref[1][..., :-1, :-crop], "bottom_error (interior)", scale_atol=True)
    for k in range(2, 6):
        if name == "single_dim_kernel" and names[k] == "dmu2":
            continue                    # not propagated by the 1-D op (dau_conv_grad_op.cpp:291-292)
        strict(ours[k], ref[k], names[k], scale_atol=True)
