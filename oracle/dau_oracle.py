"""CPU oracle for the DAU convolution hot path (TEST INFRASTRUCTURE ONLY).

This module is a NumPy restatement of the reference's own numerical model of the
DAU layer.  It is the *checker*: only ``tests/``, ``__graft_entry__.smoke()`` and
the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.
The product path (``dau-convnet_b200/``) never does, and fails loudly when the
CUDA library is missing.

Parity status: PINNED.  ``tests/golden/*.npz`` hold outputs produced by executing
the reference's ``DAUConvPython`` class *verbatim* from
``/root/reference/plugins/tensorflow/tests/dau_conv_test.py:13-295`` (see
``oracle/make_golden.py``); ``tests/test_oracle.py`` checks this restatement
against every one of them.

What it follows (paths relative to /root/reference):
  * forward / backward math ......... plugins/tensorflow/tests/dau_conv_test.py:13-295
  * filter construction ............. src/dau_conv/base_dau_conv_layer.cu:387-448, 537-710
  * correlation convention .......... src/dau_conv/util/convolve.cu:107-129
  * edge masks (ignore_edge_grads) .. dau_conv_test.py:110-136, base_dau_conv_layer.cpp:754-791
  * x w post-scaling of mu/sigma .... base_dau_conv_layer.cu:335-337, dau_conv_test.py:289-292
  * pad < (ks-1)/2 (output smaller) .. base_dau_conv_layer.cpp:20-22, 544-545, 938-949; .cu:258-266

All arithmetic is float64 unless ``dtype`` says otherwise; callers cast.
Tensor layouts are the reference's: x [N,S,H,W], w/mu1/mu2 [1,S,G,F] (F fastest).
"""
from __future__ import annotations

import math

import numpy as np

__all__ = [
    "blur_kernel_size", "get_filters", "correlate_same", "offset_and_sum",
    "offset_and_dot", "edge_mask_flags", "forward", "backward",
    "output_border", "forward_cropped", "backward_cropped",
]


def blur_kernel_size(sigma: float) -> int:
    """Kb = 2*ceil(5*sigma)+1 with the product in fp32, as the reference's Dtype=float
    arithmetic takes it (base_dau_conv_layer.cpp:146-147): 5*0.8f == 4.0f -> Kb = 9."""
    return 2 * int(math.ceil(float(np.float32(5.0) * np.float32(sigma)))) + 1


def get_filters(sigma, kb=None, unit_normalization=True, square_unit_normalization=False,
                single_dim_kernel=False, forbid_positive_dim1=False):
    """DAUKernelCompute for one unit with w=1, mu=(0,0).

    Follows BaseDAUKernelCompute::get_kernels (base_dau_conv_layer.cu:537-710) and the
    NumPy ``_get_filters`` (dau_conv_test.py:177-220).  Returns
    (blur, d_w, d_mu1, d_mu2, d_sigma, d_error), each [kb, kb] float64; axis 0 is y
    (mu2 direction), axis 1 is x (mu1 direction).
    """
    sigma = float(sigma)
    if kb is None:
        kb = blur_kernel_size(sigma)
    c = kb // 2
    ax = np.arange(kb, dtype=np.float64) - c
    dx = np.tile(ax, (kb, 1))            # varies along axis 1
    dy = dx.T
    d2 = dx * dx + dy * dy
    g = np.exp(-d2 / (2.0 * sigma * sigma))
    if single_dim_kernel:                # .cu:432-434 -> only the centre row survives
        g = g * (dy == 0)
    if forbid_positive_dim1:             # .cu:436-438 -> taps right of centre are zero
        g = g * (dx <= 0)

    d_mu1 = dx / sigma ** 2 * g
    d_mu2 = dy / sigma ** 2 * g
    d_sig = d2 / sigma ** 3 * g

    if not unit_normalization:           # .cu:612-616, 642-644
        norm = 1.0
        s_mu1 = s_mu2 = s_sig = 0.0
    elif square_unit_normalization:      # .cu:618-634, 646-651
        norm = 1.0 / np.sum(g * g)
        s_mu1 = 2.0 * np.sum(d_mu1 * g)
        s_mu2 = 2.0 * np.sum(d_mu2 * g)
        s_sig = 2.0 * np.sum(d_sig * g)
    else:                                # .cu:635-640, 652-655
        norm = 1.0 / np.sum(g)
        s_mu1 = np.sum(d_mu1)
        s_mu2 = np.sum(d_mu2)
        s_sig = np.sum(d_sig)
    # .cu:661-662: sums of mu1/mu2 with |.| <= 1e-10 are flushed to zero
    if abs(s_mu1) <= 1e-10:
        s_mu1 = 0.0
    if abs(s_mu2) <= 1e-10:
        s_mu2 = 0.0

    d_w = g * norm                                   # .cu:669
    d_mu1 = norm * d_mu1 - (s_mu1 * norm) * d_w      # .cu:676-683 (w == 1)
    d_mu2 = norm * d_mu2 - (s_mu2 * norm) * d_w
    d_sig = norm * d_sig - (s_sig * norm) * d_w
    blur = g * norm                                  # .cu:688-693 (S=G=F=1)
    d_err = blur[::-1, ::-1].copy()                  # .cu:699-704 mirror
    return blur, d_w, d_mu1, d_mu2, d_sig, d_err


def correlate_same(x, filt):
    """Zero-padded 'same' 2-D correlation over the last two axes.

    out[..., y, x] = sum_{v,u} in[..., y+v-c, x+u-c] * filt[v, u]
    (convolve.cu:107-124; identical to scipy.ndimage.correlate(mode='constant')).
    """
    kb_h, kb_w = filt.shape
    ch, cw = kb_h // 2, kb_w // 2
    H, W = x.shape[-2:]
    xp = np.zeros(x.shape[:-2] + (H + 2 * ch, W + 2 * cw), dtype=np.float64)
    xp[..., ch:ch + H, cw:cw + W] = x
    out = np.zeros(x.shape, dtype=np.float64)
    for v in range(kb_h):
        for u in range(kb_w):
            f = filt[v, u]
            if f != 0.0:
                out += f * xp[..., v:v + H, u:u + W]
    return out


def _taps(mu1, mu2, use_interpolation):
    """floor / frac split of one unit's offsets -> [(dy, dx, weight)] (dau_conv_test.py:36-56)."""
    ox = math.floor(mu1)
    oy = math.floor(mu2)
    fx = mu1 - ox
    fy = mu2 - oy
    if not use_interpolation:
        return int(ox), int(oy), [(0, 0, 1.0)]
    return int(ox), int(oy), [(0, 0, (1 - fx) * (1 - fy)), (0, 1, fx * (1 - fy)),
                              (1, 0, (1 - fx) * fy), (1, 1, fx * fy)]


def _pad_for(mu1, mu2):
    m = max(float(np.max(np.abs(mu1))), float(np.max(np.abs(mu2))))
    return int(math.ceil(m + 1)) + 1


def offset_and_sum(xb, w, mu1, mu2, num_dau_units_ignore=0, use_interpolation=True):
    """Y[n,f] = sum_{s,g} w * bilinear_shift(xb[n,s], mu)   (dau_conv_test.py:14-61)."""
    N, S, H, W = xb.shape
    _, S2, G, F = w.shape
    assert S2 == S
    pad = _pad_for(mu1, mu2)
    xp = np.zeros((N, S, H + 2 * pad, W + 2 * pad), dtype=np.float64)
    xp[:, :, pad:pad + H, pad:pad + W] = xb
    y = np.zeros((N, F, H, W), dtype=np.float64)
    for f in range(F):
        acc = y[:, f]
        for s in range(S):
            for g in range(G - num_dau_units_ignore):
                ox, oy, taps = _taps(float(mu1[0, s, g, f]), float(mu2[0, s, g, f]), use_interpolation)
                wv = float(w[0, s, g, f])
                for dy, dx, b in taps:
                    ay, ax = oy + dy + pad, ox + dx + pad
                    acc += (wv * b) * xp[:, s, ay:ay + H, ax:ax + W]
    return y


def edge_mask_flags(height, width):
    """(disable_last_row, disable_last_column) of ignore_edge_gradients
    (dau_conv_test.py:110-136, base_dau_conv_layer.cpp:763-771)."""
    def rule(n):
        for p in (64, 32, 16, 8):
            if n >= p:
                return n % p == 0
        return False
    return rule(height), rule(width)


def offset_and_dot(xk, error, mu1, mu2, num_dau_units_ignore=0, ignore_edge_gradients=True,
                   use_interpolation=True):
    """R[s,g,f] = sum_{n,y,x} err[n,f,y,x] * bilinear_shift(xk[n,s], mu)[y,x]
    (dau_conv_test.py:95-175).  ``error`` is not modified (the reference mutates it)."""
    N, S, H, W = xk.shape
    _, _, G, F = mu1.shape
    err = np.array(error, dtype=np.float64, copy=True)
    if ignore_edge_gradients:
        last_row, last_col = edge_mask_flags(H, W)
        if last_col:
            err[:, :, :, W - 1] = 0.0
        if last_row:
            err[:, :, H - 1, :] = 0.0
    pad = _pad_for(mu1, mu2)
    xp = np.zeros((N, S, H + 2 * pad, W + 2 * pad), dtype=np.float64)
    xp[:, :, pad:pad + H, pad:pad + W] = xk
    out = np.zeros(mu1.shape, dtype=np.float64)
    for f in range(F):
        e = err[:, f]
        for s in range(S):
            for g in range(G - num_dau_units_ignore):
                ox, oy, taps = _taps(float(mu1[0, s, g, f]), float(mu2[0, s, g, f]), use_interpolation)
                tot = 0.0
                for dy, dx, b in taps:
                    ay, ax = oy + dy + pad, ox + dx + pad
                    tot += b * float(np.sum(xp[:, s, ay:ay + H, ax:ax + W] * e))
                out[0, s, g, f] = tot
    return out


def forward(x, w, mu1, mu2, sigma, num_dau_units_ignore=0, do_error_backprop=False,
            single_dim_kernel=False, forbid_positive_dim1=False, use_interpolation=True,
            unit_normalization=True, square_unit_normalization=False, kb=None):
    """DAU forward: blur then displaced bilinear gather (dau_conv_test.py:64-93)."""
    filt = get_filters(float(np.ravel(sigma)[0]), kb, unit_normalization, square_unit_normalization,
                       single_dim_kernel, forbid_positive_dim1)
    k = filt[5] if do_error_backprop else filt[0]
    xb = correlate_same(np.asarray(x, dtype=np.float64), k)
    return offset_and_sum(xb, w, mu1, mu2, num_dau_units_ignore, use_interpolation)


def backward(x, error, w, mu1, mu2, sigma, num_dau_units_ignore=0, ignore_edge_gradients=True,
             single_dim_kernel=False, forbid_positive_dim1=False, use_interpolation=True,
             unit_normalization=True, square_unit_normalization=False, kb=None):
    """DAU backward (dau_conv_test.py:222-295).

    Returns (dL/dx, dL/dw, dL/dmu1, dL/dmu2, dL/dsigma); the last three already carry
    the xw factor (base_dau_conv_layer.cu:335-337) but NOT mu_learning_rate_factor.
    """
    x = np.asarray(x, dtype=np.float64)
    error = np.asarray(error, dtype=np.float64)
    sig = float(np.ravel(sigma)[0])
    kw = dict(single_dim_kernel=single_dim_kernel, forbid_positive_dim1=forbid_positive_dim1,
              unit_normalization=unit_normalization, square_unit_normalization=square_unit_normalization,
              kb=kb)
    # dL/dx: blur the error with the mirrored filter, gather with -mu, params read as [F,G,S]
    dx = forward(error, np.swapaxes(w, 1, 3), np.swapaxes(-mu1, 1, 3), np.swapaxes(-mu2, 1, 3),
                 [sig], do_error_backprop=True, use_interpolation=use_interpolation, **kw)
    filt = get_filters(sig, kb, unit_normalization, square_unit_normalization,
                       single_dim_kernel, forbid_positive_dim1)
    grads = []
    for k in range(4):
        xk = correlate_same(x, filt[1 + k])
        grads.append(offset_and_dot(xk, error, mu1, mu2, num_dau_units_ignore,
                                    ignore_edge_gradients, use_interpolation))
    wv = np.asarray(w, dtype=np.float64)
    return dx, grads[0], grads[1] * wv, grads[2] * wv, grads[3] * wv


def output_border(size_in, kernel_size, pad):
    """(border, size_out): size_out = size_in + 2*pad - kernel_size + 1 (base_dau_conv_layer.cpp:20-22,
    stride 1) and the output pixel o sits over input pixel o + border with
    border = size_in/2 - size_out/2 (base_dau_conv_layer.cpp:544-545)."""
    size_out = size_in + 2 * pad - kernel_size + 1
    return size_in // 2 - size_out // 2, size_out


def forward_cropped(x, w, mu1, mu2, sigma, kernel_size, pad, **kw):
    """Forward with pad < (kernel_size-1)/2: the centre [H_out, W_out] of the same-size result
    (the reference's kernels compute exactly those pixels, base_dau_conv_layer.cpp:540-545)."""
    y = forward(x, w, mu1, mu2, sigma, **kw)
    by, ho = output_border(y.shape[2], kernel_size, pad)
    bx, wo = output_border(y.shape[3], kernel_size, pad)
    return y[:, :, by:by + ho, bx:bx + wo]


def backward_cropped(x, error, w, mu1, mu2, sigma, kernel_size, pad, ignore_edge_gradients=True, **kw):
    """Backward for an error of the OUTPUT size: the reference zero-pads it back to the input size
    at offset `border` (caffe_gpu_pad2d, base_dau_conv_layer.cu:258-266; im2col padding
    base_dau_conv_layer.cpp:938-949) and runs the same-size pass; the edge mask of
    ignore_edge_gradients is taken on the OUTPUT grid (base_dau_conv_layer.cpp:763-791) and
    touches the parameter gradients only."""
    x = np.asarray(x, dtype=np.float64)
    error = np.asarray(error, dtype=np.float64)
    N, S, H, W = x.shape
    by, ho = output_border(H, kernel_size, pad)
    bx, wo = output_border(W, kernel_size, pad)
    assert error.shape[2:] == (ho, wo), (error.shape, ho, wo)
    full = np.zeros(error.shape[:2] + (H, W), dtype=np.float64)
    full[:, :, by:by + ho, bx:bx + wo] = error
    dx = backward(x, full, w, mu1, mu2, sigma, ignore_edge_gradients=False, **kw)[0]
    if ignore_edge_gradients:
        last_row, last_col = edge_mask_flags(ho, wo)
        if last_col:
            full[:, :, :, bx + wo - 1] = 0.0
        if last_row:
            full[:, :, by + ho - 1, :] = 0.0
    grads = backward(x, full, w, mu1, mu2, sigma, ignore_edge_gradients=False, **kw)[1:]
    return (dx,) + tuple(grads)
