"""Every BASELINE.json configuration at its REAL channel counts / map sizes / kernel size against the
pinned C oracle (oracle/dau_cpu.c), all six tensors -- the kernels' wave-split paths (`split_k` tails of
the gathers, `n_splits` partials of the gradient grid, mosaic packing) depend on the shape and the
batch, which small-shape tests do not reach.

 * 32-image batches of C4 and the AlexNet-DAU layers L2..L5 (seconds of CPU each);
 * the batch the metric is quoted on for L2..L5 (N = 256) and C2 (N = 128): the launch plans of the
   bench configurations themselves;
 * C4 at N = 256: y and dL/dx of sampled images against the oracle (per-image work), parameter
   gradients against the sum of eight 32-image passes of the SAME kernels (whose 32-image plan the first
   case holds against the oracle).

The sum over images that the reference does inside one kernel (dau_conv_backward_core.hpp:1257-1262) is
split here over CTAs (`n_splits`) and, in the multi-GPU path, over ranks: tests/test_shards.py.
"""
import numpy as np
import pytest

from _tolerance import close
from oracle import dau_cpu

pytestmark = pytest.mark.gpu

# name: (S, F, units, H, W, kernel_size, mu range)      (bench.py CONFIGS without the batch)
LAYERS = {
    "c2": (128, 128, (2, 2), 64, 64, 9, 4.0),
    "c4": (512, 512, (1, 2), 28, 28, 17, 8.0),
    "l2": (96, 256, (1, 2), 27, 27, 9, 3.0),
    "l3": (256, 384, (1, 2), 13, 13, 9, 3.0),
    "l4": (384, 384, (1, 2), 13, 13, 9, 3.0),
    "l5": (384, 256, (1, 2), 13, 13, 9, 3.0),
}
NAMES = ("dx", "dw", "dmu1", "dmu2", "dsigma")


def make_case(cfg, n, seed):
    S, F, units, H, W, ks, m = LAYERS[cfg]
    G = units[0] * units[1]
    rng = np.random.default_rng(seed)
    lim = ks // 2 - 0.01
    x = rng.random((n, S, H, W), dtype=np.float32)
    e = rng.standard_normal((n, F, H, W), dtype=np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    mu1 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    mu2 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    return x, e, w, mu1, mu2


def run_gpu(cfg, x, e, w, mu1, mu2, edge):
    import torch
    import dau_conv
    S, F, units, H, W, ks, m = LAYERS[cfg]
    dev = torch.device("cuda:0")
    sig = np.full(w.shape, 0.5, np.float32)
    t = [torch.from_numpy(a).to(dev) for a in (x, w, mu1, mu2, sig)]
    a = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=ks // 2,
             unit_testing=edge)
    y = dau_conv.dau_conv_op(*t, sigma_value=0.5, **a)
    g = dau_conv.dau_conv_grad_op(torch.from_numpy(e).to(dev), *t, sigma_value=0.5, **a)
    torch.cuda.synchronize()
    out = y.cpu().numpy(), [v.cpu().numpy() for v in g]
    del y, g, t
    torch.cuda.empty_cache()
    return out


def oracle(cfg, x, e, w, mu1, mu2, edge):
    import dau_conv
    H, W = x.shape[2:]
    lib = dau_conv._lib.lib
    mr, mc = bool(edge and lib.dau_edge_mask_rule(H)), bool(edge and lib.dau_edge_mask_rule(W))
    y = dau_cpu.forward(x, w, mu1, mu2, 0.5)
    g = dau_cpu.backward(x, e, w, mu1, mu2, 0.5, mask_last_row=mr, mask_last_col=mc)
    return y, g


CASES = [("c4", 32, False), ("l2", 32, True), ("l3", 32, False), ("l4", 32, True), ("l5", 32, False),
         ("c2", 16, True),
         ("l2", 256, False), ("l3", 256, True), ("l4", 256, False), ("l5", 256, True), ("c2", 128, False)]


@pytest.mark.parametrize("cfg,n,edge", CASES, ids=["%s_n%d_%s" % (c, n, "edgemask" if e else "exact") for c, n, e in CASES])
def test_full_channel_counts_against_c_oracle(cfg, n, edge):
    x, e, w, mu1, mu2 = make_case(cfg, n, 4000 + n + len(cfg))
    y, g = run_gpu(cfg, x, e, w, mu1, mu2, edge)
    y_ref, g_ref = oracle(cfg, x, e, w, mu1, mu2, edge)
    close(y, y_ref, "y %s" % cfg)
    for name, got, ref in zip(NAMES, g, g_ref):
        close(got, ref, "%s %s" % (name, cfg), scale_atol=name != "dx")


def test_c4_at_its_full_batch():
    cfg, n = "c4", 256
    x, e, w, mu1, mu2 = make_case(cfg, n, 4242)
    y, g = run_gpu(cfg, x, e, w, mu1, mu2, False)
    pick = [0, 1, 126, 127, 255]                      # first pair, a middle pair, the last image
    y_ref, g_ref = oracle(cfg, x[pick], e[pick], w, mu1, mu2, False)
    close(y[pick], y_ref, "y c4@256")
    close(g[0][pick], g_ref[0], "dx c4@256")
    # parameter gradients: the 256-image launch plan against eight 32-image passes of the same kernels
    acc = [np.zeros(w.shape, np.float64) for _ in range(4)]
    for lo in range(0, n, 32):
        _, gp = run_gpu(cfg, x[lo:lo + 32], e[lo:lo + 32], w, mu1, mu2, False)
        for a, v in zip(acc, gp[1:]):
            a += v
    for name, got, ref in zip(NAMES[1:], g[1:], acc):
        close(got, ref, "%s c4@256" % name, scale_atol=True)
