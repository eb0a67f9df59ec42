"""Times the REFERENCE's own CUDA kernels (oracle/_ref/libdau_ref_cuda.so, built for sm_100a by
oracle/Makefile.refcuda) on config C2 next to this repo's path, same GPU, same tensors.
Prints one JSON line; a comparison tool, not bench.py's reference arm (that one is the CPU path)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dau-convnet_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import dau_conv  # noqa: E402
import ref_cuda  # noqa: E402


def main():
    N, S, F, units, H, W = 128, 128, 128, (2, 2), 64, 64
    if len(sys.argv) > 1:
        N = int(sys.argv[1])
    G = units[0] * units[1]
    rng = np.random.default_rng(1235)
    dev = torch.device("cuda:0")
    x = torch.from_numpy(rng.random((N, S, H, W), dtype=np.float32)).to(dev)
    err = torch.from_numpy(rng.standard_normal((N, F, H, W), dtype=np.float32)).to(dev)
    w = torch.from_numpy((0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)).to(dev)
    mu1 = torch.from_numpy(rng.uniform(-3.99, 3.99, (1, S, G, F)).astype(np.float32)).to(dev)
    mu2 = torch.from_numpy(rng.uniform(-3.99, 3.99, (1, S, G, F)).astype(np.float32)).to(dev)
    sigma = torch.full((1, S, G, F), 0.5, device=dev)
    r = ref_cuda.run(x, err, w, mu1, mu2, sigma, units, kernel_size=9, ignore_edge_gradients=False, iters=6)
    info = r[6]
    attrs = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=9, pad=4,
                 unit_testing=False, mu_learning_rate_factor=1.0)
    for _ in range(3):
        y = dau_conv.dau_conv_op(x, w, mu1, mu2, sigma, **attrs)
        g = dau_conv.dau_conv_grad_op(err, x, w, mu1, mu2, sigma, **attrs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = dau_conv.dau_conv_op(x, w, mu1, mu2, sigma, **attrs)
        g = dau_conv.dau_conv_grad_op(err, x, w, mu1, mu2, sigma, **attrs)
    e1.record()
    torch.cuda.synchronize()
    ours_ms = e0.elapsed_time(e1) / 5
    ref_ms = info["fwd_ms"] + info["bwd_ms"]

    def rel(a, b):
        return float((a - b).abs().max() / b.abs().max())
    out = {"config": "C2: N=%d, S=F=128, 64x64, G=4, kernel_size=9" % N,
           "reference_cuda_sm100a": {"fwd_ms": info["fwd_ms"], "bwd_ms": info["bwd_ms"], "step_ms": ref_ms,
                                     "images_per_s": N / ref_ms * 1e3, "workspace_bytes": info["workspace_bytes"]},
           "this_repo": {"step_ms": ours_ms, "images_per_s": N / ours_ms * 1e3,
                         "note": "dau_conv_op + dau_conv_grad_op through the Python mirror (allocates outputs per call)"},
           "speedup": ref_ms / ours_ms,
           "max_rel_diff_interior": {"top": rel(y[..., :-1, :-1], r[0][..., :-1, :-1]),
                                     "bottom_error": rel(g[0][..., :-1, :-1], r[1][..., :-1, :-1]),
                                     "dw": rel(g[1], r[2]), "dmu1": rel(g[2], r[3]), "dmu2": rel(g[3], r[4]),
                                     "dsigma": rel(g[4], r[5])}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
