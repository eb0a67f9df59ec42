"""Times the REFERENCE's own CUDA kernels (oracle/_ref/libdau_ref_cuda.so, built for sm_100a by
oracle/Makefile.refcuda) next to this repo's path: same GPU, same tensors (bench.py's configs and
input generator).  Prints one JSON line per config; a comparison tool, not bench.py's reference arm
(that one is the reference's CPU path, as BASELINE.json asks).

usage: python tests/ref_cuda_time.py [--config c2 ...] [--images N]   (default: c2)"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dau-convnet_b200"), os.path.dirname(os.path.abspath(__file__))):
    sys.path.insert(0, p)
import bench  # noqa: E402
import dau_conv  # noqa: E402
import ref_cuda  # noqa: E402


def one(cfg, n_images):
    N, S, F, units, H, W, ks, m = bench.CONFIGS[cfg]
    N = n_images or N
    G = units[0] * units[1]
    dev = torch.device("cuda:0")
    x, e, w, mu1, mu2 = [torch.from_numpy(a).to(dev) for a in bench.make_inputs_numpy(cfg, N, 1234)]
    sigma = torch.full((1, S, G, F), 0.5, device=dev)
    out = {"config": bench.workload_name(cfg) if hasattr(bench, "workload_name") else cfg, "images": N}
    try:
        r = ref_cuda.run(x, e, w, mu1, mu2, sigma, units, kernel_size=ks, ignore_edge_gradients=True, iters=6)
    except RuntimeError as ex:
        out["reference_cuda_sm100a"] = {"error": str(ex)}
        return out
    info = r[6]
    attrs = dict(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks, pad=(ks - 1) // 2,
                 unit_testing=True, mu_learning_rate_factor=1.0)   # edge mask on both sides: see tests/test_reference_cuda.py
    for _ in range(3):
        y = dau_conv.dau_conv_op(x, w, mu1, mu2, sigma, **attrs)
        g = dau_conv.dau_conv_grad_op(e, x, w, mu1, mu2, sigma, **attrs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = dau_conv.dau_conv_op(x, w, mu1, mu2, sigma, **attrs)
        g = dau_conv.dau_conv_grad_op(e, x, w, mu1, mu2, sigma, **attrs)
    e1.record()
    torch.cuda.synchronize()
    ours_ms = e0.elapsed_time(e1) / 5
    ref_ms = info["fwd_ms"] + info["bwd_ms"]

    def rel(a, b):
        return float((a - b).abs().max() / b.abs().max())
    out.update({
        "reference_cuda_sm100a": {"fwd_ms": info["fwd_ms"], "bwd_ms": info["bwd_ms"], "step_ms": ref_ms,
                                  "images_per_s": N / ref_ms * 1e3, "workspace_bytes": info["workspace_bytes"],
                                  "what": "BaseDAUConvLayer<float>::Forward_gpu + Backward_gpu, layer object built once "
                                          "outside the timed region, 5 timed passes, CUDA events"},
        "this_repo": {"step_ms": ours_ms, "images_per_s": N / ours_ms * 1e3,
                      "what": "dau_conv_op + dau_conv_grad_op through the Python mirror (outputs allocated per call), "
                              "5 timed passes, CUDA events"},
        "speedup": ref_ms / ours_ms,
        "max_abs_diff_over_max_abs_reference_interior": {
            "top": rel(y[..., :-1, :-1], r[0][..., :-1, :-1]),
            "bottom_error": rel(g[0][..., :-1, :-1], r[1][..., :-1, :-1]),
            "dw": rel(g[1], r[2]), "dmu1": rel(g[2], r[3]), "dmu2": rel(g[3], r[4]), "dsigma": rel(g[4], r[5])}})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", nargs="+", default=["c2"], choices=sorted(bench.CONFIGS))
    ap.add_argument("--images", type=int, default=0)
    a = ap.parse_args()
    for cfg in a.config:
        print(json.dumps(one(cfg, a.images)), flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
