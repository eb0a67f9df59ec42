"""Per-kernel device-time breakdown of one AlexNet-DAU training step on one GPU (torch.profiler),
at the per-rank batch of the 8-GPU strong-scaling run (32) or any other.  usage: --batch 32 [--variant default]"""
import argparse
import os
import sys

import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "dau-convnet_b200"))
from alexnet_dau_train_step import AlexNetDAU  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--variant", default="default")
    ap.add_argument("--rows", type=int, default=25)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    model = AlexNetDAU((1, 2) if a.variant == "default" else (2, 2)).to(dev)
    x = torch.rand((a.batch, 3, 227, 227), device=dev)
    target = torch.randint(0, 1000, (a.batch,), device=dev)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)

    def step():
        opt.zero_grad(set_to_none=False)
        nn.functional.cross_entropy(model(x), target).backward()
        opt.step()
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        step()
    e1.record()
    torch.cuda.synchronize()
    print("batch %d: %.3f ms / step (CUDA events, 10 steps)" % (a.batch, e0.elapsed_time(e1) / 10))
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            step()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total / 5 / 1e3, e.count // 5) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    total = sum(r[1] for r in rows)
    print("sum of kernel time: %.3f ms / step over %d launches" % (total, sum(r[2] for r in rows)))
    for k, ms, n in rows[:a.rows]:
        print("%8.3f ms  x%-3d %s" % (ms, n, k[:110]))


if __name__ == "__main__":
    main()
