"""Per-kernel device-time breakdown of one AlexNet-DAU training step on one GPU (torch.profiler), at the
per-rank batch of the 8-GPU strong-scaling run (32) or any other.  usage: --batch 32 [--variant default]"""
import argparse
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "dau-convnet_b200"))
import alexnet_dau_train_step as A  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--variant", default="default")
    ap.add_argument("--rows", type=int, default=40)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    tr = A.Trainer(a.variant, a.batch, dev, graph=False)
    ms, _ = tr.timed(10, 5)
    print("batch %d: %.3f ms / step eager (CUDA events, 10 steps)" % (a.batch, ms))
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            tr.step()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total / 5 / 1e3, e.count // 5) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    total = sum(r[1] for r in rows)
    print("sum of kernel time: %.3f ms / step over %d launches" % (total, sum(r[2] for r in rows)))
    for k, t, n in rows[:a.rows]:
        print("%8.3f ms  x%-3d %s" % (t, n, k[:120]))
    tr2 = A.Trainer(a.variant, a.batch, dev, graph=True)
    ms2, _ = tr2.timed(10, 5)
    print("batch %d: %.3f ms / step as one CUDA graph" % (a.batch, ms2))


if __name__ == "__main__":
    main()
