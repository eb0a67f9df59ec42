#!/usr/bin/env python3
"""AlexNet-DAU training step (BASELINE.json configs[2] / configs[4]).

  python tools/alexnet_dau_train_step.py [--variant default|large] [--global-batch 256] [--steps K]
  torchrun --nproc-per-node N ... tools/alexnet_dau_train_step.py --global-batch 256      (batch-sharded)

The network is the AlexNet of the reference's README (README.md:33-35; the prototxt gists are not
in the tree, so the layer shapes are the derivation of SURVEY.md section 8(d), assumption A1):
conv1 11x11/4 (torch/cuDNN) -> ReLU -> LRN -> pool -> DAU L2 (96->256 @27x27) -> ReLU -> LRN ->
pool -> DAU L3 (256->384 @13x13) -> DAU L4 (384->384) -> DAU L5 (384->256) -> pool -> FC x3.
`default` has 2 units per (s,f) pair (0.74 M units), `large` 4 (1.47 M).  The four DAU layers run
on libdau_conv_b200.so through dau_conv.DAUConv2d; everything else is stock torch.  One step =
forward, loss, backward, ONE bucketed NCCL all-reduce of all gradients, SGD update, on synthetic
data.  Prints one JSON line (whole-job images/s, max over ranks, CUDA events).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dau-convnet_b200"))

import torch
import torch.distributed as dist
from torch import nn

import dau_conv


class AlexNetDAU(nn.Module):
    def __init__(self, units, num_classes=1000):
        super().__init__()
        dau = lambda cin, cout: dau_conv.DAUConv2d(cout, units, 9, activation=torch.relu, in_channels=cin)
        self.features = nn.Sequential(
            nn.Conv2d(3, 96, 11, stride=4), nn.ReLU(inplace=True), nn.LocalResponseNorm(5, 1e-4, 0.75),
            nn.MaxPool2d(3, 2),
            dau(96, 256), nn.LocalResponseNorm(5, 1e-4, 0.75), nn.MaxPool2d(3, 2),
            dau(256, 384), dau(384, 384), dau(384, 256), nn.MaxPool2d(3, 2))
        self.classifier = nn.Sequential(
            nn.Flatten(), nn.Linear(256 * 6 * 6, 4096), nn.ReLU(inplace=True),
            nn.Linear(4096, 4096), nn.ReLU(inplace=True), nn.Linear(4096, num_classes))

    def forward(self, x):
        return self.classifier(self.features(x))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="default", choices=["default", "large"])
    ap.add_argument("--global-batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--overlap", type=int, default=1,
                    help="1: gradient buckets are all-reduced while backward still runs (torch DDP); "
                         "0: one all-reduce of all gradients after backward")
    ap.add_argument("--graph", type=int, default=0,
                    help="1: capture forward + backward + SGD of one rank in a CUDA graph (single GPU only)")
    ap.add_argument("--lean", type=int, default=1,
                    help="1: gradients are dropped, not zero-filled, between steps (autograd then assigns instead of "
                         "accumulating: no fill and no add per parameter) and SGD runs as one fused kernel per "
                         "parameter group; 0: zero-filled gradients and the foreach SGD")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1234)                                    # identical parameters on every rank
    units = (1, 2) if args.variant == "default" else (2, 2)
    model = AlexNetDAU(units).to(dev)
    net = model
    if world > 1 and args.overlap:
        net = nn.parallel.DistributedDataParallel(model, device_ids=[local], bucket_cap_mb=32,
                                                  gradient_as_bucket_view=True)
    n_local = args.global_batch // world
    gen = torch.Generator(device=dev); gen.manual_seed(99 + rank)
    x = torch.rand((n_local, 3, 227, 227), device=dev, generator=gen)
    target = torch.randint(0, 1000, (n_local,), device=dev, generator=gen)
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.SGD(params, lr=1e-3, momentum=0.9, fused=bool(args.lean))
    n_dau_units = sum(m.weights.numel() for m in model.modules() if isinstance(m, dau_conv.DAUConv2d))

    def step():
        opt.zero_grad(set_to_none=bool(args.lean))
        loss = nn.functional.cross_entropy(net(x), target)
        loss.backward()
        if world > 1 and not args.overlap:                     # one bucket: all gradients, summed, / world
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat /= world
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p)); o += p.numel()
        opt.step()
        return loss

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        step()
    sync()
    if args.graph and world == 1:
        # everything in the step is stream-ordered and allocation-free after warm-up (the DAU op reads
        # no device value on the host once sigma is known), so the whole step replays as one graph
        graph = torch.cuda.CUDAGraph()
        eager_step = step
        with torch.cuda.graph(graph):
            static_loss = eager_step()

        def step():
            graph.replay()
            return static_loss
        step()
        sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    sync()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    if rank == 0:
        print(json.dumps({
            "metric": "alexnet_dau_train_step_images_per_sec", "value": args.global_batch / (ms * 1e-3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "scaling": "strong", "dtype": "f32", "data": "synthetic", "loss": float(loss),
            "config": {"workload": "AlexNet-DAU-%s, %d DAU units, global batch %d at 227x227, batch-sharded dp%d, "
                                   "%s + SGD" % (
                                       args.variant, n_dau_units, args.global_batch, world,
                                       "bucketed NCCL all-reduce overlapped with backward" if args.overlap
                                       else "one NCCL all-reduce of all gradients")}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
