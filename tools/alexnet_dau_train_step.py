#!/usr/bin/env python3
"""AlexNet-DAU training step (BASELINE.json configs[2] / configs[4]).

  python tools/alexnet_dau_train_step.py [--variant default|large] [--global-batch 256] [--steps K]
  torchrun --nproc-per-node N ... tools/alexnet_dau_train_step.py --global-batch 256      (batch-sharded)

The network is the AlexNet of the reference's README (README.md:33-35; the prototxt gists are not
in the tree, so the layer shapes are the derivation of SURVEY.md section 8(d), assumption A1):
conv1 11x11/4 (torch/cuDNN) -> ReLU -> LRN -> pool -> DAU L2 (96->256 @27x27) -> ReLU -> LRN ->
pool -> DAU L3 (256->384 @13x13) -> DAU L4 (384->384) -> DAU L5 (384->256) -> pool -> FC x3.
`default` has 2 units per (s,f) pair (0.74 M units), `large` 4 (1.47 M).  The four DAU layers run
on libdau_conv_b200.so through dau_conv.DAUConv2d; everything else is stock torch.

One step = forward, loss, backward, all-reduce of ALL gradients, SGD(momentum) update, on synthetic
data.  How the step is laid out (`Trainer`):
  * the parameters form two buckets -- "head" (the three FC layers, 235 MB, 95 % of the bytes) and "body"
    (conv1 + the DAU layers); a bucket's gradients travel in ONE coalesced NCCL all-reduce; gradients are
    dropped (not zero-filled) before every step so autograd assigns instead of accumulating, and the
    optimizer is one fused multi-tensor SGD kernel: no extra pass over the 235 MB of FC gradients;
  * the loss is sum_i CE_i / global_batch, so per-rank gradients ADD to the single-GPU gradient: the
    all-reduce is a plain SUM, uneven shards are exact, nothing is divided afterwards;
  * backward runs in two pieces: through the FC head first, whose bucket is all-reduced (NCCL, async,
    its own stream) WHILE the backward pass of the convolutional body runs; the small body bucket
    follows.  The DAU layers' [4,S,G,F] gradients are all that leaves a GPU for the path itself;
  * the FC head's SGD update (a pass over 1.2 GB that does not shrink with the per-GPU batch) runs on a side stream
    as soon as the head's all-reduce has landed, under the body's backward pass (2 GPUs x 32 images: 8.61 -> 8.52 ms);
  * with graph=True the whole step (both collectives included) is captured once in a CUDA graph and
    replayed: at 32 images per GPU the step is ~150 launches of a few tens of microseconds each.
Prints one JSON line (whole-job images/s, max over ranks, CUDA events).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (os.path.join(ROOT, "dau-convnet_b200"),):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch
import torch.distributed as dist
from torch import nn

import dau_conv

VARIANTS = {"default": (1, 2), "large": (2, 2)}


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


class _Bucket(object):
    """A group of parameters whose gradients travel in ONE (coalesced) all-reduce.  The parameters live in one flat
    buffer (views); gradients are whatever tensors autograd ASSIGNS during backward (they are dropped, not
    zero-filled, before every step, so there is no fill pass and no accumulate pass over the 235 MB of FC weights)."""

    def __init__(self, params, device):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.p = torch.empty(n, device=device)
        o = 0
        for p in self.params:
            k = p.numel()
            self.p[o:o + k].copy_(p.data.reshape(-1))
            p.data = self.p[o:o + k].view_as(p)
            o += k

    def drop_grads(self):
        for p in self.params:
            p.grad = None

    def grads(self):
        return [p.grad for p in self.params]

    @property
    def g(self):
        """Flat copy of the gradients (tests)."""
        return torch.cat([p.grad.reshape(-1) for p in self.params])


class Trainer(object):
    def __init__(self, variant="default", global_batch=256, device=None, rank=0, world=1, lr=1e-3, momentum=0.9,
                 overlap=True, graph=False, seed=1234, num_classes=1000, data_seed=99, overlap_sgd=True):
        self.rank, self.world, self.global_batch = rank, world, int(global_batch)
        self.dev = torch.device(device if device is not None else "cuda")
        self.lr, self.momentum, self.overlap = lr, momentum, bool(overlap)
        torch.manual_seed(seed)                                   # identical parameters on every rank
        self.model = AlexNetDAU(VARIANTS[variant], num_classes).to(self.dev)
        self.head = _Bucket(self.model.classifier.parameters(), self.dev)
        self.body = _Bucket(self.model.features.parameters(), self.dev)
        # one fused multi-tensor kernel per step: momentum and update in a single pass over p, g and m
        self.opt = torch.optim.SGD(self.head.params + self.body.params, lr=lr, momentum=momentum, fused=True)
        # overlap_sgd: the FC head's update (62 M parameters, a pass over 1.2 GB that does not shrink with the batch)
        # runs on a side stream as soon as the head's all-reduce has landed, under the body's backward pass
        self.overlap_sgd = bool(overlap_sgd) and bool(overlap)
        if self.overlap_sgd:
            self.opt_head = torch.optim.SGD(self.head.params, lr=lr, momentum=momentum, fused=True)
            self.opt_body = torch.optim.SGD(self.body.params, lr=lr, momentum=momentum, fused=True)
            self._side = torch.cuda.Stream(self.dev)
        # images [lo, hi) of the global batch (dau_conv.distributed.shard_range: uneven shards allowed)
        from dau_conv.distributed import shard_range
        self.lo, self.hi = shard_range(self.global_batch, rank, world)
        gen = torch.Generator(device="cpu"); gen.manual_seed(data_seed)
        # synthetic data: the SAME global batch whatever the rank count (strong scaling, comparable runs)
        self.x = None; self.target = None
        self._gen = gen
        self.n_dau_units = sum(m.weights.numel() for m in self.model.modules() if isinstance(m, dau_conv.DAUConv2d))
        self.num_classes = num_classes
        self._graph = None
        self._want_graph = bool(graph)
        self.graph_error = None
        self._loss = None

    # -- data -------------------------------------------------------------------------------------------
    def set_batch(self, x_global=None, target_global=None):
        """This rank's shard of a global batch (synthetic uniform images / random labels by default)."""
        if x_global is None:
            x_global = torch.rand((self.global_batch, 3, 227, 227), generator=self._gen)
            target_global = torch.randint(0, self.num_classes, (self.global_batch,), generator=self._gen)
        x = x_global[self.lo:self.hi].to(self.dev); t = target_global[self.lo:self.hi].to(self.dev)
        if self.x is None or self.x.shape != x.shape:
            self.x, self.target = x.contiguous(), t.contiguous()
            self._graph = None
        else:                                                     # keep the addresses a captured graph reads
            self.x.copy_(x); self.target.copy_(t)

    # -- one step -----------------------------------------------------------------------------------------
    def _reduce(self, bucket):
        """SUM all-reduce of the bucket's gradient tensors as ONE coalesced NCCL launch, asynchronous."""
        if self.world > 1 and dist.is_available() and dist.is_initialized():
            with dist._coalescing_manager(device=self.dev, async_ops=True) as cm:
                for g in bucket.grads():
                    dist.all_reduce(g, op=dist.ReduceOp.SUM)
            return cm
        return None                                               # one rank (or shards simulated on one GPU by a test)

    def compute_gradients(self, step_head=False):
        """Forward + backward of this rank's shard; leaves the SUMMED (all-reduced) gradients of the global
        batch in the parameters' .grad and returns the shard's part of the loss.  step_head: also apply the FC
        head's SGD update on the side stream while the body's backward pass runs (overlap_sgd)."""
        m = self.model
        self.head.drop_grads(); self.body.drop_grads()
        f = m.features(self.x)
        fd = f.detach().requires_grad_(True)
        loss = nn.functional.cross_entropy(m.classifier(fd), self.target, reduction="sum") / self.global_batch
        loss.backward()                                           # FC head: its bucket is complete now
        w_head = self._reduce(self.head) if self.overlap else None
        if step_head:
            main = torch.cuda.current_stream(self.dev)
            self._side.wait_stream(main)                          # the head's gradients were produced on `main`
            with torch.cuda.stream(self._side):
                if w_head is not None:
                    w_head.wait()                                 # the side stream waits for the all-reduce, not `main`
                self.opt_head.step()
            w_head = None
        f.backward(fd.grad)                                       # convolutional body, under the head's all-reduce
        if not self.overlap:
            w_head = self._reduce(self.head)
        w_body = self._reduce(self.body)
        for w in (w_head, w_body):
            if w is not None:
                w.wait()                                          # stream-ordered: the compute stream waits, not the host
        if step_head:
            torch.cuda.current_stream(self.dev).wait_stream(self._side)
        return loss.detach()

    def _eager_step(self):
        if self.overlap_sgd:
            loss = self.compute_gradients(step_head=True)
            self.opt_body.step()
            return loss
        loss = self.compute_gradients()
        self.opt.step()
        return loss

    def capture(self, warmup=3):
        """Captures the step in a CUDA graph (collectives included); falls back to eager on failure."""
        side = torch.cuda.Stream(self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._eager_step()
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._loss = self._eager_step()
            self._graph = g
        except Exception as ex:                                   # pragma: no cover  (reported by the caller)
            self._graph = None
            self.graph_error = "%s: %s" % (type(ex).__name__, str(ex)[:300])
            torch.cuda.synchronize(self.dev)

    def step(self):
        if self.x is None:
            self.set_batch()
        if self._want_graph and self._graph is None and self.graph_error is None:
            self.capture()
        if self._graph is not None:
            self._graph.replay()
            return self._loss
        return self._eager_step()

    def global_loss(self, loss):
        """Sum of the ranks' loss parts = the mean loss of the global batch."""
        t = loss.detach().clone().reshape(1)
        if self.world > 1 and dist.is_initialized():
            dist.all_reduce(t)
        return float(t.item())

    # -- timing ---------------------------------------------------------------------------------------------
    def sync(self):
        torch.cuda.synchronize(self.dev)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        torch.cuda.synchronize(self.dev)

    def timed(self, steps, warmup):
        """(ms per step as the max over ranks, last global loss): CUDA events around `steps` steps."""
        loss = None
        for _ in range(max(warmup, 1)):
            loss = self.step()
        self.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = self.step()
        e1.record()
        self.sync()
        t = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps, self.global_loss(loss)

    def describe(self, variant):
        return ("AlexNet-DAU-%s, %d DAU units, global batch %d at 227x227, batch-sharded dp%d (%d images on rank 0), "
                "SUM all-reduce of two flat gradient buckets (FC head %s the backward pass of the body) + SGD momentum, %s"
                % (variant, self.n_dau_units, self.global_batch, self.world, self.hi - self.lo,
                   "overlapped with" if self.overlap else "after",
                   "one CUDA graph per rank" if self._graph is not None else "eager launches"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="default", choices=sorted(VARIANTS))
    ap.add_argument("--global-batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--overlap", type=int, default=1,
                    help="1: the FC head's gradient bucket is all-reduced while the body's backward pass runs; "
                         "0: both buckets after backward")
    ap.add_argument("--overlap-sgd", type=int, default=1,
                    help="1: the FC head's SGD update runs on a side stream under the body's backward pass")
    ap.add_argument("--graph", type=int, default=-1,
                    help="1: replay the step as one CUDA graph per rank; 0: eager launches; -1 (default): graph on one "
                         "rank, eager on several (capturing the NCCL collectives is opt-in: tests/test_alexnet_step.py)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    graph = (world == 1) if args.graph < 0 else bool(args.graph)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if graph:       # torch's CUDA-graph notes: no NCCL async error handling (its watchdog polls events) around a capture
            os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
        dist.init_process_group("nccl", device_id=dev)
    tr = Trainer(args.variant, args.global_batch, dev, rank, world, overlap=bool(args.overlap), graph=graph,
                 overlap_sgd=bool(args.overlap_sgd))
    ms, loss = tr.timed(args.steps, args.warmup)
    if rank == 0:
        print(json.dumps({
            "metric": "alexnet_dau_train_step_images_per_sec", "value": args.global_batch / (ms * 1e-3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "scaling": "strong", "dtype": "f32", "data": "synthetic", "loss": loss,
            "graph": tr._graph is not None, "graph_error": tr.graph_error,
            "config": {"workload": tr.describe(args.variant)}}))
    teardown(tr, world)


def teardown(tr, world):
    """Leaves the process group.  A captured graph holds NCCL work: it is released first, and a watchdog ends the
    process if the communicator teardown still blocks (seen with torch 2.11 / NCCL 2.28 after graph replays)."""
    import gc
    import threading
    sys.stdout.flush()
    if world <= 1:
        return
    threading.Timer(30.0, lambda: os._exit(0)).start() if tr._graph is not None else None
    torch.cuda.synchronize(tr.dev)
    dist.barrier()
    tr._graph = None
    tr._loss = None
    gc.collect()
    torch.cuda.synchronize(tr.dev)
    dist.destroy_process_group()
    if tr is not None and tr.graph_error is None and threading.active_count() > 1:
        os._exit(0)                                               # do not wait for the watchdog timer


if __name__ == "__main__":
    main()
