"""The batch-sharded (multi-GPU) path on hardware: "parity for multi-GPU = same
tolerance vs the single-GPU result".  The reference sums the parameter gradients over the images inside
one kernel (include/dau_conv/dau_conv_impl/dau_conv_backward_core.hpp:1257-1262); here the sum is split
over ranks and finished by ONE all-reduce of the flat [4,S,G,F] buffer (dau_conv/distributed.py).

 * one GPU, two handles, images [0,N/2) and [N/2,N): y / dL/dx concatenated == the single pass, the
   summed [4,S,G,F] == the single pass (same tolerance as against the oracle);
 * the same through torch.multiprocessing + NCCL on 2 ranks with dau_conv.distributed.GradBucket
   (skipped on a one-GPU box);
 * ONE process driving handles on cuda:0 and cuda:1 (the kernels' shared-memory opt-in is a per-device
   attribute; the caller's current device must be left alone).
"""
import os
import socket

import numpy as np
import pytest

from _tolerance import close

pytestmark = pytest.mark.gpu

NAMES = ("dx", "dw", "dmu1", "dmu2", "dsigma")
# (N, S, F, units, H, W, kernel_size, mu range)
SHARD_SHAPES = [
    (6, 5, 6, (1, 2), 20, 24, 9, 3.9),          # S < 8: no gather tail split possible -> bit-identical y / dL/dx
    (32, 128, 128, (2, 2), 64, 64, 9, 4.0),     # a quarter of C2
    (64, 384, 384, (1, 2), 13, 13, 9, 3.0),     # AlexNet-DAU L4 (mosaic of 8 images, 8 images per mosaic pair)
    (7, 16, 32, (1, 2), 27, 27, 9, 3.0),        # odd batch: the shards are 4 + 3 images
]


def _case(shape, seed):
    N, S, F, units, H, W, ks, m = shape
    G = units[0] * units[1]
    rng = np.random.default_rng(seed)
    lim = ks // 2 - 0.01
    x = rng.random((N, S, H, W), dtype=np.float32)
    e = rng.standard_normal((N, F, H, W), dtype=np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    mu1 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    mu2 = np.clip(rng.uniform(-m, m, (1, S, G, F)), -lim, lim).astype(np.float32)
    return x, e, w, mu1, mu2


def _handle(dev, shape):
    import dau_conv
    from dau_conv import _lib
    N, S, F, units, H, W, ks, m = shape
    st = _lib.default_settings(num_output=F, number_units_x=units[0], number_units_y=units[1], kernel_size=ks,
                               pad=ks // 2)
    h = dau_conv.DAULayerHandle(st, dev)
    h.set_sigma_hint(0.5)
    return h


def _run(h, dev, x, e, w, mu1, mu2, bucket=None):
    import torch
    t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x, w, mu1, mu2, np.full(w.shape, 0.5, np.float32))]
    y = h.forward(*t)
    out = h.backward(t[0], torch.from_numpy(np.ascontiguousarray(e)).to(dev), *t[1:],
                     grads=bucket.views if bucket is not None else None)
    return y, out


@pytest.mark.parametrize("shape", SHARD_SHAPES, ids=["x".join(str(v) for v in s[:3]) + "_%dx%d" % s[4:6] for s in SHARD_SHAPES])
def test_two_shards_on_one_gpu_equal_the_single_pass(shape):
    import torch
    from dau_conv.distributed import GradBucket, shard_range
    dev = torch.device("cuda:0")
    x, e, w, mu1, mu2 = _case(shape, 31)
    N = shape[0]
    y_full, g_full = _run(_handle(dev, shape), dev, x, e, w, mu1, mu2)
    bucket = GradBucket(w.shape, dev)                   # both shards ACCUMULATE into the same flat buffer
    ys, dxs = [], []
    streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
    for r in range(2):
        lo, hi = shard_range(N, r, 2)
        with torch.cuda.stream(streams[r]):
            if r == 1:
                streams[1].wait_stream(streams[0])      # the accumulation into the bucket is stream-ordered
            y, out = _run(_handle(dev, shape), dev, x[lo:hi], e[lo:hi], w, mu1, mu2, bucket)
        ys.append(y); dxs.append(out[0])
    torch.cuda.synchronize()
    y_cat = torch.cat(ys).cpu().numpy(); dx_cat = torch.cat(dxs).cpu().numpy()
    y_ref, dx_ref = y_full.cpu().numpy(), g_full[0].cpu().numpy()
    if shape[1] < 8:
        assert np.array_equal(y_cat, y_ref) and np.array_equal(dx_cat, dx_ref)       # per-image work, same plan
    close(y_cat, y_ref, "y shards")
    close(dx_cat, dx_ref, "dx shards")
    for k, name in enumerate(NAMES[1:]):
        close(bucket.flat[k].cpu().numpy(), g_full[1 + k].cpu().numpy(), "%s shards" % name, scale_atol=True)


def test_one_process_two_devices():
    """Handles on cuda:0 and cuda:1 driven from ONE process (a Caffe-style host): same results on both,
    and the caller's current device is where it was."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    shape = (4, 16, 32, (1, 2), 40, 64, 9, 3.9)
    x, e, w, mu1, mu2 = _case(shape, 8)
    torch.cuda.set_device(0)
    res = []
    for d in (1, 0, 1):                                   # cuda:1 FIRST: its kernel attributes were never set
        dev = torch.device("cuda", d)
        y, g = _run(_handle(dev, shape), dev, x, e, w, mu1, mu2)
        assert torch.cuda.current_device() == 0          # every entry point restores the caller's device
        torch.cuda.synchronize(dev)
        res.append([y.cpu().numpy()] + [v.cpu().numpy() for v in g])
    for other in res[1:]:
        for a, b in zip(res[0], other):
            assert np.array_equal(a, b)
    from oracle import dau_oracle as O
    close(res[0][0], O.forward(x, w, mu1, mu2, [0.5]), "y two devices")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _nccl_worker(rank, world, port, shape, out_dir):
    import sys
    import torch
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "dau-convnet_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from dau_conv.distributed import GradBucket, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    x, e, w, mu1, mu2 = _case(shape, 31)
    lo, hi = shard_range(shape[0], rank, world)
    bucket = GradBucket(w.shape, dev)
    y, out = _run(_handle(dev, shape), dev, x[lo:hi], e[lo:hi], w, mu1, mu2, bucket)
    bucket.all_reduce()                                    # the ONE exchange step of the path
    torch.cuda.synchronize()
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), y=y.cpu().numpy(), dx=out[0].cpu().numpy(),
             flat=bucket.flat.cpu().numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("shape", SHARD_SHAPES[1:3], ids=["c2_quarter", "alexnet_l4"])
def test_two_ranks_nccl_equal_the_single_gpu_result(shape, tmp_path):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    mp.spawn(_nccl_worker, args=(world, _free_port(), shape, str(tmp_path)), nprocs=world, join=True)
    dev = torch.device("cuda:0")
    x, e, w, mu1, mu2 = _case(shape, 31)
    y_full, g_full = _run(_handle(dev, shape), dev, x, e, w, mu1, mu2)
    torch.cuda.synchronize()
    r = [np.load(tmp_path / ("r%d.npz" % k)) for k in range(world)]
    assert np.array_equal(r[0]["flat"], r[1]["flat"])     # every rank holds the same reduced buffer
    close(np.concatenate([r[0]["y"], r[1]["y"]]), y_full.cpu().numpy(), "y 2 ranks")
    close(np.concatenate([r[0]["dx"], r[1]["dx"]]), g_full[0].cpu().numpy(), "dx 2 ranks")
    for k, name in enumerate(NAMES[1:]):
        close(r[0]["flat"][k], g_full[1 + k].cpu().numpy(), "%s 2 ranks" % name, scale_atol=True)
