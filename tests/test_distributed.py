"""World-size-2 gloo test of the batch-sharded path's host logic, on CPU.

The layer math is stood in for by the oracle (this is a test of sharding + the single all-reduce,
not of the CUDA kernels): the sum of per-shard parameter gradients must equal the full-batch
gradient, and dL/dx / y of a shard must equal the corresponding slice of the full-batch result.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import dau_oracle as O


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _case():
    rng = np.random.default_rng(5)
    N, S, F, G, H, W = 5, 2, 3, 2, 9, 10
    x = rng.random((N, S, H, W)); e = rng.standard_normal((N, F, H, W))
    w = 0.1 * rng.standard_normal((1, S, G, F))
    mu1 = rng.uniform(-3, 3, w.shape); mu2 = rng.uniform(-3, 3, w.shape)
    return x, e, w, mu1, mu2


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dau-convnet_b200"))
    from dau_conv.distributed import GradBucket, shard_range
    x, e, w, mu1, mu2 = _case()
    lo, hi = shard_range(x.shape[0], rank, world)
    y = O.forward(x[lo:hi], w, mu1, mu2, [0.5])
    dx, dw, dm1, dm2, ds = O.backward(x[lo:hi], e[lo:hi], w, mu1, mu2, [0.5], ignore_edge_gradients=False)
    bucket = GradBucket(w.shape, "cpu", torch.float64)
    for v, g in zip(bucket.views, (dw, dm1, dm2, ds)):
        v.copy_(torch.from_numpy(g))
    bucket.all_reduce()
    np.save(os.path.join(out, "g%d.npy" % rank), bucket.flat.numpy())
    np.save(os.path.join(out, "y%d.npy" % rank), y); np.save(os.path.join(out, "dx%d.npy" % rank), dx)
    dist.destroy_process_group()


def test_shard_range_covers_batch():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dau-convnet_b200"))
    from dau_conv.distributed import shard_range
    for n in (1, 5, 8, 128):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def test_two_rank_allreduce_equals_full_batch(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    x, e, w, mu1, mu2 = _case()
    full = O.backward(x, e, w, mu1, mu2, [0.5], ignore_edge_gradients=False)
    g0 = np.load(tmp_path / "g0.npy"); g1 = np.load(tmp_path / "g1.npy")
    np.testing.assert_array_equal(g0, g1)                       # every rank holds the reduced sum
    for k in range(4):
        np.testing.assert_allclose(g0[k], full[1 + k], rtol=1e-10, atol=1e-12)
    y = np.concatenate([np.load(tmp_path / "y0.npy"), np.load(tmp_path / "y1.npy")])
    dx = np.concatenate([np.load(tmp_path / "dx0.npy"), np.load(tmp_path / "dx1.npy")])
    np.testing.assert_allclose(y, O.forward(x, w, mu1, mu2, [0.5]), rtol=1e-12)
    np.testing.assert_allclose(dx, full[0], rtol=1e-12)
