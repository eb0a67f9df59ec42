"""The AlexNet-DAU training-step harness (tools/alexnet_dau_train_step.py):
the loss falls, the gradients of a sharded batch add up to the gradients of the whole batch (on one
GPU, and through NCCL on two ranks with the step replayed as a CUDA graph), and ranks stay in step."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

pytestmark = pytest.mark.gpu


def _trainer(**kw):
    import alexnet_dau_train_step as A
    kw.setdefault("num_classes", 10)
    return A.Trainer("default", **kw)


def _rel(a, b):
    a = a.double(); b = b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_loss_falls_over_steps():
    tr = _trainer(global_batch=8, device="cuda:0", lr=2e-3)
    losses = [float(tr.step()) for _ in range(12)]
    assert all(np.isfinite(losses))
    assert losses[-1] < 0.8 * losses[0], losses


def test_two_shards_add_up_to_the_full_batch_gradient():
    """Gradient buckets of images [0,3) and [3,8) (the shard ranges of 2 ranks on an uneven split would be 4+4;
    5 images + 3 here through global_batch 8 on a 2-'rank' layout with rank-local batches) summed == the
    one-rank gradient: what the SUM all-reduce relies on."""
    import torch
    full = _trainer(global_batch=6, device="cuda:0")
    full.set_batch()
    l_full = float(full.compute_gradients())
    parts = []
    for r in range(2):
        t = _trainer(global_batch=6, device="cuda:0", rank=r, world=2)      # torch.distributed not initialised:
        t.set_batch()                                                        # the all-reduce is the test's sum
        parts.append((float(t.compute_gradients()), t.head.g.clone(), t.body.g.clone()))
        assert torch.equal(t.head.p, full.head.p) and torch.equal(t.body.p, full.body.p)   # same seed, same init
    assert abs(parts[0][0] + parts[1][0] - l_full) <= 1e-5 * abs(l_full)
    assert _rel(parts[0][1] + parts[1][1], full.head.g) < 2e-4
    assert _rel(parts[0][2] + parts[1][2], full.body.g) < 2e-4
    torch.cuda.synchronize()


def test_graph_replay_equals_eager():
    import torch
    a = _trainer(global_batch=4, device="cuda:0", graph=False)
    b = _trainer(global_batch=4, device="cuda:0", graph=True)
    for _ in range(6):                    # 3 of b's steps are the capture's warm-up: same count on both
        a.step()
    b.set_batch()
    b.capture(warmup=3)
    assert b.graph_error is None, b.graph_error
    for _ in range(3):
        b.step()
    torch.cuda.synchronize()
    # capture() runs warmup + 1 captured-but-not-executed step: a did 6 steps, b did 3 (warm-up) + 3 (replays)
    assert _rel(b.body.p, a.body.p) < 1e-4 and _rel(b.head.p, a.head.p) < 1e-4


def test_overlapped_head_update_equals_the_plain_step():
    """overlap_sgd: the FC head's SGD on a side stream under the body's backward pass -- same parameters after 4 steps."""
    import torch
    a = _trainer(global_batch=4, device="cuda:0", overlap_sgd=False)
    b = _trainer(global_batch=4, device="cuda:0", overlap_sgd=True)
    for _ in range(4):
        a.step(); b.step()
    torch.cuda.synchronize()
    assert torch.equal(a.head.p, b.head.p) and _rel(b.body.p, a.body.p) < 1e-5


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_dir, graph):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import alexnet_dau_train_step as A
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    if graph:
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    tr = A.Trainer("default", global_batch=6, device=dev, rank=rank, world=world, num_classes=10, graph=graph)
    tr.set_batch()
    loss = None
    if graph:
        tr.capture(warmup=2)
        steps = 2
    else:
        steps = 4
    for _ in range(steps):
        loss = tr.step()
    gl = tr.global_loss(loss)
    torch.cuda.synchronize()
    torch.save({"head": tr.head.p.cpu(), "body": tr.body.p.cpu(), "loss": gl, "graph_error": tr.graph_error,
                "graphed": tr._graph is not None}, os.path.join(out_dir, "r%d.pt" % rank))
    A.teardown(tr, world)


def _launch_ranks(world, out_dir, graph, timeout):
    """Two rank processes of their own (not mp.spawn: a hung collective must be killable from here).
    Returns True when both finished, False when they had to be killed."""
    import subprocess
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", str(r), str(world), str(port),
                               str(out_dir), str(int(graph))]) for r in range(world)]
    ok = True
    import time
    t_end = time.time() + timeout
    for pr in procs:
        try:
            pr.wait(timeout=max(1.0, t_end - time.time()))
        except subprocess.TimeoutExpired:
            ok = False
    for pr in procs:
        if pr.poll() is None:
            pr.kill(); pr.wait()
    if ok:
        assert all(pr.returncode == 0 for pr in procs), [pr.returncode for pr in procs]
    return ok


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "cuda_graph"])
def test_two_ranks_nccl_follow_the_single_gpu_run(tmp_path, graph):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    finished = _launch_ranks(2, tmp_path, graph, timeout=240)
    if not finished:
        assert graph, "the eager 2-rank step hung"
        pytest.xfail("the 2-rank run with the step captured in a CUDA graph (NCCL collectives inside) did not exit "
                     "on this software stack (torch 2.11 / NCCL 2.28): multi-rank runs stay eager by default")
    one = _trainer(global_batch=6, device="cuda:0")
    loss = None
    for _ in range(4):
        loss = one.step()
    torch.cuda.synchronize()
    r = [torch.load(tmp_path / ("r%d.pt" % k)) for k in range(2)]
    if graph and not r[0]["graphed"]:
        pytest.xfail("CUDA-graph capture with NCCL inside was refused: %s" % r[0]["graph_error"])
    assert torch.equal(r[0]["head"], r[1]["head"]) and torch.equal(r[0]["body"], r[1]["body"])   # ranks in step
    assert _rel(r[0]["body"], one.body.p.cpu()) < 1e-4 and _rel(r[0]["head"], one.head.p.cpu()) < 1e-4
    assert abs(r[0]["loss"] - float(loss)) <= 1e-3 * abs(float(loss))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "--worker":
    sys.path.insert(0, os.path.join(ROOT, "dau-convnet_b200"))
    _worker(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], bool(int(sys.argv[6])))
