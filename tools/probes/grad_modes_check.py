"""Development check: the AlexNet-DAU gradients are the same whether .grad is zero-filled and accumulated into
or dropped and assigned, and a repeated step reproduces them (guards the --lean option of the harness)."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "dau-convnet_b200"))
from alexnet_dau_train_step import AlexNetDAU  # noqa: E402

torch.manual_seed(0)
m = AlexNetDAU((1, 2)).cuda()
x = torch.rand(8, 3, 227, 227, device="cuda")
t = torch.randint(0, 1000, (8,), device="cuda")


def grads(set_none):
    for p in m.parameters():
        if set_none:
            p.grad = None
        elif p.grad is not None:
            p.grad.zero_()
    torch.nn.functional.cross_entropy(m(x), t).backward()
    return [p.grad.clone() for p in m.parameters() if p.grad is not None]


def diff(a, b):
    return max(((u - v).abs().max() / (v.abs().max() + 1e-20)).item() for u, v in zip(a, b))


g0, g1, g2, g3 = grads(True), grads(False), grads(True), grads(False)
print("assign vs accumulate: %.3e   assign vs assign: %.3e   accumulate vs accumulate: %.3e" % (
    diff(g0, g1), diff(g0, g2), diff(g1, g3)))
