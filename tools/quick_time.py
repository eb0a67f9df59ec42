"""Quick per-kernel timing of one DAU layer config on cuda:0 (development aid)."""
import argparse
import os
import sys
import ctypes as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dau-convnet_b200"))
import torch
import dau_conv

ap = argparse.ArgumentParser()
ap.add_argument("--N", type=int, default=128); ap.add_argument("--S", type=int, default=128)
ap.add_argument("--F", type=int, default=128); ap.add_argument("--G", type=int, default=4)
ap.add_argument("--H", type=int, default=64); ap.add_argument("--W", type=int, default=64)
ap.add_argument("--ks", type=int, default=9); ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()
dev = torch.device("cuda:0")
tf, ms = C.c_double(0), C.c_double(0)
dau_conv._lib.check(dau_conv._lib.lib.dau_measure_fma_peak(0, 20000, C.byref(tf), C.byref(ms)))
print("fma_peak_tflops %.2f (%.2f ms)" % (tf.value, ms.value))
N, S, F, G, H, W = a.N, a.S, a.F, a.G, a.H, a.W
m = a.ks // 2
x = torch.rand((N, S, H, W), device=dev); e = torch.randn((N, F, H, W), device=dev)
w = 0.1 * torch.randn((1, S, G, F), device=dev)
mu1 = (torch.rand((1, S, G, F), device=dev) * 2 * m - m).clamp(-m + 0.01, m - 0.01)
mu2 = (torch.rand((1, S, G, F), device=dev) * 2 * m - m).clamp(-m + 0.01, m - 0.01)
sg = torch.full((1, S, G, F), 0.5, device=dev)
h = dau_conv.get_handle(dev, num_output=F, number_units_x=1, number_units_y=G, kernel_size=a.ks, pad=m)
h.set_sigma_hint(0.5)
h.enable_timing(True)
U = N * F * S * G * H * W
for it in range(a.iters):
    y = h.forward(x, w, mu1, mu2, sg)
    g = h.backward(x, e, w, mu1, mu2, sg)
    torch.cuda.synchronize()
    tfw, tbw = h.timing(0), h.timing(1)
    tot = sum(tfw.values()) + sum(tbw.values())
    print("iter %d total %.3f ms  img/s %.1f  TFLOPS(32U) %.2f" % (it, tot, N / tot * 1e3, 32 * U / tot / 1e9))
    print("   fwd", {k: round(v, 3) for k, v in tfw.items()})
    print("   bwd", {k: round(v, 3) for k, v in tbw.items()})
print("gather_forward TFLOPS %.2f  gather_backprop %.2f  grad_units %.2f" % (
    8 * U / tfw["gather_forward"] / 1e9, 8 * U / tbw["gather_backprop"] / 1e9, 16 * U / tbw["grad_units"] / 1e9))
