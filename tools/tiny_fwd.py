import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dau-convnet_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch, dau_conv
from oracle import dau_oracle as O
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
N, S, F, G, H, W = int(os.environ.get("TN", 2)), 3, 2, 2, 8, int(os.environ.get("TW", 65))
x = rng.random((N, S, H, W)).astype(np.float32)
w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
mu1 = rng.uniform(-3, 3, w.shape).astype(np.float32); mu2 = rng.uniform(-3, 3, w.shape).astype(np.float32)
sg = np.full(w.shape, 0.5, np.float32)
t = [torch.from_numpy(a).to(dev) for a in (x, w, mu1, mu2, sg)]
y = dau_conv.dau_conv_op(*t, num_output=F, number_units_x=1, number_units_y=G)
torch.cuda.synchronize()
ref = O.forward(x, w, mu1, mu2, [0.5])
print("max abs err", float(np.abs(y.cpu().numpy() - ref).max()))
