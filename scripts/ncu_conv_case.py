import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_models_b200.ops import conv_native as cn
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
def r(*s): return torch.randn(*s, generator=g).to(dev)
case = sys.argv[1] if len(sys.argv) > 1 else "res16"
if case == "res16":
    x = r(128, 32, 32, 16); w = r(16, 3, 3, 16) * 0.1
elif case == "res64":
    x = r(128, 8, 8, 64); w = r(64, 3, 3, 64) * 0.1
elif case == "inc35":            # mixed_35x35x288: 3x3 96 -> 96 on 35x35, batch 32
    x = r(32, 35, 35, 96); w = r(96, 3, 3, 96) * 0.05
elif case == "inc8":             # mixed_8x8x2048: 1x1 2048 -> 448 on 8x8
    x = r(32, 8, 8, 2048); w = r(448, 1, 1, 2048) * 0.02
else:                            # inc17 — mixed_17x17x768: 1x7 192 -> 192 on 17x17
    x = r(32, 17, 17, 192); w = r(192, 1, 7, 192) * 0.05
pad = ((w.shape[1] - 1) // 2, (w.shape[2] - 1) // 2)
dy = torch.randn_like(cn.conv_fprop(x, w, (1, 1), pad))
for _ in range(6):
    cn.conv_fprop(x, w, (1, 1), pad)
    cn.conv_wgrad(x, dy, w.shape, (1, 1), pad)
torch.cuda.synchronize()
print("ok")
