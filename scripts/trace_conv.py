import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_models_b200 import _native as nat
from distributed_tensorflow_models_b200.ops import conv_native as cn
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
def r(*s): return torch.randn(*s, generator=g).to(dev)
for name, (n, h, c, co) in {"64ch_8x8_b128": (128, 8, 64, 64), "16ch_32x32_b128": (128, 32, 16, 16)}.items():
    x = r(n, h, h, c); w = r(co, 3, 3, c) * 0.1
    for _ in range(3): cn.conv_fprop(x, w, (1, 1), (1, 1))
    buf = torch.zeros(4096, dtype=torch.int64, device=dev)
    nat.lib().dtf_debug_set_trace(C.c_void_p(buf.data_ptr()))
    cn.conv_fprop(x, w, (1, 1), (1, 1)); torch.cuda.synchronize()
    nat.lib().dtf_debug_set_trace(None)
    t = buf.cpu().tolist()
    t0 = t[3]
    print(f"== {name}: kernel start->post-setup {t[4]-t0} cyc; producer tile start {t[0]-t0}; epilogue begin {t[1]-t0} end {t[2]-t0}")
    print(" kb | P:loads_issued  empty_ok  stored  fenced | M:full_ok  committed")
    for kb in range(20):
        b = 16 + 8 * kb
        if t[b] == 0: break
        print(f" {kb:2d} | {t[b]-t0:8d} {t[b+1]-t0:8d} {t[b+2]-t0:8d} {t[b+3]-t0:8d} | {t[b+4]-t0:8d} {t[b+5]-t0:8d}")
    mm = [(t[2048 + 2 * i] - t0, t[2049 + 2 * i] - t0) for i in range(24) if t[2048 + 2 * i]]
    print(" MMA per k-block (full_ok, committed):", mm)
