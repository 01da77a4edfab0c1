"""Pipeline timeline of CTA 0 of dtf_conv_gemm (SM clock stamps written by the kernel when armed through
dtf_debug_set_trace): when the TMA issuer got each slot, when the MMA warp saw each stage full and committed it,
and when the epilogue started / finished each tile."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_models_b200 import _native as nat
from distributed_tensorflow_models_b200.ops import conv_native as cn
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
def r(*s): return torch.randn(*s, generator=g).to(dev)
CASES = {"16ch_32x32_b128": (128, 32, 16, 16, False), "16ch_32x32_b128_inaff": (128, 32, 16, 16, True),
         "64ch_8x8_b128": (128, 8, 64, 64, False), "32ch_16x16_b128_inaff": (128, 16, 32, 32, True)}
for name, (n, h, c, co, aff) in CASES.items():
    x = r(n, h, h, c); w = r(co, 3, 3, c) * 0.1
    kw = dict(in_scale=torch.rand(c, device=dev) + 0.5, in_shift=torch.randn(c, device=dev) * 0.1, in_relu=True) if aff else {}
    for _ in range(3): cn.conv_fprop(x, w, (1, 1), (1, 1), **kw)
    buf = torch.zeros(4096, dtype=torch.int64, device=dev)
    nat.lib().dtf_debug_set_trace(C.c_void_p(buf.data_ptr()))
    cn.conv_fprop(x, w, (1, 1), (1, 1), **kw); torch.cuda.synchronize()
    nat.lib().dtf_debug_set_trace(None)
    t = buf.cpu().tolist()
    t0 = t[3]
    print(f"== {name}: setup done {t[4]-t0} cyc; kernel end {t[5]-t0} cyc")
    iss = [t[1024 + i] - t0 for i in range(512) if t[1024 + i]]
    mm = [(t[2048 + 2 * i] - t0, t[2049 + 2 * i] - t0) for i in range(512) if t[2048 + 2 * i]]
    ep = [(t[512 + 2 * i] - t0, t[513 + 2 * i] - t0) for i in range(128) if t[512 + 2 * i]]
    if len(iss) and t[1025]:      # window kernel: (slot free, fill issued) per window; MMA (window full, tile committed)
        print(" fill (slot free, issued):", [(t[1024 + 2 * i] - t0, t[1025 + 2 * i] - t0) for i in range(64) if t[1024 + 2 * i]])
    else:
        print(" TMA issue times      :", iss[:48])
    print(" MMA (full, committed):", mm[:48])
    print(" epilogue (begin, end):", ep)
    print(" epilogue detail (acc loaded, stored, stats):", [tuple(t[3072 + 4 * i + k] - t0 for k in range(3)) for i in range(16) if t[3072 + 4 * i]])
