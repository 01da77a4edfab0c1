"""ncu target: the depthwise forward / stride-2 input-gradient / filter-gradient kernels once each on a MobileNet-v1
shape (64 x 112 x 112 x 64, 3x3).  `ncu --set full --clock-control none --import-source on -k regex:dwconv ...`"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_tensorflow_models_b200 import _native as nat  # noqa: E402
from distributed_tensorflow_models_b200.ops import functional as Fn  # noqa: E402
from distributed_tensorflow_models_b200.ops import zoo_native as zn  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
lib = nat.lib()
vp = lambda t: C.c_void_p(t.data_ptr())
for stride in (1, 2):
    n, h, w, c, k = 64, 112, 112, 64, 3
    x = torch.randn(n, h, w, c, generator=g).to(dev)
    wt = torch.randn(k, k, c, 1, generator=g).to(dev)
    oh, ow, ph, pw = Fn.resolve_padding(h, w, k, k, (stride, stride), "SAME")
    p = zn.DwConvParams(n, h, w, c, 1, k, k, stride, stride, ph[0], pw[0], oh, ow)
    y = torch.empty(n, oh, ow, c, device=dev)
    dy = torch.randn(n, oh, ow, c, generator=g).to(dev)
    dx = torch.empty_like(x)
    dw = torch.zeros_like(wt)
    lib.dtf_dwconv_fwd(C.byref(p), vp(x), vp(wt), None, vp(y), nat.stream_ptr())
    lib.dtf_dwconv_dgrad(C.byref(p), vp(dy), vp(wt), vp(dx), nat.stream_ptr())
    lib.dtf_dwconv_wgrad(C.byref(p), vp(x), vp(dy), vp(dw), nat.stream_ptr())
torch.cuda.synchronize()
print("ok")
