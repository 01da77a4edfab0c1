"""Timing of the zoo-only bandwidth kernels (csrc/zoo_ops.cu) against their byte roofline and against PyTorch.

For every case: graph-replayed CUDA-event time of the native forward / input-gradient / filter-gradient kernels,
the bytes each must move at minimum (read input + write output, filters are noise), the fraction of the measured
HBM copy bandwidth (MEASURED_PEAKS.json, fallback 6.4 TB/s) that corresponds to, and the same op through PyTorch
(cuDNN / ATen, channels-last, TF32 off) for scale.  Inputs are sized well beyond L2 (126 MB) where the real
networks' shapes allow; smaller layers are reported as they are (they are L2-resident in the real networks too).

    python scripts/gpu_check_zoo_ops.py  > gpurun_out/zoo_ops_perf.jsonl
"""
from __future__ import annotations

import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from distributed_tensorflow_models_b200 import _native as nat  # noqa: E402
from distributed_tensorflow_models_b200.ops import functional as Fn  # noqa: E402
from distributed_tensorflow_models_b200.ops import zoo_native as zn  # noqa: E402

# name, N, H, W, C, k, stride        (MobileNet-v1 224 at batch 64 / NASNet-A mobile and large cells)
DW_CASES = [
    ("mnv1_dw_112x112x32", 64, 112, 112, 32, 3, 1),
    ("mnv1_dw_112x112x64_s2", 64, 112, 112, 64, 3, 2),
    ("mnv1_dw_56x56x128", 64, 56, 56, 128, 3, 1),
    ("mnv1_dw_28x28x256", 64, 28, 28, 256, 3, 1),
    ("mnv1_dw_14x14x512", 64, 14, 14, 512, 3, 1),
    ("mnv1_dw_7x7x1024", 64, 7, 7, 1024, 3, 1),
    ("mnv2_dw_56x56x144", 64, 56, 56, 144, 3, 1),
    ("nasnet_sep5x5_28x28x88", 32, 28, 28, 88, 5, 1),
    ("nasnet_sep7x7_42x42x168_s2", 16, 42, 42, 168, 7, 2),
    ("nasnet_sep5x5_42x42x336", 16, 42, 42, 336, 5, 1),
]


def graph_time(fn, iters: int = 20) -> float:
    """microseconds per call, CUDA-graph replay of `iters` back-to-back calls (no Python launch gaps)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) * 1000.0 / iters)
    return best


def main() -> None:
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6400.0
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(0)
    lib = nat.lib()
    for name, n, h, w, c, k, st in DW_CASES:
        x = torch.randn(n, h, w, c, generator=gen).to(dev)
        wt = (torch.randn(k, k, c, 1, generator=gen) * 0.2).to(dev)
        oh, ow, ph, pw = Fn.resolve_padding(h, w, k, k, (st, st), "SAME")
        p = zn.DwConvParams(n, h, w, c, 1, k, k, st, st, ph[0], pw[0], oh, ow)
        y = torch.empty(n, oh, ow, c, device=dev)
        dy = torch.randn(n, oh, ow, c, generator=gen).to(dev)
        dx = torch.empty_like(x)
        dw = torch.zeros_like(wt)
        sp = nat.stream_ptr
        vp = lambda t: C.c_void_p(t.data_ptr())
        t_f = graph_time(lambda: lib.dtf_dwconv_fwd(C.byref(p), vp(x), vp(wt), None, vp(y), sp()))
        t_d = graph_time(lambda: lib.dtf_dwconv_dgrad(C.byref(p), vp(dy), vp(wt), vp(dx), sp()))
        t_w = graph_time(lambda: lib.dtf_dwconv_wgrad(C.byref(p), vp(x), vp(dy), vp(dw), sp()))
        # PyTorch: grouped convolution, channels-last memory format
        xt = x.permute(0, 3, 1, 2)                      # NCHW view of the NHWC buffer == channels_last
        wtt = wt.permute(2, 3, 0, 1).contiguous(memory_format=torch.channels_last)
        pad_t = (pw[0], pw[1], ph[0], ph[1])
        xp = F.pad(xt, pad_t).contiguous(memory_format=torch.channels_last)
        t_tf = graph_time(lambda: F.conv2d(xp, wtt, None, stride=st, groups=c))
        yt = F.conv2d(xp, wtt, None, stride=st, groups=c)
        err = float((yt.permute(0, 2, 3, 1)[:, :oh, :ow] - y).abs().max() / yt.abs().max())
        dyt = dy.permute(0, 3, 1, 2)
        t_td = graph_time(lambda: torch.nn.grad.conv2d_input(xp.shape, wtt, dyt, stride=st, groups=c))
        t_tw = graph_time(lambda: torch.nn.grad.conv2d_weight(xp, wtt.shape, dyt, stride=st, groups=c))
        bytes_io = 4 * (x.numel() + y.numel())
        rec = {"case": name, "shape": [n, h, w, c], "k": k, "stride": st, "in_out_MB": round(bytes_io / 1e6, 1),
               "fwd_us": round(t_f, 1), "dgrad_us": round(t_d, 1), "wgrad_us": round(t_w, 1),
               "fwd_GBs": round(bytes_io / t_f / 1e3, 0), "fwd_frac_of_hbm_copy": round(bytes_io / t_f / 1e3 / hbm, 3),
               "dgrad_frac_of_hbm_copy": round(bytes_io / t_d / 1e3 / hbm, 3),
               "wgrad_frac_of_hbm_copy": round(bytes_io / t_w / 1e3 / hbm, 3),
               "torch_fwd_us": round(t_tf, 1), "torch_dgrad_us": round(t_td, 1), "torch_wgrad_us": round(t_tw, 1),
               "fwd_max_rel_err_vs_torch": err}
        print(json.dumps(rec), flush=True)
        del x, y, dy, dx, xp, yt
    # the pointwise kernels: activation / instance norm / resize / pad on a 256 MB tensor
    x = torch.randn(16, 256, 256, 64, generator=gen).to(dev)
    dyx = torch.randn_like(x)
    nbytes = 4 * x.numel()
    y = torch.empty_like(x)
    vp = lambda t: C.c_void_p(t.data_ptr())
    sp = nat.stream_ptr
    t = graph_time(lambda: lib.dtf_act_fwd(vp(x), vp(y), C.c_longlong(x.numel()), 0, C.c_float(0.0), sp()), 5)
    print(json.dumps({"case": "relu6_fwd_256MB", "us": round(t, 1), "frac_of_hbm_copy": round(2 * nbytes / t / 1e3 / hbm, 3),
                      "torch_us": round(graph_time(lambda: F.relu6(x), 5), 1)}), flush=True)
    t = graph_time(lambda: lib.dtf_act_bwd(vp(x), vp(dyx), vp(y), C.c_longlong(x.numel()), 1, C.c_float(0.2), sp()), 5)
    print(json.dumps({"case": "leaky_relu_bwd_256MB", "us": round(t, 1),
                      "frac_of_hbm_copy": round(3 * nbytes / t / 1e3 / hbm, 3)}), flush=True)
    sums = torch.empty(16, 2, 64, device=dev)
    gamma, beta = torch.ones(64, device=dev), torch.zeros(64, device=dev)
    t = graph_time(lambda: lib.dtf_instnorm_fwd(vp(x), vp(gamma), vp(beta), vp(y), vp(sums), 16, 256 * 256, 64,
                                                C.c_float(1e-5), sp()), 5)
    xt = x.permute(0, 3, 1, 2)
    print(json.dumps({"case": "instance_norm_fwd_256MB (2 reads + 1 write)", "us": round(t, 1),
                      "frac_of_hbm_copy": round(3 * nbytes / t / 1e3 / hbm, 3),
                      "torch_us": round(graph_time(lambda: F.instance_norm(xt, weight=gamma, bias=beta), 5), 1)}),
          flush=True)


if __name__ == "__main__":
    main()
