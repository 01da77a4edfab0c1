"""GPU bring-up check for the tcgen05 conv / GEMM kernels against PyTorch fp32 references.

Every case runs fprop, dgrad and wgrad.  A device fault is sticky for the process, so the driver
mode (default) re-launches itself as a child per *group of cases*, resuming after the case that
faulted or hung:  ``python scripts/gpu_check_conv.py``  (writes gpurun_out/conv_check.log).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# name, N, H, W, Cin, Cout, R, S, stride, pad(h,w), extras
CASES = [
    ("fc_64x10", 128, 1, 1, 64, 10, 1, 1, 1, (0, 0), {}),
    ("fc_2304x384_bias_relu", 128, 1, 1, 2304, 384, 1, 1, 1, (0, 0), {"bias": True, "relu": True}),
    ("res3x3_16_16", 8, 32, 32, 16, 16, 3, 3, 1, (1, 1), {}),
    ("res3x3_16_16_stats_res", 8, 32, 32, 16, 16, 3, 3, 1, (1, 1), {"stats": True, "residual": True}),
    ("res3x3_16_16_inaff", 8, 32, 32, 16, 16, 3, 3, 1, (1, 1), {"in_affine": True}),
    ("stem3x3_4_16", 8, 32, 32, 4, 16, 3, 3, 1, (1, 1), {}),
    ("res3x3_s2_16_32", 8, 32, 32, 16, 32, 3, 3, 2, (1, 1), {}),
    ("res1x1_s2_16_32", 8, 32, 32, 16, 32, 1, 1, 2, (0, 0), {}),
    ("res3x3_64_64_full", 128, 8, 8, 64, 64, 3, 3, 1, (1, 1), {"stats": True}),
    ("cnn5x5_64_64", 16, 12, 12, 64, 64, 5, 5, 1, (2, 2), {"bias": True, "relu": True}),
    ("inc1x7_128", 4, 17, 17, 128, 128, 1, 7, 1, (0, 3), {}),
    ("inc3x3_s2_valid", 4, 35, 35, 32, 48, 3, 3, 2, (0, 0), {}),
    ("vgg3x3_256_512", 4, 4, 4, 256, 512, 3, 3, 1, (1, 1), {"bias": True, "relu": True}),
    ("wide_192", 4, 8, 8, 64, 192, 1, 1, 1, (0, 0), {}),
    ("wide_384", 4, 8, 8, 192, 384, 3, 3, 1, (1, 1), {}),
    ("fc_1001", 32, 1, 1, 2048, 1001, 1, 1, 1, (0, 0), {"bias": True}),
    ("alex_fc6_dead_taps", 16, 1, 1, 256, 512, 5, 5, 1, (2, 2), {"bias": True}),
    ("alex_11x11_valid", 4, 32, 32, 4, 64, 11, 11, 1, (0, 0), {}),
    ("big_res_batch128", 128, 32, 32, 16, 16, 3, 3, 1, (1, 1), {"stats": True}),
    # hardware-im2col coverage: both swizzle widths, fused input transform, strides, odd unit counts, tails
    ("b128_res16_inaff", 128, 32, 32, 16, 16, 3, 3, 1, (1, 1), {"in_affine": True, "stats": True}),
    ("b128_res32_inaff", 128, 16, 16, 32, 32, 3, 3, 1, (1, 1), {"in_affine": True, "stats": True}),
    ("b128_res64_inaff", 128, 8, 8, 64, 64, 3, 3, 1, (1, 1), {"in_affine": True, "stats": True}),
    ("b128_res16_32_s2_inaff", 128, 32, 32, 16, 32, 3, 3, 2, (1, 1), {"in_affine": True}),
    ("b128_res32_64_s2_inaff", 128, 16, 16, 32, 64, 3, 3, 2, (1, 1), {"in_affine": True}),
    ("inc_c48_5x5_tail", 3, 35, 35, 48, 64, 5, 5, 1, (2, 2), {"in_affine": True}),
    ("inc_c80_3x3_valid_tail", 3, 73, 73, 80, 192, 3, 3, 1, (0, 0), {}),
    ("inc7x1_160_inaff", 5, 17, 17, 160, 192, 7, 1, 1, (3, 0), {"in_affine": True}),
    ("vgg3x3_64_64_224", 2, 112, 112, 64, 64, 3, 3, 1, (1, 1), {"bias": True, "relu": True}),
    ("alex_5x5_s1_64_192", 8, 27, 27, 64, 192, 5, 5, 1, (2, 2), {"bias": True, "relu": True}),
]


def run_cases(start: int, end: int) -> None:
    import torch
    import torch.nn.functional as F
    from distributed_tensorflow_models_b200 import _native as nat
    from distributed_tensorflow_models_b200.ops import conv_native as cn

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(1234)
    WIN_MODE = int(os.environ.get("DTF_CHECK_WIN_MODE", "2"))     # 2: every eligible shape takes the window kernel
    nat.lib().dtf_debug_set_conv_win(WIN_MODE)
    WINT = [int(v) for v in os.environ.get("DTF_CHECK_WINT", "1,64,1").split(",")]
    nat.lib().dtf_debug_set_conv_wint(*WINT)

    def rel(a, b):
        return float((a - b).abs().max() / (b.abs().max() + 1e-12))

    only = os.environ.get("DTF_CHECK_ONLY")
    for idx in range(start, min(end, len(CASES))):
        name, n, h, w, ci, co, r, s, st, pad, ex = CASES[idx]
        if only and only not in name:
            continue
        print(json.dumps({"begin": idx, "name": name}), flush=True)
        x = torch.randn(n, h, w, ci, generator=g).to(dev)
        wt = (torch.randn(co, r, s, ci, generator=g) / (r * s * ci) ** 0.5).to(dev)
        bias = torch.randn(co, generator=g).to(dev) if ex.get("bias") else None
        in_scale = (torch.rand(ci, generator=g) + 0.5).to(dev) if ex.get("in_affine") else None
        in_shift = (torch.randn(ci, generator=g) * 0.3).to(dev) if ex.get("in_affine") else None
        xa = x
        if in_scale is not None:
            xa = torch.relu(x * in_scale + in_shift)
        ref = F.conv2d(xa.permute(0, 3, 1, 2), wt.permute(0, 3, 1, 2), bias, stride=st, padding=pad)
        ref = ref.permute(0, 2, 3, 1).contiguous()
        residual = torch.randn(ref.shape, generator=g).to(dev) if ex.get("residual") else None
        if residual is not None:
            ref = ref + residual
        if ex.get("relu"):
            ref = torch.relu(ref)
        stats = torch.zeros(2 * co, device=dev) if ex.get("stats") else None
        t0 = time.time()
        y = cn.conv_fprop(x, wt, (st, st), pad, bias=bias, residual=residual, relu=bool(ex.get("relu")),
                          in_scale=in_scale, in_shift=in_shift, in_relu=in_scale is not None, stats=stats)
        torch.cuda.synchronize()
        res = {"idx": idx, "name": name, "fprop_rel": rel(y, ref), "ms_first": (time.time() - t0) * 1e3}
        if in_scale is not None:
            y3, act = cn.conv_fprop(x, wt, (st, st), pad, in_scale=in_scale, in_shift=in_shift, in_relu=True,
                                    want_act=True)
            res["act_written"] = act is not None
            if act is not None:
                res["act_rel"] = rel(act, xa)
                res["fprop_rel_act"] = rel(y3, F.conv2d(xa.permute(0, 3, 1, 2), wt.permute(0, 3, 1, 2), None, stride=st,
                                                        padding=pad).permute(0, 2, 3, 1))
        if stats is not None:
            s1 = ref.reshape(-1, co).sum(0)
            s2 = (ref.reshape(-1, co) ** 2).sum(0)
            res["stats_sum_rel"] = rel(stats[:co], s1)
            res["stats_sq_rel"] = rel(stats[co:], s2)
        # dgrad / wgrad against autograd of the plain conv (no epilogue)
        dy = torch.randn(ref.shape, generator=g).to(dev)
        xr = xa.detach().clone().requires_grad_(True)
        wr = wt.detach().clone().requires_grad_(True)
        yr = F.conv2d(xr.permute(0, 3, 1, 2), wr.permute(0, 3, 1, 2), None, stride=st, padding=pad)
        yr.backward(dy.permute(0, 3, 1, 2))
        w_t = cn.weight_transpose(wt)
        res["wtrans_ok"] = bool(torch.equal(w_t, wt.permute(3, 1, 2, 0).contiguous()))
        dx = cn.conv_dgrad(dy, w_t, (h, w), (st, st), pad)
        torch.cuda.synchronize()
        res["dgrad_rel"] = rel(dx, xr.grad)
        dw = cn.conv_wgrad(x, dy, wt.shape, (st, st), pad, in_scale=in_scale, in_shift=in_shift,
                           in_relu=in_scale is not None)
        torch.cuda.synchronize()
        res["wgrad_rel"] = rel(dw, wr.grad)
        # timing (device events) of the three kernels
        def timeit(fn, iters=20):
            # replay a captured graph of `iters` launches: the Python / ctypes launch cost (~10 us) would
            # otherwise hide kernels this short
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                fn()
                with torch.cuda.graph(graph, stream=side):
                    for _ in range(iters):
                        fn()
            torch.cuda.synchronize()
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters * 1e3
        fp = lambda: cn.conv_fprop(x, wt, (st, st), pad, out=y, in_scale=in_scale, in_shift=in_shift,
                                   in_relu=in_scale is not None)
        dg = lambda: cn.conv_dgrad(dy, w_t, (h, w), (st, st), pad, out=dx)
        res["us_fprop"] = timeit(fp)
        res["us_dgrad"] = timeit(dg)
        wg = lambda: cn.conv_wgrad(x, dy, wt.shape, (st, st), pad, out=dw, in_scale=in_scale, in_shift=in_shift,
                                   in_relu=in_scale is not None)
        res["us_wgrad"] = timeit(wg)
        nat.lib().dtf_debug_set_wgrad_mn(0)
        dw0 = cn.conv_wgrad(x, dy, wt.shape, (st, st), pad, in_scale=in_scale, in_shift=in_shift,
                            in_relu=in_scale is not None)
        res["wgrad_rel_reg"] = rel(dw0, wr.grad)
        res["us_wgrad_reg"] = timeit(wg)
        nat.lib().dtf_debug_set_wgrad_mn(1)
        # A/B: the same two kernels with the sliding-window kernel off (hardware-im2col feed), then with the
        # software-gather producers
        nat.lib().dtf_debug_set_conv_wint(0, WINT[1], WINT[2])
        yw = cn.conv_fprop(x, wt, (st, st), pad, bias=bias, residual=residual, relu=bool(ex.get("relu")),
                           in_scale=in_scale, in_shift=in_shift, in_relu=in_scale is not None)
        dxw = cn.conv_dgrad(dy, w_t, (h, w), (st, st), pad)
        res["fprop_rel_win"] = rel(yw, ref)
        res["dgrad_rel_win"] = rel(dxw, xr.grad)
        res["us_fprop_win"] = timeit(fp)
        res["us_dgrad_win"] = timeit(dg)
        nat.lib().dtf_debug_set_conv_win(0)
        y1 = cn.conv_fprop(x, wt, (st, st), pad, bias=bias, residual=residual, relu=bool(ex.get("relu")),
                           in_scale=in_scale, in_shift=in_shift, in_relu=in_scale is not None)
        dx1 = cn.conv_dgrad(dy, w_t, (h, w), (st, st), pad)
        res["fprop_rel_im2col"] = rel(y1, ref)
        res["dgrad_rel_im2col"] = rel(dx1, xr.grad)
        res["us_fprop_im2col"] = timeit(fp)
        res["us_dgrad_im2col"] = timeit(dg)
        nat.lib().dtf_debug_set_conv_tma(0)
        y2 = cn.conv_fprop(x, wt, (st, st), pad, bias=bias, residual=residual, relu=bool(ex.get("relu")),
                           in_scale=in_scale, in_shift=in_shift, in_relu=in_scale is not None)
        dx2 = cn.conv_dgrad(dy, w_t, (h, w), (st, st), pad)
        res["fprop_rel_gather"] = rel(y2, ref)
        res["dgrad_rel_gather"] = rel(dx2, xr.grad)
        res["us_fprop_gather"] = timeit(fp)
        res["us_dgrad_gather"] = timeit(dg)
        nat.lib().dtf_debug_set_conv_tma(1)
        nat.lib().dtf_debug_set_conv_win(WIN_MODE)
        nat.lib().dtf_debug_set_conv_wint(*WINT)
        res["us_cudnn_fprop"] = timeit(lambda: F.conv2d(xa.permute(0, 3, 1, 2), wt.permute(0, 3, 1, 2), None,
                                                         stride=st, padding=pad))
        ok = all(res[k] < 5e-3 for k in res if "_rel" in k)
        res["ok"] = bool(ok)
        print(json.dumps(res), flush=True)


def driver() -> int:
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    log_path = os.path.join(out_dir, os.environ.get("DTF_CHECK_LOG", "conv_check.log"))
    start = 0
    results = {}
    with open(log_path, "w") as log:
        while start < len(CASES):
            try:
                r = subprocess.run([sys.executable, __file__, "--child", str(start), str(len(CASES))],
                                   capture_output=True, text=True, timeout=300)
                out, err, rc = r.stdout, r.stderr, r.returncode
            except subprocess.TimeoutExpired as e:
                out = e.stdout.decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
                err = "TIMEOUT"
                rc = -9
            last_begin = None
            for line in out.splitlines():
                log.write(line + "\n")
                try:
                    d = json.loads(line)
                except Exception:
                    continue
                if "begin" in d:
                    last_begin = d["begin"]
                elif "idx" in d:
                    results[d["idx"]] = d
            if rc == 0:
                break
            failed = last_begin if last_begin is not None else start
            if failed not in results:
                results[failed] = {"idx": failed, "name": CASES[failed][0], "ok": False, "crash": err[-600:]}
                log.write(json.dumps(results[failed]) + "\n")
            start = failed + 1
        n_ok = sum(1 for v in results.values() if v.get("ok"))
        summary = {"cases": len(CASES), "ok": n_ok}
        log.write(json.dumps(summary) + "\n")
    for i in sorted(results):
        print(json.dumps(results[i]))
    print(json.dumps(summary))
    return 0 if n_ok == len(CASES) else 1


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        run_cases(int(sys.argv[2]), int(sys.argv[3]))
    else:
        sys.exit(driver())
