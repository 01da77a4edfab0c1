"""Numerics of every native sm_100a op against the plain PyTorch fp32 reference of the same op
(forward and gradients).  tf32 tensor-core ops are compared at tf32 tolerance, the rest at fp32."""
import pytest
import torch

from distributed_tensorflow_models_b200.ops import functional as Fn

pytestmark = pytest.mark.gpu

TF32_TOL = 5e-3
FP32_TOL = 2e-5


def _rel(a, b):
    """Relative Frobenius error (robust to the few ReLU-mask flips tf32 rounding causes near zero)."""
    return float((a - b).norm() / (b.norm() + 1e-12))


def _both(fn, inputs, tol, grad_inputs=None):
    """Run fn(*inputs) on the native and the reference backend; compare outputs and input gradients."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    outs, grads = [], []
    g = torch.Generator(device="cpu").manual_seed(7)
    dy = None
    for backend in ("auto", "torch"):
        Fn.set_backend(backend)
        try:
            ins = [t.detach().clone().requires_grad_(t.is_floating_point()) if isinstance(t, torch.Tensor) else t
                   for t in inputs]
            y = fn(*ins)
            if dy is None:
                dy = torch.randn(y.shape, generator=g).to(y.device)
            y.backward(dy)
            outs.append(y.detach())
            grads.append([t.grad for t in ins if isinstance(t, torch.Tensor) and t.requires_grad])
        finally:
            Fn.set_backend("auto")
    assert _rel(outs[0], outs[1]) < tol, f"forward mismatch {_rel(outs[0], outs[1])}"
    for a, b in zip(grads[0], grads[1]):
        if a is None and b is None:
            continue
        assert _rel(a, b) < tol, f"grad mismatch {_rel(a, b)}"


def _r(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("cfg", [
    (4, 32, 32, 3, 16, 3, 1, "SAME", False, False),     # stem: 3 input channels (padded to 4 internally)
    (4, 32, 32, 16, 32, 3, 2, ((1, 1), (1, 1)), False, False),
    (4, 24, 24, 3, 64, 5, 1, "SAME", True, True),
    (2, 35, 35, 48, 64, 5, 1, "SAME", False, False),
    (2, 12, 12, 64, 64, 3, 2, "SAME", True, True),      # asymmetric SAME padding
    (2, 32, 32, 3, 64, 11, 1, "VALID", True, True),
])
def test_conv2d(cfg):
    n, h, w, ci, co, k, s, pad, bias, relu = cfg
    x = _r(n, h, w, ci, seed=1)
    wt = _r(co, k, k, ci, seed=2, scale=(k * k * ci) ** -0.5)
    b = (torch.sign(_r(co, seed=3)) * 6.0) if bias else None     # keeps pre-activations away from the ReLU kink
    if bias:
        _both(lambda x, wt, b: Fn.conv2d(x, wt, b, s, pad, relu=relu), [x, wt, b], TF32_TOL)
    else:
        _both(lambda x, wt: Fn.conv2d(x, wt, None, s, pad, relu=relu), [x, wt], TF32_TOL)


def test_conv_rect_kernels():
    x = _r(2, 17, 17, 128, seed=1)
    for kh, kw in ((1, 7), (7, 1), (1, 3), (3, 1)):
        wt = _r(96, kh, kw, 128, seed=2, scale=(kh * kw * 128) ** -0.5)
        _both(lambda x, wt: Fn.conv2d(x, wt, None, 1, "SAME"), [x, wt], TF32_TOL)


@pytest.mark.parametrize("n,k,m", [(128, 2304, 384), (128, 64, 10), (32, 2048, 1001), (8, 192, 10)])
def test_linear(n, k, m):
    x, w, b = _r(n, k, seed=1), _r(m, k, seed=2, scale=k ** -0.5), torch.sign(_r(m, seed=3)) * 6.0
    _both(lambda x, w, b: Fn.linear(x, w, b, relu=True), [x, w, b], TF32_TOL)


@pytest.mark.parametrize("c,relu,has_gamma", [(16, True, True), (64, False, True), (80, True, False), (448, True, False)])
def test_batch_norm_train(c, relu, has_gamma):
    x = _r(8, 9, 9, c, seed=1) * 2 + 0.5
    gamma = (_r(c, seed=2) * 0.1 + 1.0) if has_gamma else None
    beta = _r(c, seed=3) * 0.1
    mms, mvs = [], []

    def run(x, *gb):
        mm = torch.zeros(c, device="cuda")
        mv = torch.ones(c, device="cuda")
        gam = gb[0] if has_gamma else None
        bet = gb[-1]
        y = Fn.batch_norm(x, gam, bet, mm, mv, training=True, decay=0.9, eps=1e-3, relu=relu,
                          unbiased_moving_var=True)
        mms.append(mm)
        mvs.append(mv)
        return y
    ins = [x] + ([gamma] if has_gamma else []) + [beta]
    _both(run, ins, 2e-4)
    assert _rel(mms[0], mms[1]) < 1e-4 and _rel(mvs[0], mvs[1]) < 1e-4


def test_batch_norm_eval():
    c = 32
    x = _r(4, 5, 5, c, seed=1)
    mm, mv = _r(c, seed=4) * 0.2, (_r(c, seed=5).abs() + 0.5)
    beta = _r(c, seed=3)
    _both(lambda x, beta: Fn.batch_norm(x, None, beta, mm, mv, training=False, decay=0.9, eps=1e-3, relu=True),
          [x, beta], 1e-4)


@pytest.mark.parametrize("k,s,pad", [(3, 2, "SAME"), (2, 2, "VALID"), (3, 2, "VALID"), (3, 1, "SAME")])
def test_pools(k, s, pad):
    x = _r(3, 13, 13, 32, seed=1)
    _both(lambda x: Fn.max_pool(x, k, s, pad), [x], FP32_TOL)
    _both(lambda x: Fn.avg_pool(x, k, s, pad), [x], FP32_TOL)


def test_global_avg_pool():
    x = _r(16, 8, 8, 64, seed=1)
    _both(lambda x: Fn.global_avg_pool(x), [x], FP32_TOL)


def test_lrn():
    x = _r(4, 12, 12, 64, seed=1) * 3
    _both(lambda x: Fn.lrn(x, 4, 1.0, 0.001 / 9.0, 0.75), [x], 1e-4)


@pytest.mark.parametrize("c,smooth,weight", [(10, 0.0, 1.0), (1001, 0.1, 0.4)])
def test_softmax_xent(c, smooth, weight):
    logits = _r(32, c, seed=1) * 3
    labels = torch.randint(0, c, (32,), generator=torch.Generator().manual_seed(1)).cuda()
    _both(lambda z: Fn.softmax_cross_entropy(z, labels, smooth, weight), [logits], 1e-4)


def test_dropout_statistics_and_backward_mask():
    from distributed_tensorflow_models_b200.ops import native_ops
    x = torch.ones(1 << 16, device="cuda", requires_grad=True)
    y = Fn.dropout(x, 0.8, True)
    kept = (y > 0).float().mean().item()
    assert abs(kept - 0.8) < 0.01
    assert torch.allclose(y[y > 0], torch.full_like(y[y > 0], 1 / 0.8))
    y.sum().backward()
    assert torch.equal(x.grad > 0, y > 0)        # same mask regenerated in backward
    native_ops.advance_dropout_seed(x.device)
    y2 = Fn.dropout(x, 0.8, True)
    assert not torch.equal(y2 > 0, y > 0)


def test_fused_update_matches_reference():
    from distributed_tensorflow_models_b200.optim import EmaConfig, OptimizerConfig, ShardOptimizer
    for kind in ("sgd", "rmsprop"):
        cfg = OptimizerConfig(kind=kind, weight_decay=2e-3, ema=EmaConfig(0.9999))
        w0 = _r(5000, seed=1)
        opts = []
        for dev in ("cuda", "cpu"):
            w = w0.to(dev).clone()
            o = ShardOptimizer(cfg, w, 0, 5000, 3000)
            for step in range(3):
                g = _r(5000, seed=10 + step).to(dev)
                o.apply(g, 0.05, 1.0, num_updates=step)
            opts.append(o)
        assert _rel(opts[0].w.cpu(), opts[1].w) < 1e-5
        assert _rel(opts[0].ema.cpu(), opts[1].ema) < 1e-5


def test_dgrad_epilogue_bn_backward_sums_match_reduce_kernel():
    """The fused BN-backward reductions of the dgrad epilogue equal the stand-alone reduce kernel."""
    import ctypes as C
    from distributed_tensorflow_models_b200 import _native as nat
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    n, h, w, ci, co = 8, 16, 16, 32, 64
    x = _r(n, h, w, ci, seed=1)
    wt = _r(co, 3, 3, ci, seed=2, scale=0.05)
    dy = _r(n, h, w, co, seed=3)
    scale, shift = _r(ci, seed=4).abs() + 0.5, _r(ci, seed=5) * 0.3
    w_t = cn.weight_transpose(wt)
    sums = torch.zeros(2 * ci, device="cuda")
    da = cn.conv_dgrad(dy, w_t, (h, w), (1, 1), (1, 1), bn_bwd=(x, scale, shift, sums))
    ref = torch.zeros(2 * ci, device="cuda")
    vp = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    nat.check(nat.lib().dtf_bn_bwd_reduce(vp(da), vp(x), vp(scale), vp(shift), 1, C.c_longlong(n * h * w), ci, ci, ci,
                                          vp(ref), nat.stream_ptr()), "reduce")
    torch.cuda.synchronize()
    assert _rel(sums, ref) < 1e-4
    mask = (x * scale + shift > 0).float()
    assert _rel(sums[:ci], (da * mask).reshape(-1, ci).sum(0)) < 1e-4


# ---------------------------------------------------------------------------------------------------
# every convolution kernel path against the fp32 reference (the dispatcher's choices can be forced off)
# ---------------------------------------------------------------------------------------------------
def _conv_ref(x, wt, stride, pad, scale=None, shift=None):
    import torch.nn.functional as F
    xa = x if scale is None else torch.relu(x * scale + shift)
    y = F.conv2d(xa.permute(0, 3, 1, 2), wt.permute(0, 3, 1, 2), None, stride=stride, padding=pad)
    return xa, y.permute(0, 2, 3, 1).contiguous()


class _Paths:
    """Context manager: select the fprop/dgrad kernel (wint | win | tma | gather) through the debug switches."""

    def __init__(self, path):
        self.path = path

    def __enter__(self):
        from distributed_tensorflow_models_b200 import _native as nat
        lib = nat.lib()
        lib.dtf_debug_set_conv_wint(1 if self.path == "wint" else 0, 64, 1)
        lib.dtf_debug_set_conv_win(2 if self.path == "win" else 0)
        lib.dtf_debug_set_conv_tma(0 if self.path == "gather" else 1)
        return self

    def __exit__(self, *exc):
        from distributed_tensorflow_models_b200 import _native as nat
        lib = nat.lib()
        lib.dtf_debug_set_conv_wint(1, 64, 1)
        lib.dtf_debug_set_conv_win(1)
        lib.dtf_debug_set_conv_tma(1)


@pytest.mark.parametrize("path", ["wint", "win", "tma", "gather"])
@pytest.mark.parametrize("shape", [(8, 32, 32, 16, 16, 3), (8, 16, 16, 32, 32, 3), (9, 8, 8, 64, 64, 3),
                                   (3, 12, 12, 64, 32, 5)])
def test_conv_kernel_paths_fprop_dgrad(path, shape):
    """Sliding-window (channel-major and pixel-major), hardware-im2col and software-gather kernels: same result."""
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    torch.backends.cudnn.allow_tf32 = False
    n, h, w, ci, co, k = shape
    pad = (k - 1) // 2
    x = _r(n, h, w, ci, seed=1)
    wt = _r(co, k, k, ci, seed=2, scale=(k * k * ci) ** -0.5)
    scale, shift = _r(ci, seed=4).abs() + 0.5, _r(ci, seed=5) * 0.3
    dy = _r(n, h, w, co, seed=3)
    for aff in (False, True):
        xa, ref = _conv_ref(x, wt, 1, pad, scale if aff else None, shift if aff else None)
        xr = xa.clone().requires_grad_(True)
        _conv_ref(xr, wt, 1, pad)[1].backward(dy)
        with _Paths(path):
            kw = dict(in_scale=scale, in_shift=shift, in_relu=True) if aff else {}
            stats = torch.zeros(2 * co, device="cuda")
            res = _r(n, h, w, co, seed=6)
            y = cn.conv_fprop(x, wt, (1, 1), (pad, pad), residual=res, stats=stats, **kw)
            dx = cn.conv_dgrad(dy, cn.weight_transpose(wt), (h, w), (1, 1), (pad, pad))
        torch.cuda.synchronize()
        assert _rel(y, ref + res) < TF32_TOL
        assert _rel(stats[:co], (ref + res).reshape(-1, co).sum(0)) < TF32_TOL
        assert _rel(stats[co:], ((ref + res) ** 2).reshape(-1, co).sum(0)) < TF32_TOL
        assert _rel(dx, xr.grad) < TF32_TOL


def test_window_kernel_act_side_output():
    """The window kernels return relu(scale * x + shift) as a by-product (exact: fp32 FMA + max)."""
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    for (n, h, c) in ((128, 32, 16), (16, 16, 32), (5, 8, 64)):
        x = _r(n, h, h, c, seed=1)
        wt = _r(c, 3, 3, c, seed=2, scale=0.1)
        scale, shift = _r(c, seed=4).abs() + 0.5, _r(c, seed=5) * 0.3
        y, act = cn.conv_fprop(x, wt, (1, 1), (1, 1), in_scale=scale, in_shift=shift, in_relu=True, want_act=True)
        assert act is not None, "the sliding-window kernel should take this shape"
        torch.cuda.synchronize()
        xa, ref = _conv_ref(x, wt, 1, 1, scale, shift)
        assert _rel(act, xa) < 1e-6
        assert _rel(y, ref) < TF32_TOL


@pytest.mark.parametrize("mode", ["mn", "reg"])
@pytest.mark.parametrize("shape", [(8, 32, 32, 16, 16, 3, 1), (8, 16, 16, 32, 64, 3, 2), (4, 8, 8, 64, 64, 3, 1),
                                   (2, 17, 17, 128, 96, 1, 1), (3, 12, 12, 48, 64, 5, 1)])
def test_wgrad_kernels(mode, shape):
    """MN-major (copy-only producers) and K-major (register-transposing) weight-gradient kernels vs autograd,
    with and without the fused input transform."""
    from distributed_tensorflow_models_b200 import _native as nat
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    torch.backends.cudnn.allow_tf32 = False
    n, h, w, ci, co, k, st = shape
    pad = (k - 1) // 2
    x = _r(n, h, w, ci, seed=1)
    wt = _r(co, k, k, ci, seed=2, scale=0.1)
    scale, shift = _r(ci, seed=4).abs() + 0.5, _r(ci, seed=5) * 0.3
    nat.lib().dtf_debug_set_wgrad_mn(1 if mode == "mn" else 0)
    try:
        for aff in (False, True):
            xa = torch.relu(x * scale + shift) if aff else x
            wr = wt.clone().requires_grad_(True)
            yr = _conv_ref(xa, wr, st, pad)[1]
            dy = _r(*yr.shape, seed=3)
            yr.backward(dy)
            kw = dict(in_scale=scale, in_shift=shift, in_relu=True) if aff else {}
            dw = cn.conv_wgrad(x, dy, wt.shape, (st, st), (pad, pad), **kw)
            torch.cuda.synchronize()
            assert _rel(dw, wr.grad) < TF32_TOL
    finally:
        nat.lib().dtf_debug_set_wgrad_mn(1)


def test_strided_dgrad_zero_insertion_path():
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    torch.backends.cudnn.allow_tf32 = False
    for (n, h, ci, co) in ((8, 32, 16, 32), (8, 16, 32, 64)):
        x = _r(n, h, h, ci, seed=1).requires_grad_(True)
        wt = _r(co, 3, 3, ci, seed=2, scale=0.1)
        y = _conv_ref(x, wt, 2, 1)[1]
        dy = _r(*y.shape, seed=3)
        y.backward(dy)
        dx = cn.conv_dgrad(dy, cn.weight_transpose(wt), (h, h), (2, 2), (1, 1))
        torch.cuda.synchronize()
        assert _rel(dx, x.grad) < TF32_TOL


def test_weight_transpose_batched():
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    ws = [_r(16, 3, 3, 16, seed=1), _r(64, 1, 1, 32, seed=2), _r(10, 1, 1, 64, seed=3), _r(48, 5, 5, 20, seed=4)]
    batch = cn.WeightTransposeBatch()
    for w in ws:
        batch.register(w)
    outs = [batch.get(w) for w in ws]
    torch.cuda.synchronize()
    for w, o in zip(ws, outs):
        assert torch.equal(o, w.permute(3, 1, 2, 0).contiguous())
    ws[0].mul_(2.0)                       # a new forward pass re-registers -> the next get() refreshes everything
    batch.register(ws[0])
    assert torch.equal(batch.get(ws[0]), ws[0].permute(3, 1, 2, 0).contiguous())


def test_deferred_bn_finalize_in_window_kernel():
    """A lazy BnCoef is finalized by the sliding-window convolution itself: same output, same published
    coefficients and moving statistics as the stand-alone finalize kernel."""
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    from distributed_tensorflow_models_b200.ops import native_ops as no
    n, h, c = 16, 16, 32
    x = _r(n, h, h, c, seed=1) * 1.7 + 0.3
    wt = _r(c, 3, 3, c, seed=2, scale=0.1)
    gamma, beta = _r(c, seed=3).abs() + 0.5, _r(c, seed=4) * 0.2
    outs = []
    for lazy in (False, True):
        mm, mv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
        st = no.bn_stats(x)
        coef = no.bn_prepare(st, gamma, beta, mm, mv, n * h * h, 1e-5, 0.9, True, True, True, lazy=lazy)
        assert (coef.pending is not None) == lazy
        y = cn.conv_fprop(x, wt, (1, 1), (1, 1), in_scale=coef.scale, in_shift=coef.shift, in_relu=True,
                          bn_pending=coef)
        torch.cuda.synchronize()
        assert coef.pending is None
        outs.append((y, coef.scale.clone(), coef.shift.clone(), coef.mean.clone(), coef.invstd.clone(), mm, mv))
    for a, b in zip(outs[0], outs[1]):
        assert _rel(b, a) < 1e-5


@pytest.mark.parametrize("k,pad", [(5, "SAME"), (7, "SAME"), (3, ((1, 1), (1, 1)))])
def test_one_live_tap_convolution_on_1x1_maps(k, pad):
    """AlexNet / VGG fc6: a k x k convolution on a 1 x 1 map runs as the GEMM of its single live tap
    (native_ops._OneTapFn); values and all gradients equal the full convolution's, dead taps get exact zeros."""
    from distributed_tensorflow_models_b200.ops import native_ops
    n, ci, co = 24, 64, 96
    x = _r(n, 1, 1, ci, seed=1)
    wt = _r(co, k, k, ci, seed=2, scale=ci ** -0.5)
    b = _r(co, seed=3)
    _both(lambda x, wt, b: Fn.conv2d(x, wt, b, 1, pad, relu=True), [x, wt, b], TF32_TOL)
    # into a pre-allocated .grad slot (the flat PS gradient buffer): accumulates, dead taps untouched
    w = torch.nn.Parameter(wt.clone())
    w.grad = torch.full_like(w, 0.5)
    xi = x.clone().requires_grad_(True)
    y = Fn.conv2d(xi, w, None, 1, pad)
    dy = _r(*y.shape, seed=4)
    y.backward(dy)
    native_ops.one_tap_enabled[0] = False
    try:
        w2 = torch.nn.Parameter(wt.clone())
        w2.grad = torch.full_like(w2, 0.5)
        x2 = x.clone().requires_grad_(True)
        Fn.conv2d(x2, w2, None, 1, pad).backward(dy)
    finally:
        native_ops.one_tap_enabled[0] = True
    assert _rel(w.grad, w2.grad) < TF32_TOL and _rel(xi.grad, x2.grad) < TF32_TOL
    c = k // 2
    dead = w.grad.clone()
    dead[:, c, c, :] = 0.5
    assert torch.equal(dead, torch.full_like(dead, 0.5))


def test_tf32x3_precision_brings_convolutions_to_fp32_accuracy():
    """--precision tf32x3: operands split into a TF32-exact part and the remainder, three accumulating passes of the
    same tcgen05 kernels.  Against an fp64 reference the error drops from ~1e-3 (TF32) to ~1e-6 (an fp32 multiply)."""
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    x = _r(8, 16, 16, 32, seed=1)
    w = _r(48, 3, 3, 32, seed=2, scale=(9 * 32) ** -0.5)
    dy = _r(8, 16, 16, 48, seed=3)
    x64, w64, dy64 = x.double(), w.double(), dy.double()
    xn = x64.permute(0, 3, 1, 2).requires_grad_(True)
    wn = w64.permute(0, 3, 1, 2).requires_grad_(True)
    yr = torch.nn.functional.conv2d(xn, wn, padding=1)
    yr.backward(dy64.permute(0, 3, 1, 2))
    ref = (yr.permute(0, 2, 3, 1), xn.grad.permute(0, 2, 3, 1), wn.grad.permute(0, 2, 3, 1))
    errs = {}
    for prec in ("tf32", "tf32x3"):
        cn.set_precision(prec)
        try:
            y = cn.conv_fprop(x, w, (1, 1), (1, 1))
            dx = cn.conv_dgrad(dy, cn.weight_transpose(w), (16, 16), (1, 1), (1, 1))
            dw = cn.conv_wgrad(x, dy, w.shape, (1, 1), (1, 1))
        finally:
            cn.set_precision("tf32")
        errs[prec] = [_rel(a.double(), b) for a, b in zip((y, dx, dw), ref)]
    assert all(e < 3e-3 for e in errs["tf32"]), errs
    assert all(e < 5e-6 for e in errs["tf32x3"]), errs
    assert all(a < b / 50 for a, b in zip(errs["tf32x3"], errs["tf32"])), errs
