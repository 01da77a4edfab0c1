"""Numerics of the zoo-only native kernels (csrc/zoo_ops.cu) and of the transposed convolution (tcgen05 dgrad
kernel run forwards) against the plain PyTorch fp32 reference of the same op — forward and every gradient."""
import pytest
import torch

from distributed_tensorflow_models_b200 import _native as nat
from distributed_tensorflow_models_b200.ops import functional as Fn

pytestmark = pytest.mark.gpu

TF32_TOL = 5e-3
FP32_TOL = 3e-5


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _both(fn, inputs, tol):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    outs, grads = [], []
    g = torch.Generator(device="cpu").manual_seed(11)
    dy = None
    launches = 0
    for backend in ("auto", "torch"):
        Fn.set_backend(backend)
        try:
            ins = [t.detach().clone().requires_grad_(True) if isinstance(t, torch.Tensor) else t for t in inputs]
            before = nat.launch_count()
            y = fn(*ins)
            if dy is None:
                dy = torch.randn(y.shape, generator=g).to(y.device)
            y.backward(dy)
            if backend == "auto":
                launches = nat.launch_count() - before
            outs.append(y.detach())
            grads.append([t.grad for t in ins if isinstance(t, torch.Tensor)])
        finally:
            Fn.set_backend("auto")
    assert launches >= 1, "the native kernels did not run"
    assert outs[0].shape == outs[1].shape
    assert _rel(outs[0], outs[1]) < tol, f"forward mismatch {_rel(outs[0], outs[1])}"
    for i, (a, b) in enumerate(zip(grads[0], grads[1])):
        assert _rel(a, b) < tol, f"grad {i} mismatch {_rel(a, b)}"


def _r(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("cfg", [
    # n, h, w, c, mult, k, stride, padding, bias
    (4, 28, 28, 32, 1, 3, 1, "SAME", False),       # vector kernel, 4 outputs per thread
    (4, 28, 28, 32, 1, 3, 2, "SAME", True),        # vector kernel, stride 2 (asymmetric SAME padding)
    (2, 17, 19, 64, 1, 5, 1, "SAME", False),       # odd extents: partial output tiles
    (2, 21, 21, 44, 1, 7, 2, "SAME", False),       # NASNet 7x7 separable, stride 2
    (2, 14, 14, 24, 1, 3, 1, "VALID", True),
    (2, 23, 17, 40, 1, 5, 2, "SAME", False),       # 5x5 stride 2: parity-walking dgrad, sliding-window wgrad
    (3, 15, 15, 36, 1, 7, 1, "SAME", True),        # 7x7 stride 1 (49 accumulators per thread in the filter gradient)
    (2, 16, 16, 72, 1, 3, 2, "VALID", False),      # three channel groups, last one partial
    # enough (image, row-group, channel-group) work that the filter gradient sweeps 4 (7x7: 2) output rows at once;
    # heights that leave a partial last group
    (16, 38, 11, 256, 1, 3, 1, "SAME", False),
    (16, 75, 9, 256, 1, 3, 2, "SAME", False),
    (16, 38, 9, 256, 1, 5, 1, "SAME", True),
    (8, 21, 9, 512, 1, 7, 1, "SAME", False),
    (8, 41, 9, 512, 1, 7, 2, "SAME", False),
    (2, 12, 12, 6, 2, 3, 1, "SAME", True),         # multiplier 2 -> generic kernels
    (2, 12, 12, 10, 1, 3, 2, "VALID", False),      # C % 4 != 0 -> generic kernels
    (2, 9, 9, 16, 1, (1, 3), 1, "SAME", False),    # single filter row
    (2, 9, 9, 16, 1, (3, 1), 1, "SAME", False),    # filter width 1 -> generic kernels
    (2, 9, 9, 16, 1, (4, 4), 2, "SAME", False),
])
def test_depthwise_conv2d(cfg):
    n, h, w, c, mult, k, s, pad, bias = cfg
    kh, kw = (k, k) if isinstance(k, int) else k
    x = _r(n, h, w, c, seed=1)
    wt = _r(kh, kw, c, mult, seed=2, scale=0.3)
    b = _r(c * mult, seed=3) if bias else None
    if bias:
        _both(lambda x_, w_, b_: Fn.depthwise_conv2d(x_, w_, b_, s, pad), [x, wt, b], FP32_TOL)
    else:
        _both(lambda x_, w_: Fn.depthwise_conv2d(x_, w_, None, s, pad), [x, wt], FP32_TOL)


@pytest.mark.parametrize("kind", ["relu6", "leaky_relu", "tanh", "sigmoid"])
@pytest.mark.parametrize("shape", [(3, 7, 5, 13), (2, 16, 16, 32)])
def test_activations(kind, shape):
    x = _r(*shape, seed=4, scale=4.0)
    fn = {"relu6": Fn.relu6, "leaky_relu": lambda t: Fn.leaky_relu(t, 0.2), "tanh": Fn.tanh, "sigmoid": Fn.sigmoid}[kind]
    _both(fn, [x], 1e-5 if kind in ("relu6", "leaky_relu") else 1e-4)


@pytest.mark.parametrize("shape,affine", [((2, 16, 16, 64), True), ((3, 9, 7, 20), True), ((1, 64, 64, 32), False)])
def test_instance_norm(shape, affine):
    x = _r(*shape, seed=5, scale=2.0) + 0.5
    c = shape[-1]
    if affine:
        gamma, beta = _r(c, seed=6) * 0.5 + 1.0, _r(c, seed=7)
        _both(lambda x_, g_, b_: Fn.instance_norm(x_, g_, b_, 1e-3), [x, gamma, beta], 1e-4)
    else:
        _both(lambda x_: Fn.instance_norm(x_, None, None, 1e-3), [x], 1e-4)


@pytest.mark.parametrize("bilinear", [False, True])
@pytest.mark.parametrize("align", [False, True])
@pytest.mark.parametrize("geom", [((2, 8, 8, 16), (16, 16)), ((2, 7, 5, 12), (10, 13)), ((1, 16, 16, 8), (8, 8))])
def test_resize(bilinear, align, geom):
    shape, out_hw = geom
    x = _r(*shape, seed=8)
    fn = Fn.resize_bilinear if bilinear else Fn.resize_nearest
    _both(lambda x_: fn(x_, out_hw, align), [x], 1e-5)


@pytest.mark.parametrize("mode", ["constant", "reflect"])
@pytest.mark.parametrize("pads", [(1, 1, 1, 1), (3, 3, 3, 3), (0, 2, 1, 0)])
def test_pad_spatial(mode, pads):
    x = _r(2, 9, 11, 12, seed=9)
    _both(lambda x_: Fn.pad_spatial(x_, pads, mode), [x], 1e-6)


@pytest.mark.parametrize("cfg", [
    # n, h, c_in, c_out, k, stride, padding, bias
    (4, 8, 64, 32, 4, 2, "SAME", True),        # pix2pix / DCGAN up-sampling block
    (4, 8, 32, 16, 3, 2, "SAME", False),       # CycleGAN: 3x3 stride 2 (strided window path: 16/32/64 channels)
    (8, 1, 64, 128, 4, 1, "VALID", False),     # DCGAN: noise -> 4x4
    (2, 16, 128, 3, 4, 2, "SAME", True),       # image head: 3 output channels
    (2, 7, 24, 40, 5, 2, "VALID", False),
])
def test_conv2d_transpose(cfg):
    n, h, ci, co, k, s, pad, bias = cfg
    x = _r(n, h, h, ci, seed=10)
    w = _r(co, k, k, ci, seed=11, scale=0.1)
    if bias:
        _both(lambda x_, w_, b_: Fn.conv2d_transpose(x_, w_, b_, s, pad), [x, w, _r(co, seed=12)], TF32_TOL)
    else:
        _both(lambda x_, w_: Fn.conv2d_transpose(x_, w_, None, s, pad), [x, w], TF32_TOL)


# ---- guard-band checks: the pool has no compute-sanitizer, so out-of-bounds accesses are hunted with canaries ----------
def _guarded(shape, fill, gen=None, guard=4096):
    """A tensor of ``shape`` carved out of the middle of a larger buffer whose margins hold NaN: an out-of-bounds
    read of a neighbour poisons the result, an out-of-bounds write destroys a canary."""
    n = 1
    for s in shape:
        n *= s
    n_al = (n + 3) // 4 * 4
    buf = torch.full((guard + n_al + guard,), float("nan"), device="cuda")
    view = buf[guard:guard + n].view(shape)
    if fill == "randn":
        view.copy_(torch.randn(shape, generator=gen))
    elif fill == "zeros":
        view.zero_()
    return buf, view, (guard, n)


def _canaries_intact(buf, where):
    guard, n = where
    return bool(torch.isnan(buf[:guard]).all()) and bool(torch.isnan(buf[guard + (n + 3) // 4 * 4:]).all())


@pytest.mark.parametrize("cfg", [(2, 9, 7, 8, 3, 1), (2, 9, 7, 8, 3, 2), (2, 11, 13, 12, 5, 1), (2, 11, 13, 12, 5, 2),
                                 (1, 9, 9, 36, 7, 1), (1, 10, 9, 36, 7, 2), (3, 8, 8, 4, 3, 1),
                                 (16, 38, 9, 256, 3, 1), (16, 39, 9, 256, 5, 2), (8, 21, 9, 512, 7, 1)])
def test_depthwise_kernels_stay_in_bounds(cfg):
    import ctypes as C

    from distributed_tensorflow_models_b200.ops import zoo_native as zn
    n, h, w, c, k, s = cfg
    g = torch.Generator().manual_seed(3)
    oh, ow, ph, pw = Fn.resolve_padding(h, w, k, k, (s, s), "SAME")
    p = zn.DwConvParams(n, h, w, c, 1, k, k, s, s, ph[0], pw[0], oh, ow)
    bx, x, wx = _guarded((n, h, w, c), "randn", g)
    bw, wt, ww = _guarded((k, k, c, 1), "randn", g)
    bdy, dy, wdy = _guarded((n, oh, ow, c), "randn", g)
    by, y, wy = _guarded((n, oh, ow, c), None)
    bdx, dx, wdx = _guarded((n, h, w, c), None)
    bdw, dw, wdw = _guarded((k, k, c, 1), "zeros")
    vp = lambda t: C.c_void_p(t.data_ptr())
    lib = nat.lib()
    nat.check(lib.dtf_dwconv_fwd(C.byref(p), vp(x), vp(wt), None, vp(y), nat.stream_ptr()), "fwd")
    nat.check(lib.dtf_dwconv_dgrad(C.byref(p), vp(dy), vp(wt), vp(dx), nat.stream_ptr()), "dgrad")
    nat.check(lib.dtf_dwconv_wgrad(C.byref(p), vp(x), vp(dy), vp(dw), nat.stream_ptr()), "wgrad")
    torch.cuda.synchronize()
    for buf, where in ((by, wy), (bdx, wdx), (bdw, wdw), (bx, wx), (bw, ww), (bdy, wdy)):
        assert _canaries_intact(buf, where), "a kernel wrote outside its tensor"
    finite = torch.isfinite(y).all() and torch.isfinite(dx).all() and torch.isfinite(dw).all()
    assert finite, "a kernel read outside its tensor (NaN margin leaked into the result)"
    Fn.set_backend("torch")
    try:
        xr, wr = x.clone().requires_grad_(True), wt.clone().requires_grad_(True)
        yr = Fn.depthwise_conv2d(xr, wr, None, s, "SAME")
        yr.backward(dy.clone())
    finally:
        Fn.set_backend("auto")
    assert _rel(y, yr.detach()) < FP32_TOL and _rel(dx, xr.grad) < FP32_TOL and _rel(dw, wr.grad) < FP32_TOL


def test_pointwise_zoo_kernels_stay_in_bounds():
    import ctypes as C
    g = torch.Generator().manual_seed(4)
    lib = nat.lib()
    vp = lambda t: C.c_void_p(t.data_ptr())
    n, h, w, c = 2, 7, 5, 12
    bx, x, wx = _guarded((n, h, w, c), "randn", g)
    # activation (odd element count: scalar tail), instance norm, resize up / down, reflect pad
    b1, x1, w1 = _guarded((3, 5, 7, 13), "randn", g)
    by1, y1, wy1 = _guarded((3, 5, 7, 13), None)
    nat.check(lib.dtf_act_fwd(vp(x1), vp(y1), C.c_longlong(x1.numel()), 2, C.c_float(0.0), nat.stream_ptr()), "act")
    bs, sums, ws = _guarded((n, 2, c), None)
    by2, y2, wy2 = _guarded((n, h, w, c), None)
    nat.check(lib.dtf_instnorm_fwd(vp(x), None, None, vp(y2), vp(sums), n, h * w, c, C.c_float(1e-3), nat.stream_ptr()),
              "instnorm")
    by3, y3, wy3 = _guarded((n, 10, 13, c), None)
    nat.check(lib.dtf_resize_fwd(vp(x), vp(y3), n, h, w, c, 10, 13, 1, 0, nat.stream_ptr()), "resize")
    bdx, dx, wdx = _guarded((n, h, w, c), None)
    nat.check(lib.dtf_resize_bwd(vp(y3), vp(dx), n, h, w, c, 10, 13, 1, 0, nat.stream_ptr()), "resize bwd")
    by4, y4, wy4 = _guarded((n, h + 5, w + 3, c), None)
    nat.check(lib.dtf_pad_fwd(vp(x), vp(y4), n, h, w, c, 3, 2, 1, 2, 1, nat.stream_ptr()), "pad")
    bdx2, dx2, wdx2 = _guarded((n, h, w, c), None)
    nat.check(lib.dtf_pad_bwd(vp(y4), vp(dx2), n, h, w, c, 3, 2, 1, 2, 1, nat.stream_ptr()), "pad bwd")
    torch.cuda.synchronize()
    for buf, where in ((by1, wy1), (bs, ws), (by2, wy2), (by3, wy3), (bdx, wdx), (by4, wy4), (bdx2, wdx2), (bx, wx),
                       (b1, w1)):
        assert _canaries_intact(buf, where), "a kernel wrote outside its tensor"
    for t in (y1, sums, y2, y3, dx, y4, dx2):
        assert torch.isfinite(t).all(), "a kernel read outside its tensor"
