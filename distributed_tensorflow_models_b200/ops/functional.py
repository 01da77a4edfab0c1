"""Functional NHWC ops with TensorFlow semantics (padding rules, LRN formula, pooling divisors).

Two implementations sit behind each function:

* **reference** – plain PyTorch fp32 ops (runs anywhere; it is the numerical oracle of the test-suite and
  the engine of the CPU plumbing mode);
* **native** – the hand-written sm_100a kernels of ``libdtfb200.so`` wrapped in ``torch.autograd.Function``
  (``ops/native_ops.py``), selected automatically for CUDA tensors.

``set_backend('torch')`` forces the reference path on CUDA too — that is the cuDNN/cuBLAS baseline arm
(BASELINE.md B0/B1), never the product path.

Layouts: activations NHWC, conv weights OHWI ``[C_out, R, S, C_in]``, FC weights ``[out, in]``.
Reference semantics being reproduced are cited per function (paths relative to the reference repo).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple, Union

import torch
import torch.nn.functional as F

_STATE = {"backend": "auto"}

Padding = Union[str, Tuple[Tuple[int, int], Tuple[int, int]]]


def set_backend(name: str) -> None:
    """'auto' (native kernels on CUDA, reference on CPU), 'torch' (reference everywhere) or 'native'."""
    assert name in ("auto", "torch", "native")
    _STATE["backend"] = name


def get_backend() -> str:
    return _STATE["backend"]


def use_native(x: torch.Tensor) -> bool:
    b = _STATE["backend"]
    if b == "torch":
        return False
    if b == "native" and not x.is_cuda:
        raise RuntimeError("native backend requested for a CPU tensor")
    return x.is_cuda


# ----------------------------------------------------------------------------------------------
# padding arithmetic (TensorFlow rules)
# ----------------------------------------------------------------------------------------------
def same_pad_1d(size: int, k: int, stride: int) -> Tuple[int, int, int]:
    """TF 'SAME': out = ceil(size/stride); total pad split with the extra pixel *after*."""
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    before = total // 2
    return out, before, total - before


def resolve_padding(h: int, w: int, r: int, s: int, stride: Tuple[int, int], padding: Padding):
    """-> (out_h, out_w, (pad_top, pad_bottom), (pad_left, pad_right))"""
    if isinstance(padding, str):
        p = padding.upper()
        if p == "SAME":
            oh, pt, pb = same_pad_1d(h, r, stride[0])
            ow, pl, pr = same_pad_1d(w, s, stride[1])
            return oh, ow, (pt, pb), (pl, pr)
        if p == "VALID":
            return (h - r) // stride[0] + 1, (w - s) // stride[1] + 1, (0, 0), (0, 0)
        raise ValueError(f"unknown padding {padding}")
    (pt, pb), (pl, pr) = padding
    return (h + pt + pb - r) // stride[0] + 1, (w + pl + pr - s) // stride[1] + 1, (pt, pb), (pl, pr)


def _pair(v) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else (int(v[0]), int(v[1]))


# ----------------------------------------------------------------------------------------------
# convolution / FC
# ----------------------------------------------------------------------------------------------
def conv2d(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, stride=1,
           padding: Padding = "SAME", relu: bool = False) -> torch.Tensor:
    """NHWC convolution, OHWI weights (tf.nn.conv2d semantics — cnn/cifar10.py:221,
    inception/slim/ops.py:232, resnet/resnet_model.py:87)."""
    stride = _pair(stride)
    n, h, wd, c = x.shape
    co, r, s, ci = w.shape
    oh, ow, ph, pw = resolve_padding(h, wd, r, s, stride, padding)
    if use_native(x):
        from . import native_ops
        return native_ops.conv2d(x, w, bias, stride, (ph[0], pw[0]), (oh, ow), relu)
    xn = x.permute(0, 3, 1, 2)
    if ph[0] != ph[1] or pw[0] != pw[1]:
        xn = F.pad(xn, (pw[0], pw[1], ph[0], ph[1]))
        pad = (0, 0)
    else:
        pad = (ph[0], pw[0])
    y = F.conv2d(xn, w.permute(0, 3, 1, 2), bias, stride=stride, padding=pad)
    y = y[:, :, :oh, :ow]
    if relu:
        y = F.relu(y)
    return y.permute(0, 2, 3, 1)


def _space_to_batch(x: torch.Tensor, rate: int, ph: Tuple[int, int], pw: Tuple[int, int]):
    """Zero-pad NHWC ``x`` by (ph, pw) (plus what rounds both extents up to a multiple of ``rate``) and move the
    ``rate x rate`` sampling phases into the batch: ``[N, H, W, C] -> [rate*rate*N, H'/rate, W'/rate, C]``
    (tf.space_to_batch, which is how tf.nn.atrous_conv2d / convolution(dilation_rate=) lower a dilated filter)."""
    n, h, wd, c = x.shape
    hp, wp = h + ph[0] + ph[1], wd + pw[0] + pw[1]
    eh, ew = (-hp) % rate, (-wp) % rate
    x = F.pad(x, (0, 0, pw[0], pw[1] + ew, ph[0], ph[1] + eh))
    hq, wq = (hp + eh) // rate, (wp + ew) // rate
    x = x.reshape(n, hq, rate, wq, rate, c).permute(2, 4, 0, 1, 3, 5)
    return x.reshape(rate * rate * n, hq, wq, c).contiguous()


def _batch_to_space(y: torch.Tensor, rate: int, n: int, oh: int, ow: int) -> torch.Tensor:
    """Inverse of :func:`_space_to_batch` on the convolution's output, cropped to ``oh x ow``."""
    _, hq, wq, c = y.shape
    y = y.reshape(rate, rate, n, hq, wq, c).permute(2, 3, 0, 4, 1, 5).reshape(n, hq * rate, wq * rate, c)
    return y[:, :oh, :ow, :].contiguous()


def _by_phases(conv, x, w, bias, rate: int, ph, pw, oh: int, ow: int) -> torch.Tensor:
    """A stride-1 convolution dilated by ``rate`` as one dense VALID ``conv`` over the stacked sampling phases."""
    yb = conv(_space_to_batch(x, rate, ph, pw), w, bias, 1, "VALID")
    return _batch_to_space(yb, rate, x.shape[0], oh, ow)


def atrous_conv2d(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], rate: int,
                  padding: str = "SAME") -> torch.Tensor:
    """Dilated stride-1 convolution (slim.conv2d(rate=...) under ``output_stride``, vgg/nets/resnet_utils.py:123).

    Native path: the dilated filter samples every ``rate``-th pixel, so each of the ``rate x rate`` sampling
    phases of the padded image is an ordinary dense VALID convolution.  The phases are stacked into the batch
    (:func:`_space_to_batch`), run through the tcgen05 implicit-GEMM kernel as ONE stride-1 convolution with a
    ``rate^2`` times larger GEMM M, and interleaved back.  The weights are used as they are (no zero-stuffed
    ``(k-1)*rate+1`` filter), and autograd carries dgrad / wgrad through the same two permutations."""
    co, r, s, ci = w.shape
    er, es = (r - 1) * rate + 1, (s - 1) * rate + 1
    oh, ow, ph, pw = resolve_padding(x.shape[1], x.shape[2], er, es, (1, 1), padding)
    if use_native(x):
        if rate == 1:
            return conv2d(x, w, bias, 1, (ph, pw))
        return _by_phases(conv2d, x, w, bias, rate, ph, pw, oh, ow)
    xn = F.pad(x.permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1]))
    y = F.conv2d(xn, w.permute(0, 3, 1, 2), bias, dilation=rate)
    return y.permute(0, 2, 3, 1)


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    """x[N, in] @ w[out, in]^T + b (tf.matmul/xw_plus_b — cnn/cifar10.py:262, slim/ops.py:305)."""
    if use_native(x):
        from . import native_ops
        return native_ops.linear(x, w, bias, relu)
    y = F.linear(x, w, bias)
    return F.relu(y) if relu else y


def depthwise_conv2d(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, stride=1,
                     padding: Padding = "SAME", rate: int = 1) -> torch.Tensor:
    """tf.nn.depthwise_conv2d: filter ``[R, S, C, multiplier]`` (TensorFlow's own layout), output channel
    ``c*multiplier + m`` (vgg/nets/mobilenet_v1.py:260-277 through slim.separable_conv2d)."""
    stride = _pair(stride)
    n, h, wd, c = x.shape
    r, s, c2, mult = w.shape
    assert c2 == c
    er, es = (r - 1) * rate + 1, (s - 1) * rate + 1
    oh, ow, ph, pw = resolve_padding(h, wd, er, es, stride, padding)
    if use_native(x):
        if rate != 1:
            # same lowering as atrous_conv2d: sampling phases into the batch, dense depthwise kernel, back
            if stride != (1, 1):
                raise NotImplementedError("depthwise convolution: rate > 1 needs stride 1 (as in TensorFlow)")
            return _by_phases(depthwise_conv2d, x, w, bias, rate, ph, pw, oh, ow)
        from . import zoo_native
        return zoo_native.depthwise_conv2d(x, w, bias, stride, (ph[0], pw[0]), (oh, ow))
    xn = F.pad(x.permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1]))
    wt = w.permute(2, 3, 0, 1).reshape(c * mult, 1, r, s)
    y = F.conv2d(xn, wt, bias, stride=stride, dilation=rate, groups=c)[:, :, :oh, :ow]
    return y.permute(0, 2, 3, 1)


def conv2d_transpose_output_size(size: int, k: int, stride: int, padding: str) -> int:
    """tf.contrib.layers.conv2d_transpose output extent: SAME -> size*stride, VALID -> size*stride + max(k-stride, 0)."""
    return size * stride if padding.upper() == "SAME" else size * stride + max(k - stride, 0)


def conv2d_transpose(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, stride=2,
                     padding: str = "SAME") -> torch.Tensor:
    """slim.conv2d_transpose (vgg/nets/dcgan.py:169, pix2pix.py:90, cyclegan.py:110): the gradient of a forward
    convolution with respect to its input.  ``w`` is stored ``[C_out, R, S, C_in]``."""
    stride = _pair(stride)
    n, h, wd, c = x.shape
    co, r, s, ci = w.shape
    assert ci == c
    oh = conv2d_transpose_output_size(h, r, stride[0], padding)
    ow = conv2d_transpose_output_size(wd, s, stride[1], padding)
    # padding of the forward convolution (oh, ow) -> (h, wd) this op transposes
    fh, fw, ph, pw = resolve_padding(oh, ow, r, s, stride, padding)
    assert (fh, fw) == (h, wd), "transposed convolution geometry does not invert"
    if use_native(x):
        from . import zoo_native
        return zoo_native.conv2d_transpose(x, w, bias, stride, (ph[0], pw[0]), (oh, ow))
    # conv_transpose2d wants [C_in, C_out, R, S]; build the full output then cut the forward padding off
    y = F.conv_transpose2d(x.permute(0, 3, 1, 2), w.permute(3, 0, 1, 2), None, stride=stride)
    full_h, full_w = y.shape[2], y.shape[3]
    y = F.pad(y, (0, max(0, pw[0] + ow - full_w), 0, max(0, ph[0] + oh - full_h)))
    y = y[:, :, ph[0]:ph[0] + oh, pw[0]:pw[0] + ow]
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y.permute(0, 2, 3, 1)


# ----------------------------------------------------------------------------------------------
# batch norm
# ----------------------------------------------------------------------------------------------
def batch_norm(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor],
               moving_mean: torch.Tensor, moving_var: torch.Tensor, *, training: bool, decay: float,
               eps: float, relu: bool = False, update_moving: bool = True,
               unbiased_moving_var: bool = False) -> torch.Tensor:
    """Batch normalisation over N,H,W of an NHWC tensor.

    Training: normalise with the batch mean / *biased* variance (tf.nn.moments, slim/ops.py:117-124) and,
    when ``update_moving``, ``moving = moving*decay + batch*(1-decay)`` in place.  Eval: moving statistics
    (slim/ops.py:125-128).  ``gamma is None`` reproduces Inception's centre-only BN (ops.py:49 scale=False).
    """
    c = x.shape[-1]
    if use_native(x):
        from . import native_ops
        return native_ops.batch_norm(x, gamma, beta, moving_mean, moving_var, training, decay, eps, relu,
                                     update_moving, unbiased_moving_var)
    if training:
        xf = x.reshape(-1, c)
        mean = xf.mean(0)
        var = xf.var(0, unbiased=False)
        if update_moving:
            with torch.no_grad():
                m = xf.shape[0]
                v_upd = var * (m / max(m - 1, 1)) if unbiased_moving_var else var
                moving_mean.mul_(decay).add_(mean.detach(), alpha=1 - decay)
                moving_var.mul_(decay).add_(v_upd.detach(), alpha=1 - decay)
    else:
        mean, var = moving_mean, moving_var
    inv = torch.rsqrt(var + eps)
    scale = inv if gamma is None else inv * gamma
    shift = -mean * scale
    if beta is not None:
        shift = shift + beta
    y = x * scale + shift
    return F.relu(y) if relu else y


# ----------------------------------------------------------------------------------------------
# pooling
# ----------------------------------------------------------------------------------------------
def _pool_geometry(x, k, stride, padding):
    k, stride = _pair(k), _pair(stride)
    n, h, w, c = x.shape
    oh, ow, ph, pw = resolve_padding(h, w, k[0], k[1], stride, padding)
    return k, stride, oh, ow, ph, pw


def max_pool(x: torch.Tensor, k, stride, padding: Padding = "VALID") -> torch.Tensor:
    """tf.nn.max_pool (padding contributes -inf) — cnn/cifar10.py:228, slim/ops.py:370."""
    k, stride, oh, ow, ph, pw = _pool_geometry(x, k, stride, padding)
    if use_native(x):
        from . import native_ops
        return native_ops.max_pool(x, k, stride, (ph[0], pw[0]), (oh, ow))
    xn = x.permute(0, 3, 1, 2)
    if any(ph) or any(pw):
        xn = F.pad(xn, (pw[0], pw[1], ph[0], ph[1]), value=float("-inf"))
    y = F.max_pool2d(xn, k, stride)[:, :, :oh, :ow]
    return y.permute(0, 2, 3, 1)


def avg_pool(x: torch.Tensor, k, stride, padding: Padding = "VALID") -> torch.Tensor:
    """tf.nn.avg_pool: padded positions are excluded from the divisor — slim/ops.py:400."""
    k, stride, oh, ow, ph, pw = _pool_geometry(x, k, stride, padding)
    if use_native(x):
        from . import native_ops
        return native_ops.avg_pool(x, k, stride, (ph[0], pw[0]), (oh, ow))
    xn = x.permute(0, 3, 1, 2)
    if any(ph) or any(pw):
        ones = torch.ones((1, 1, x.shape[1], x.shape[2]), dtype=x.dtype, device=x.device)
        xs = F.avg_pool2d(F.pad(xn, (pw[0], pw[1], ph[0], ph[1])), k, stride, divisor_override=1)
        cnt = F.avg_pool2d(F.pad(ones, (pw[0], pw[1], ph[0], ph[1])), k, stride, divisor_override=1)
        y = (xs / cnt)[:, :, :oh, :ow]
    else:
        y = F.avg_pool2d(xn, k, stride)[:, :, :oh, :ow]
    return y.permute(0, 2, 3, 1)


def global_avg_pool(x: torch.Tensor) -> torch.Tensor:
    """Mean over H,W -> [N, C] (resnet/resnet_model.py:271 with an 8x8 window on an 8x8 map)."""
    return avg_pool(x, (x.shape[1], x.shape[2]), 1, "VALID").reshape(x.shape[0], x.shape[3])


# ----------------------------------------------------------------------------------------------
# LRN
# ----------------------------------------------------------------------------------------------
def lrn(x: torch.Tensor, depth_radius: int = 4, bias: float = 1.0, alpha: float = 0.001 / 9.0,
        beta: float = 0.75) -> torch.Tensor:
    """tf.nn.lrn: x / (bias + alpha * sum_{|d'-d|<=r} x[d']^2)^beta — cnn/cifar10.py:231 (note: unlike
    torch's LocalResponseNorm, alpha is *not* divided by the window size)."""
    if use_native(x):
        from . import native_ops
        return native_ops.lrn(x, depth_radius, bias, alpha, beta)
    c = x.shape[-1]
    sq = (x * x).reshape(-1, 1, c)
    k = 2 * depth_radius + 1              # windowed sum = k * zero-padded windowed mean (twice as fast as conv1d)
    ssum = (F.avg_pool1d(sq, k, 1, depth_radius, count_include_pad=True) * float(k)).reshape(x.shape)
    return x * torch.pow(bias + alpha * ssum, -beta)


# ----------------------------------------------------------------------------------------------
# dropout / misc
# ----------------------------------------------------------------------------------------------
def dropout(x: torch.Tensor, keep_prob: float, training: bool) -> torch.Tensor:
    """tf.nn.dropout(keep_prob): scale kept units by 1/keep_prob; identity when not training
    (slim/ops.py:420-424)."""
    if not training or keep_prob >= 1.0:
        return x
    if use_native(x):
        from . import native_ops
        return native_ops.dropout(x, keep_prob)
    return F.dropout(x, 1.0 - keep_prob, training=True)


def relu(x: torch.Tensor) -> torch.Tensor:
    return F.relu(x)


def _act(x: torch.Tensor, kind: str, alpha: float = 0.2) -> torch.Tensor:
    if use_native(x) and x.dtype == torch.float32:
        from . import zoo_native
        return zoo_native.activation(x, kind, alpha)
    if kind == "relu6":
        return F.relu6(x)
    if kind == "leaky_relu":
        return F.leaky_relu(x, alpha)
    if kind == "tanh":
        return torch.tanh(x)
    return torch.sigmoid(x)


def relu6(x: torch.Tensor) -> torch.Tensor:
    """tf.nn.relu6 (vgg/nets/mobilenet_v1.py:430)."""
    return _act(x, "relu6")


def leaky_relu(x: torch.Tensor, alpha: float = 0.2) -> torch.Tensor:
    """tf.nn.leaky_relu, default slope 0.2 (vgg/nets/dcgan.py:89, pix2pix.py:172)."""
    return _act(x, "leaky_relu", alpha)


def tanh(x: torch.Tensor) -> torch.Tensor:
    """tf.tanh (vgg/nets/pix2pix.py:220, cyclegan.py:271)."""
    return _act(x, "tanh")


def sigmoid(x: torch.Tensor) -> torch.Tensor:
    return _act(x, "sigmoid")


# ----------------------------------------------------------------------------------------------
# fake quantisation (tf.contrib.quantize behind --quantize)
# ----------------------------------------------------------------------------------------------
def nudged_range(mn: torch.Tensor, mx: torch.Tensor, bits: int, narrow: bool):
    """TensorFlow's range nudging: move [min, max] so that 0.0 is exactly one of the quantisation steps.
    -> (nudged_min, nudged_max, step)"""
    qmin, qmax = (1.0 if narrow else 0.0), float(2 ** bits - 1)
    # tensor / tensor: a correctly rounded fp32 division like TensorFlow's and the kernel's.  tensor / python-scalar
    # is evaluated by the CUDA backend as a multiplication by fl(1/255), which lands on the other side of the
    # 127.5 rounding tie of the default (-6, 6) range and moves every clamped value by one step.
    scale = (mx - mn) / torch.full_like(mx, qmax - qmin)
    zp = qmin - mn / scale
    # std::round: halves away from zero (torch.round is half-to-even)
    zp_r = torch.sign(zp) * torch.floor(zp.abs() + 0.5)
    zp_n = torch.where(zp < qmin, torch.full_like(zp, qmin), torch.where(zp > qmax, torch.full_like(zp, qmax), zp_r))
    return (qmin - zp_n) * scale, (qmax - zp_n) * scale, scale


class _FakeQuantRef(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, rng, gate, bits, narrow):
        mn, mx = rng[0], rng[1]
        nmin, nmax, scale = nudged_range(mn, mx, bits, narrow)
        on = (mx > mn) & (scale > 0)
        if gate is not None:
            on = on & (gate.reshape(()) != 0)
        c = torch.minimum(torch.maximum(x, nmin), nmax)
        q = torch.floor((c - nmin) * (1.0 / scale) + 0.5) * scale + nmin
        inside = (x >= nmin) & (x <= nmax)
        ctx.save_for_backward(inside | ~on)
        return torch.where(on, q, x)

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        return dy * mask, None, None, None, None


def fake_quant(x: torch.Tensor, rng: torch.Tensor, gate: Optional[torch.Tensor] = None, bits: int = 8,
               narrow_range: bool = False) -> torch.Tensor:
    """tf.fake_quant_with_min_max_vars: round ``x`` onto the ``2**bits`` (narrow: one fewer) steps of the nudged
    range ``rng = [min, max]``; the gradient passes where ``x`` lies inside that range.  ``gate`` (a scalar
    tensor): 0 makes the op the identity — the device-side form of ``quant_delay``
    (vgg/nets/mobilenet_v1_train.py:66-74,135-139)."""
    if use_native(x) and x.dtype == torch.float32:
        from . import zoo_native
        return zoo_native.fake_quant(x, rng, gate, bits, narrow_range)
    return _FakeQuantRef.apply(x, rng, gate, bits, narrow_range)


@torch.no_grad()
def quant_range_update(x: torch.Tensor, rng: torch.Tensor, decay: float, scratch: Optional[torch.Tensor] = None):
    """In place: ``rng <- [min(x), max(x)]`` when ``decay < 0`` (tf.contrib.quantize LastValueQuantize, weights),
    else ``rng <- rng*decay + [min(min(x), 0), max(max(x), 0)]*(1-decay)`` (MovingAvgQuantize, activations)."""
    if use_native(x) and x.dtype == torch.float32:
        from . import zoo_native
        assert scratch is not None
        return zoo_native.quant_range_update(x, rng, decay, scratch)
    lo, hi = torch.aminmax(x.detach())
    if decay < 0:
        rng[0], rng[1] = lo, hi
    else:
        rng[0] = rng[0] * decay + torch.clamp(lo, max=0.0) * (1 - decay)
        rng[1] = rng[1] * decay + torch.clamp(hi, min=0.0) * (1 - decay)


def instance_norm(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor],
                  eps: float = 1e-6) -> torch.Tensor:
    """tf.contrib.layers.instance_norm: per-image, per-channel moments over H, W (vgg/nets/cyclegan.py:47-60)."""
    if use_native(x):
        from . import zoo_native
        return zoo_native.instance_norm(x, gamma, beta, eps)
    mean = x.mean((1, 2), keepdim=True)
    var = x.var((1, 2), unbiased=False, keepdim=True)
    y = (x - mean) * torch.rsqrt(var + eps)
    if gamma is not None:
        y = y * gamma
    if beta is not None:
        y = y + beta
    return y


def _resize_index(out: int, size: int, align_corners: bool, device):
    scale = (size - 1) / (out - 1) if (align_corners and out > 1) else size / out
    return torch.arange(out, dtype=torch.float32, device=device) * scale


def resize_nearest(x: torch.Tensor, out_hw, align_corners: bool = False) -> torch.Tensor:
    """tf.image.resize_nearest_neighbor (TF1: source index floor(dst * in/out)) — vgg/nets/cyclegan.py:96."""
    oh, ow = int(out_hw[0]), int(out_hw[1])
    if use_native(x):
        from . import zoo_native
        return zoo_native.resize(x, (oh, ow), False, align_corners)
    sh = _resize_index(oh, x.shape[1], align_corners, x.device)
    sw = _resize_index(ow, x.shape[2], align_corners, x.device)
    ih = (torch.round(sh) if align_corners else torch.floor(sh)).long().clamp_(max=x.shape[1] - 1)
    iw = (torch.round(sw) if align_corners else torch.floor(sw)).long().clamp_(max=x.shape[2] - 1)
    return x[:, ih][:, :, iw]


def resize_bilinear(x: torch.Tensor, out_hw, align_corners: bool = False) -> torch.Tensor:
    """tf.image.resize_bilinear with TF1's sampling grid (src = dst * in/out, no half-pixel shift) —
    vgg/nets/cyclegan.py:101."""
    oh, ow = int(out_hw[0]), int(out_hw[1])
    if use_native(x):
        from . import zoo_native
        return zoo_native.resize(x, (oh, ow), True, align_corners)
    h, w = x.shape[1], x.shape[2]
    sh = _resize_index(oh, h, align_corners, x.device)
    sw = _resize_index(ow, w, align_corners, x.device)
    h0 = torch.floor(sh).long().clamp_(max=h - 1)
    w0 = torch.floor(sw).long().clamp_(max=w - 1)
    h1 = (h0 + 1).clamp_(max=h - 1)
    w1 = (w0 + 1).clamp_(max=w - 1)
    fh = (sh - h0).view(1, oh, 1, 1)
    fw = (sw - w0).view(1, 1, ow, 1)
    top = x[:, h0][:, :, w0] + (x[:, h0][:, :, w1] - x[:, h0][:, :, w0]) * fw
    bot = x[:, h1][:, :, w0] + (x[:, h1][:, :, w1] - x[:, h1][:, :, w0]) * fw
    return top + (bot - top) * fh


def pad_spatial(x: torch.Tensor, pads, mode: str = "constant") -> torch.Tensor:
    """tf.pad on the two spatial axes of NHWC; ``pads = (top, bottom, left, right)``, mode 'constant' (zeros)
    or 'reflect' (vgg/nets/cyclegan.py:201, resnet fixed padding)."""
    pt, pb, pl, pr = (int(v) for v in pads)
    if use_native(x):
        from . import zoo_native
        return zoo_native.pad(x, (pt, pb, pl, pr), mode)
    return F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb), mode=mode).permute(0, 2, 3, 1)


def drop_path(x: torch.Tensor, keep_prob: float, training: bool, generator=None) -> torch.Tensor:
    """NASNet's drop-path: zero a whole example's branch with probability 1-keep_prob and rescale the rest
    (vgg/nets/nasnet/nasnet_utils.py:153-163)."""
    if not training or keep_prob >= 1.0:
        return x
    mask = torch.floor(keep_prob + torch.rand((x.shape[0], 1, 1, 1), device=x.device, generator=generator))
    return x * (mask / keep_prob)


def flatten(x: torch.Tensor) -> torch.Tensor:
    """[N, ...] -> [N, prod] in NHWC order, the order TF's reshape produces (cnn/cifar10.py:255)."""
    return x.reshape(x.shape[0], -1)


def concat_channels(xs: Sequence[torch.Tensor]) -> torch.Tensor:
    """tf.concat(axis=3) (slim/inception_model.py:127)."""
    return torch.cat(list(xs), dim=3)


def one_hot(labels: torch.Tensor, num_classes: int) -> torch.Tensor:
    return F.one_hot(labels.long(), num_classes).to(torch.float32)


# ----------------------------------------------------------------------------------------------
# losses / metrics
# ----------------------------------------------------------------------------------------------
def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor, label_smoothing: float = 0.0,
                          weight: float = 1.0) -> torch.Tensor:
    """weight * mean_n( -sum_c y'[n,c] log softmax(logits)[n,c] ),  y' = y(1-eps) + eps/C.

    ``labels`` are class indices [N] (sparse form, cnn/cifar10.py:299) or dense targets [N, C]
    (tf.losses.softmax_cross_entropy; slim/losses.py:142-174 for smoothing and weight)."""
    if use_native(logits) and labels.dim() == 1:
        from . import native_ops
        return native_ops.softmax_xent(logits, labels, label_smoothing, weight)
    logp = F.log_softmax(logits.float(), dim=1)
    c = logits.shape[1]
    if labels.dim() == 1:
        target = F.one_hot(labels.long(), c).to(logp.dtype)
    else:
        target = labels.to(logp.dtype)
    if label_smoothing > 0:
        target = target * (1.0 - label_smoothing) + label_smoothing / c
    return -(target * logp).sum(1).mean() * weight


def in_top_k(logits: torch.Tensor, labels: torch.Tensor, k: int) -> torch.Tensor:
    """tf.nn.in_top_k (cnn/cifar10_eval.py:129): target's score is among the k largest
    (ties resolved as TF does: counts entries strictly greater than the target's score)."""
    tgt = logits.gather(1, labels.long().view(-1, 1))
    return (logits > tgt).sum(1) < k


def zero_fraction(x: torch.Tensor) -> torch.Tensor:
    """tf.nn.zero_fraction (cnn/cifar10.py:121)."""
    return (x == 0).float().mean()


def variance_scaling_std(fan_in: int) -> float:
    """tf.variance_scaling_initializer() default: truncated normal, scale=1, fan_in
    (stddev = sqrt(1/fan_in) / 0.87962566 after truncation correction)."""
    return math.sqrt(1.0 / max(1, fan_in)) / 0.87962566103423978
