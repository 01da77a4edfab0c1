"""Autograd wiring of the native sm_100a kernels (``libdtfb200.so``).

Each public function mirrors one entry of ``ops.functional`` (same arguments, NHWC fp32) and is what that
module dispatches to for CUDA tensors.  Forward and backward both run hand-written kernels:

=====================  ==========================================================================
conv2d / linear        tcgen05 implicit GEMM: fprop, dgrad (same kernel, transposed index map), wgrad
batch_norm             stats (conv epilogue or ``bn_stats``) -> ``bn_finalize`` -> ``affine_act``;
                       backward = ``bn_bwd_reduce`` + ``bn_bwd_apply``
bn_relu_conv           BN-apply + ReLU fused into the conv's operand gather (ResNet pre-activation units)
conv_bn_act            conv with statistics epilogue + BN (Inception's conv->BN->ReLU triple)
max_pool / avg_pool    NHWC vector kernels, argmax byte map for the max-pool gradient
lrn / dropout          channel-window kernel / counter-based mask regenerated in backward
softmax_xent           loss and d(logits) in one pass
fused_update           SGD / TF1-RMSProp / weight decay / EMA on a flat shard
=====================  ==========================================================================

Everything launches on the current torch stream and allocates through the torch caching allocator, so whole
steps are CUDA-graph capturable.  Small per-channel scratch (BN sums) is carved from a :class:`ScratchArena`
that is zeroed once per step.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from .. import _native as nat
from . import conv_native as cn

_vp = C.c_void_p


def _p(t: Optional[torch.Tensor]):
    return _vp(t.data_ptr()) if t is not None else None


class PoolParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("N", "H", "W", "C", "OH", "OW", "kh", "kw", "sh", "sw", "ph", "pw",
                                       "ld_in", "ld_out")]


# ---------------------------------------------------------------------------------------------------
# scratch arena for per-channel accumulators (zeroed once per step, graph friendly)
# ---------------------------------------------------------------------------------------------------
class ScratchArena:
    def __init__(self, device, nfloats: int = 1 << 20):
        self.buf = torch.zeros(nfloats, dtype=torch.float32, device=device)
        self.off = 0

    def reset(self) -> None:
        self.buf.zero_()
        self.off = 0

    def take(self, n: int) -> torch.Tensor:
        n4 = (n + 3) // 4 * 4
        if self.off + n4 > self.buf.numel():
            return torch.zeros(n, dtype=torch.float32, device=self.buf.device)
        t = self.buf[self.off:self.off + n]
        self.off += n4
        return t


_ARENAS = {}


def _dev_index(device) -> int:
    return device.index if device.index is not None else torch.cuda.current_device()


def arena(device) -> ScratchArena:
    key = _dev_index(device)
    a = _ARENAS.get(key)
    if a is None:
        a = _ARENAS[key] = ScratchArena(device)
    return a


def begin_step(device) -> None:
    """Zero the per-step scratch (call once before forward)."""
    arena(device).reset()


def _zeros(n: int, device) -> torch.Tensor:
    return arena(device).take(n)


# ---------------------------------------------------------------------------------------------------
# convolution / linear
# ---------------------------------------------------------------------------------------------------
def _pad4(x: torch.Tensor, w: torch.Tensor):
    """The gather reads 16-byte channel chunks: pad C_in to a multiple of 4 (3-channel stems)."""
    c = x.shape[-1]
    if c % 4 == 0:
        return x, w
    padc = (-c) % 4
    return torch.nn.functional.pad(x, (0, padc)), torch.nn.functional.pad(w, (0, padc)).contiguous()


one_tap_enabled = [True]      # tests switch the single-live-tap shortcut off to compare against the full sweep


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, out_hw, relu, stats):
        x = cn.as_pixel_dense(x)
        y = cn.conv_fprop(x, w, stride, pad, bias=bias, relu=relu, stats=stats, out_hw=out_hw)
        if ctx.needs_input_grad[0]:
            cn.wt_register(w)          # its dgrad operand comes from the step's batched transpose
        ctx.save_for_backward(x, w, y if relu else None)
        ctx.cfg = (stride, pad, relu, bias is not None)
        ctx.params = (w, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        stride, pad, relu, has_bias = ctx.cfg
        dy = dy.contiguous()
        if relu:
            dy = torch.ops.aten.threshold_backward(dy, y, 0.0)
        dx = dw = db = None
        wparam, bparam = ctx.params
        wrote = []
        if has_bias and ctx.needs_input_grad[2]:
            co = dy.shape[-1]
            slot = _grad_slot(bparam)
            if slot is not None:
                with _wgrad_stream(dy):        # a leaf of the backward pass like the weight gradient: off the chain
                    nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(slot),
                                                   nat.stream_ptr()), "dtf_colsum")
                    _grad_written(bparam)
            else:
                db = _zeros(co, dy.device)
                nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(db),
                                               nat.stream_ptr()), "dtf_colsum")
        if ctx.needs_input_grad[1]:
            slot = _grad_slot(wparam)
            if slot is not None:
                with _wgrad_stream(x, dy):
                    cn.conv_wgrad(x, dy, w.shape, stride, pad, out=slot)  # red.add straight into the flat buffer
                    _grad_written(wparam)
            else:
                dw = cn.conv_wgrad(x, dy, w.shape, stride, pad)
        if ctx.needs_input_grad[0]:
            dx = cn.conv_dgrad(dy, cn.wt_get(w), (x.shape[1], x.shape[2]), stride, pad)
        _grad_written(*wrote)
        return dx, dw, db, None, None, None, None, None


class _OneTapFn(torch.autograd.Function):
    """A k x k convolution whose input AND output are 1 x 1 maps: exactly one filter tap ever meets data, all the
    others multiply padding (AlexNet's fc6 = 5x5 SAME and VGG's fc6 = 7x7 'same' on the 1x1 map that CIFAR images
    leave, alexnet/alexnet.py:119, vgg/nets/vgg.py:202-203 — 24/25 and 48/49 of those tensors).  The full tensor stays
    a variable (it is pushed, pulled, decayed and checkpointed like any other: SURVEY §7.4 item 10), but the compute
    runs as the [C_out, C_in] GEMM of the live tap: forward, input gradient and weight gradient touch 1/(R*S) of the
    weights instead of sweeping 400 MB three times per step; the dead taps' gradient is the zero the buffer holds."""

    @staticmethod
    def forward(ctx, x, w, bias, tap, relu):
        co, r, s, ci = w.shape
        nat.run_pre_launch()                      # the slice below reads w with a PyTorch kernel: pull gate first
        wc = w.detach()[:, tap[0], tap[1], :].contiguous().view(co, 1, 1, ci)
        x = cn.as_pixel_dense(x)
        y = cn.conv_fprop(x, wc, (1, 1), (0, 0), bias=bias, relu=relu, out_hw=(1, 1))
        ctx.save_for_backward(x, wc, y if relu else None)
        ctx.cfg = (tap, relu, bias is not None)
        ctx.params = (w, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, wc, y = ctx.saved_tensors
        tap, relu, has_bias = ctx.cfg
        wparam, bparam = ctx.params
        dy = dy.contiguous()
        if relu:
            dy = torch.ops.aten.threshold_backward(dy, y, 0.0)
        dx = dw = db = None
        wrote = []
        if has_bias and ctx.needs_input_grad[2]:
            co = dy.shape[-1]
            slot = _grad_slot(bparam)
            if slot is not None:
                with _wgrad_stream(dy):        # a leaf of the backward pass like the weight gradient: off the chain
                    nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(slot),
                                                   nat.stream_ptr()), "dtf_colsum")
                    _grad_written(bparam)
            else:
                db = _zeros(co, dy.device)
                nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(db),
                                               nat.stream_ptr()), "dtf_colsum")
        if ctx.needs_input_grad[0]:
            dx = cn.conv_dgrad(dy, cn.weight_transpose(wc), (1, 1), (1, 1), (0, 0))
        if ctx.needs_input_grad[1]:
            dwc = cn.conv_wgrad(x, dy, wc.shape, (1, 1), (0, 0)).view(wc.shape[0], wc.shape[3])
            slot = _grad_slot(wparam)
            if slot is not None:
                slot[:, tap[0], tap[1], :].add_(dwc)
                wrote.append(wparam)
            else:
                dw = torch.zeros_like(wparam)
                dw[:, tap[0], tap[1], :] = dwc
        _grad_written(*wrote)
        return dx, dw, db, None, None


def conv2d(x, w, bias, stride, pad_before, out_hw, relu, stats=None):
    cin = x.shape[-1]
    co, r, s, _ = w.shape
    if (r * s > 1 and x.shape[1] == 1 and x.shape[2] == 1 and tuple(out_hw) == (1, 1) and stats is None
            and cin % 4 == 0 and 0 <= pad_before[0] < r and 0 <= pad_before[1] < s and one_tap_enabled[0]):
        return _OneTapFn.apply(x, w, bias, (int(pad_before[0]), int(pad_before[1])), bool(relu))
    if cin % 4 != 0:
        nat.run_pre_launch()      # the pad below reads w with a PyTorch kernel: a pull gate still owed goes first
        x, w = _pad4(x, w)        # differentiable pads: the weight gradient is sliced back by autograd
    return _ConvFn.apply(x, w, bias, tuple(stride), tuple(pad_before), tuple(out_hw), bool(relu), stats)


def linear(x, w, bias, relu):
    n, k = x.shape
    y = conv2d(x.reshape(n, 1, 1, k), w.reshape(w.shape[0], 1, 1, k), bias, (1, 1), (0, 0), (1, 1), relu)
    return y.reshape(n, w.shape[0])


# ---------------------------------------------------------------------------------------------------
# batch norm
# ---------------------------------------------------------------------------------------------------
class _BnTrainSpec:
    """Arguments of a training-mode BN whose statistics are still to be computed."""

    def __init__(self, mm, mv, eps, decay, update_moving, unbiased):
        self.mm, self.mv, self.eps, self.decay = mm, mv, eps, decay
        self.update_moving, self.unbiased = update_moving, unbiased


class BnCoef:
    """Per-channel (scale, shift, mean, invstd) of one batch-norm application."""

    def __init__(self, scale, shift, mean, invstd, training):
        self.scale, self.shift, self.mean, self.invstd, self.training = scale, shift, mean, invstd, training
        self.pending = None      # deferred finalize: (stats, gamma, beta, mm, mv, count, eps, decay, update, unbiased)

    def materialize(self):
        """Run the stand-alone finalize kernel if the statistics have not been turned into coefficients yet."""
        if self.pending is not None:
            stats, gamma, beta, mm, mv, count, eps, decay, update_moving, unbiased = self.pending
            self.pending = None
            nat.check(nat.lib().dtf_bn_finalize(_p(stats), _p(gamma), _p(beta), C.c_longlong(count), C.c_float(eps),
                                                C.c_float(decay), int(update_moving), int(unbiased), _p(mm), _p(mv),
                                                _p(self.scale), _p(self.shift), _p(self.mean), _p(self.invstd),
                                                mm.numel(), nat.stream_ptr()), "dtf_bn_finalize")
        return self


@torch.no_grad()
def bn_prepare(stats: Optional[torch.Tensor], gamma, beta, mm, mv, count: int, eps: float, decay: float,
               training: bool, update_moving: bool, unbiased: bool, lazy: bool = False) -> BnCoef:
    """``lazy``: leave the finalize to the first consumer — a sliding-window convolution derives the coefficients
    in its prologue (``bn_relu_conv``); anything else calls ``coef.materialize()`` first."""
    c = mm.numel()
    dev = mm.device
    buf = torch.empty(4 * c, dtype=torch.float32, device=dev)
    scale, shift, mean, invstd = buf[:c], buf[c:2 * c], buf[2 * c:3 * c], buf[3 * c:]
    if training and lazy:
        coef = BnCoef(scale, shift, mean, invstd, training)
        coef.pending = (stats, gamma, beta, mm, mv, count, eps, decay, update_moving, unbiased)
        return coef
    if training:
        nat.check(nat.lib().dtf_bn_finalize(_p(stats), _p(gamma), _p(beta), C.c_longlong(count), C.c_float(eps),
                                            C.c_float(decay), int(update_moving), int(unbiased), _p(mm), _p(mv),
                                            _p(scale), _p(shift), _p(mean), _p(invstd), c, nat.stream_ptr()),
                  "dtf_bn_finalize")
    else:
        nat.check(nat.lib().dtf_bn_eval_affine(_p(gamma), _p(beta), _p(mm), _p(mv), C.c_float(eps), _p(scale),
                                               _p(shift), c, nat.stream_ptr()), "dtf_bn_eval_affine")
    return BnCoef(scale, shift, mean, invstd, training)


def bn_stats(x: torch.Tensor) -> torch.Tensor:
    c = x.shape[-1]
    st = _zeros(2 * c, x.device)
    m = x.numel() // c
    nat.check(nat.lib().dtf_bn_stats(_p(x), C.c_longlong(m), c, c, _p(st), nat.stream_ptr()), "dtf_bn_stats")
    return st


# Weight gradients are leaves of the backward pass: nothing downstream waits for them except the push.  They are
# issued on a side stream, so that the chain  dgrad -> batch-norm backward -> dgrad ...
# of the layers below does not queue behind them (ResNet-32 74.1k -> 78.2k, Inception-v3 2,393 -> 2,541, VGG-16 20.2k ->
# 22.8k img/s on one B200); DTF_WGRAD_STREAM=0 puts them back in line.
import contextlib as _contextlib
import os as _os

_wgrad_side = {"on": _os.environ.get("DTF_WGRAD_STREAM", "1") == "1", "streams": {}, "used": {}}


@_contextlib.contextmanager
def _wgrad_stream(*reads):
    """Run the enclosed launches on the device's weight-gradient stream, ordered after everything issued so far on
    the current stream; ``reads``: tensors the launches read (kept alive for that stream)."""
    if not (_wgrad_side["on"] and reads and reads[0].is_cuda):
        yield
        return
    dev = reads[0].device.index
    side = _wgrad_side["streams"].get(dev)
    if side is None:
        side = _wgrad_side["streams"][dev] = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream(reads[0].device)
    ev = torch.cuda.Event()
    ev.record(main)
    side.wait_event(ev)
    with torch.cuda.stream(side):
        yield
    for t in reads:
        if t is not None:
            t.record_stream(side)
    _wgrad_side["used"][dev] = side


def join_wgrad_stream(device) -> None:
    """The current stream waits for the weight gradients issued on the side stream (end of the backward pass)."""
    dev = device.index if device.index is not None else torch.cuda.current_device()
    side = _wgrad_side["used"].pop(dev, None)
    if side is not None:
        ev = torch.cuda.Event()
        ev.record(side)
        torch.cuda.current_stream(device).wait_event(ev)


_grad_listener = [None]


def set_grad_listener(fn) -> None:
    """``fn(param)`` is called right after a kernel that wrote ``param``'s gradient straight into the flat PS
    gradient buffer has been launched (gradients that travel through autograd are reported by the engine's
    post-accumulate hooks instead).  The bucketed push starts a bucket when all its tensors have reported."""
    _grad_listener[0] = fn


def _grad_written(*params) -> None:
    fn = _grad_listener[0]
    if fn is not None:
        for p in params:
            if p is not None:
                fn(p)


# K11 — concat elimination.  A block that concatenates branch outputs along the channels allocates the block's
# output once and hands every branch the channel slice it owns; the branch's LAST op (the batch-norm apply behind its
# final convolution, or a pooling kernel) takes the slice here and writes into it with the block's pixel pitch, so no
# concat kernel runs (the reference concatenates with tf.concat(axis=3, ...), inception/slim/inception_model.py:127).
_out_slot = [None]


def set_output_slot(view: Optional[torch.Tensor]) -> None:
    _out_slot[0] = view


def take_output_slot(shape) -> Optional[torch.Tensor]:
    s = _out_slot[0]
    if s is not None and tuple(s.shape) == tuple(shape):
        _out_slot[0] = None
        return s
    return None


def output_slot_pending() -> bool:
    return _out_slot[0] is not None


class _ConcatViewFn(torch.autograd.Function):
    """The block buffer whose channel slices the branches have already written: forward is free, backward hands each
    branch its slice of the incoming gradient as a strided view (the consumers read it with the buffer's pitch)."""

    @staticmethod
    def forward(ctx, buf, *outs):
        ctx.widths = [o.shape[-1] for o in outs]
        return buf.as_strided(buf.shape, buf.stride())

    @staticmethod
    def backward(ctx, dy):
        grads, c0 = [], 0
        for w in ctx.widths:
            grads.append(dy[..., c0:c0 + w])
            c0 += w
        return (None, *grads)


def concat_views(buf: torch.Tensor, outs) -> torch.Tensor:
    return _ConcatViewFn.apply(buf, *outs)


def _grad_slot(p):
    """The parameter's pre-allocated .grad (a view of the flat PS gradient buffer) when it can be accumulated
    into directly — saves one elementwise add kernel per parameter per step."""
    if p is None or not isinstance(p, torch.nn.Parameter):
        return None
    g = p.grad
    if g is None or not g.is_contiguous() or g.shape != p.shape:
        return None
    return g


def _bn_backward(dy, x, coef: BnCoef, relu: bool, has_gamma: bool, has_beta: bool, add=None, gamma=None,
                 beta=None, sums=None):
    """Shared BN(+ReLU) backward: returns (dx, dgamma, dbeta); dgamma/dbeta are None when they were accumulated
    straight into the parameters' .grad."""
    c = x.shape[-1]
    m = x.numel() // c
    dx = torch.empty_like(x)
    ld_dy = cn._pitch(dy) if dy.dim() == 4 else c      # dy may be a channel slice of a wider gradient (K11)
    if not coef.training:
        nat.check(nat.lib().dtf_affine_act_bwd(_p(dy), _p(x), _p(coef.scale), _p(coef.shift), int(relu),
                                               C.c_longlong(m), c, c, ld_dy, c, _p(dx), nat.stream_ptr()),
                  "dtf_affine_act_bwd")
        if add is not None:
            dx = dx + add
        # frozen statistics: dgamma/dbeta via the same reduction
    if sums is None:          # (the dgrad epilogue may already have produced the reductions)
        sums = _zeros(2 * c, x.device)
        nat.check(nat.lib().dtf_bn_bwd_reduce(_p(dy), _p(x), _p(coef.scale), _p(coef.shift), int(relu),
                                              C.c_longlong(m), c, c, ld_dy, _p(sums), nat.stream_ptr()),
                  "dtf_bn_bwd_reduce")
    gslot, bslot = _grad_slot(gamma), _grad_slot(beta)
    direct = coef.training and (not has_gamma or gslot is not None) and (not has_beta or bslot is not None)
    if direct:
        dg, db = gslot if has_gamma else None, bslot if has_beta else None
    else:
        dg = torch.zeros(c, dtype=torch.float32, device=x.device) if has_gamma else None
        db = torch.zeros(c, dtype=torch.float32, device=x.device) if has_beta else None
    if coef.training:
        nat.check(nat.lib().dtf_bn_bwd_apply(_p(dy), _p(x), _p(coef.scale), _p(coef.shift), _p(coef.mean),
                                             _p(coef.invstd), _p(sums), int(relu), C.c_longlong(m), c, c, ld_dy, c,
                                             _p(add), c if add is not None else 0, _p(dx), _p(dg), _p(db),
                                             nat.stream_ptr()), "dtf_bn_bwd_apply")
    else:
        if db is not None:
            db.copy_(sums[:c])
        if dg is not None:
            # d/dgamma of gamma*invstd*(x-mean): invstd * (sum g*x - mean * sum g), with moving statistics
            dg.copy_(coef.invstd_eval * (sums[c:] - coef.mean_eval * sums[:c]))
    if direct:
        _grad_written(gamma if has_gamma else None, beta if has_beta else None)
        return dx, None, None
    return dx, dg, db


class _BnApplyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, coef, relu):
        x = x.contiguous()
        c = x.shape[-1]
        m = x.numel() // c
        y = take_output_slot(x.shape)            # a channel slice of the enclosing block's output, or None
        if y is None:
            y = torch.empty_like(x)
        ld_y = cn._pitch(y)
        if isinstance(coef, _BnTrainSpec):
            # statistics, moving-average update and apply in one native call
            sp = coef
            buf = torch.empty(4 * c, dtype=torch.float32, device=x.device)
            scratch = torch.empty(2 * c, dtype=torch.float32, device=x.device)
            nat.check(nat.lib().dtf_bn_forward_train(_p(x), C.c_longlong(m), c, c, _p(gamma), _p(beta),
                                                     C.c_float(sp.eps), C.c_float(sp.decay), int(sp.update_moving),
                                                     int(sp.unbiased), _p(sp.mm), _p(sp.mv), int(relu), _p(scratch),
                                                     _p(buf), _p(y), ld_y, nat.stream_ptr()), "dtf_bn_forward_train")
            coef = BnCoef(buf[:c], buf[c:2 * c], buf[2 * c:3 * c], buf[3 * c:], True)
        else:
            nat.check(nat.lib().dtf_affine_act(_p(x), _p(coef.scale), _p(coef.shift), int(relu), C.c_longlong(m), c, c,
                                               ld_y, _p(y), nat.stream_ptr()), "dtf_affine_act")
        ctx.save_for_backward(x)
        ctx.coef, ctx.relu = coef, relu
        ctx.has = (gamma is not None, beta is not None)
        ctx.params = (gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dx, dg, db = _bn_backward(cn.as_pixel_dense(dy), x, ctx.coef, ctx.relu, ctx.has[0], ctx.has[1],
                                  gamma=ctx.params[0], beta=ctx.params[1])
        return dx, dg, db, None, None


def batch_norm(x, gamma, beta, mm, mv, training, decay, eps, relu, update_moving, unbiased, stats=None):
    """Stand-alone BN (+ReLU).  ``stats`` may carry sums already produced by a conv epilogue."""
    shape = x.shape
    c = shape[-1]
    x4 = x if x.dim() == 4 else x.reshape(shape[0], 1, 1, c)
    x4 = x4.contiguous()
    count = x4.numel() // c
    if training and stats is None:
        y = _BnApplyFn.apply(x4, gamma, beta, _BnTrainSpec(mm, mv, eps, decay, update_moving, unbiased), bool(relu))
        return y.reshape(shape)
    coef = bn_prepare(stats, gamma, beta, mm, mv, count, eps, decay, training, update_moving, unbiased)
    if not training:
        coef.mean_eval = mm
        coef.invstd_eval = torch.rsqrt(mv + eps)
    y = _BnApplyFn.apply(x4, gamma, beta, coef, bool(relu))
    return y.reshape(shape)


def conv_bn_act(x, w, stride, padding, gamma, beta, mm, mv, decay, eps, relu):
    """conv -> BN(train) -> ReLU with the statistics taken from the conv epilogue (one pass less over y)."""
    from .functional import resolve_padding
    co, r, s, _ = w.shape
    oh, ow, ph, pw = resolve_padding(x.shape[1], x.shape[2], r, s, stride, padding)
    stats = _zeros(2 * co, x.device)
    y = conv2d(x, w, None, stride, (ph[0], pw[0]), (oh, ow), False, stats=stats)
    return batch_norm(y, gamma, beta, mm, mv, True, decay, eps, relu, True, False, stats=stats)


# ---------------------------------------------------------------------------------------------------
# BN-ReLU fused into the consumer convolution (ResNet pre-activation units)
# ---------------------------------------------------------------------------------------------------
class _BnReluConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, w, residual, coef, stride, pad, out_hw, stats_out, res_link=None,
                link_first=False, link_last=False):
        # res_link: identity-shortcut units hand the shortcut's gradient from the unit's last convolution (which
        # received it as `residual`) to its first one, whose batch-norm backward adds it while writing dx —
        # instead of autograd summing the two contributions to x with a separate elementwise kernel.
        ctx.res_link = res_link
        ctx.link_first, ctx.link_last = link_first, link_last
        # the window kernels hand back relu(bn(x)) as a by-product: the weight gradient then copies it instead of
        # re-applying the transform to every im2col element
        y, act = cn.conv_fprop(x, w, stride, pad, residual=residual, in_scale=coef.scale, in_shift=coef.shift,
                               in_relu=True, stats=stats_out, out_hw=out_hw,
                               want_act=any(ctx.needs_input_grad), bn_pending=coef)
        if ctx.needs_input_grad[0]:
            cn.wt_register(w)
        ctx.save_for_backward(x, w)
        ctx.act = act
        ctx.coef = coef
        ctx.cfg = (stride, pad, residual is not None)
        ctx.params = (gamma, beta, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        stride, pad, has_res = ctx.cfg
        coef = ctx.coef
        gamma, beta, wparam = ctx.params
        dy = dy.contiguous()
        slot = _grad_slot(wparam)
        dw = None
        if ctx.act is not None:
            wsrc, waff = ctx.act, dict()
        else:
            wsrc, waff = x, dict(in_scale=coef.scale, in_shift=coef.shift, in_relu=True)
        if slot is not None:
            with _wgrad_stream(wsrc, dy):
                cn.conv_wgrad(wsrc, dy, w.shape, stride, pad, out=slot, **waff)
                _grad_written(wparam)
        else:
            dw = cn.conv_wgrad(wsrc, dy, w.shape, stride, pad, **waff)
        ctx.act = None
        sums = _zeros(2 * x.shape[-1], x.device)
        da = cn.conv_dgrad(dy, cn.wt_get(w), (x.shape[1], x.shape[2]), stride, pad,
                           bn_bwd=(x, coef.scale, coef.shift, sums))
        add = None
        d_res = dy if has_res else None
        if ctx.res_link is not None:
            if ctx.link_last:
                ctx.res_link["g"] = dy          # the shortcut is the unit's own input: pass its gradient on
                d_res = None
            if ctx.link_first:
                add = ctx.res_link.pop("g", None)
        dx, dg, db = _bn_backward(da, x, coef, True, True, True, add=add, gamma=gamma, beta=beta, sums=sums)
        return dx, dg, db, dw, d_res, None, None, None, None, None, None, None, None


def bn_relu_conv(x, gamma, beta, w, coef: BnCoef, stride, padding, residual=None, want_stats=True, res_link=None,
                 link_first=False, link_last=False):
    from .functional import resolve_padding
    co, r, s, _ = w.shape
    stride = (stride, stride) if isinstance(stride, int) else tuple(stride)
    oh, ow, ph, pw = resolve_padding(x.shape[1], x.shape[2], r, s, stride, padding)
    stats = _zeros(2 * co, x.device) if want_stats else None
    y = _BnReluConvFn.apply(x, gamma, beta, w, residual, coef, stride, (ph[0], pw[0]), (oh, ow), stats, res_link,
                            link_first, link_last)
    return y, stats


def resnet_unit(ctx, x_and_stats, convs, project, bn_vars, conv_var, fixed_padding, decay, eps):
    """One pre-activation residual unit on the fused path.  ``x_and_stats`` = (x, stats of x)."""
    x, st = x_and_stats
    store = ctx.store
    count = x.numel() // x.shape[-1]
    gamma, beta, mm, mv = bn_vars(store, x.shape[-1])
    # identity units: the only consumer of this batch-norm is the first convolution, which finalizes it itself
    coef = bn_prepare(st, gamma, beta, mm, mv, count, eps, decay, True, ctx.update_moving, True,
                      lazy=project is None)
    shortcut = x
    if project is not None:
        k, f, s = project
        wp = conv_var(store, x.shape[-1], k, f)
        shortcut, _ = bn_relu_conv(x, gamma, beta, wp, coef, s, fixed_padding(k, s), want_stats=False)
    y, yst = x, st
    n = len(convs)
    # identity shortcut: its gradient rides from the last convolution's backward to the first one's (see
    # _BnReluConvFn); needs autograd to be recording and the unit input to want a gradient
    link = {} if (project is None and torch.is_grad_enabled() and x.requires_grad) else None
    for i, (k, f, s) in enumerate(convs):
        if i > 0:
            gamma, beta, mm, mv = bn_vars(store, y.shape[-1])
            coef = bn_prepare(yst, gamma, beta, mm, mv, y.numel() // y.shape[-1], eps, decay, True,
                              ctx.update_moving, True, lazy=True)
        w = conv_var(store, y.shape[-1], k, f)
        last = i == n - 1
        y, yst = bn_relu_conv(y, gamma, beta, w, coef, s, fixed_padding(k, s),
                              residual=shortcut if last else None, want_stats=True, res_link=link,
                              link_first=(i == 0), link_last=last)
    return y, yst


# ---------------------------------------------------------------------------------------------------
# pooling
# ---------------------------------------------------------------------------------------------------
def _pool_params(x, k, stride, pad, out_hw, c_out_ld):
    p = PoolParams()
    n, h, w, c = x.shape
    p.N, p.H, p.W, p.C = n, h, w, c
    p.OH, p.OW = out_hw
    p.kh, p.kw = k
    p.sh, p.sw = stride
    p.ph, p.pw = pad
    p.ld_in, p.ld_out = c, c_out_ld
    return p


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad, out_hw):
        x = x.contiguous()
        n, h, w, c = x.shape
        y = take_output_slot((n, out_hw[0], out_hw[1], c))
        if y is None:
            y = torch.empty((n, out_hw[0], out_hw[1], c), dtype=torch.float32, device=x.device)
        arg = torch.empty((n, out_hw[0], out_hw[1], c), dtype=torch.uint8, device=x.device)
        p = _pool_params(x, k, stride, pad, out_hw, cn._pitch(y))
        nat.check(nat.lib().dtf_max_pool_fwd(C.byref(p), _p(x), _p(y), _p(arg), nat.stream_ptr()), "dtf_max_pool_fwd")
        ctx.save_for_backward(arg)
        ctx.p = p
        ctx.in_shape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        dy = cn.as_pixel_dense(dy)
        c = ctx.in_shape[-1]
        dx = torch.empty(ctx.in_shape, dtype=torch.float32, device=dy.device)
        nat.check(nat.lib().dtf_max_pool_bwd(C.byref(ctx.p), _p(dy), cn._pitch(dy), _p(arg), _p(dx), c, nat.stream_ptr()),
                  "dtf_max_pool_bwd")
        return dx, None, None, None, None


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad, out_hw):
        x = x.contiguous()
        n, h, w, c = x.shape
        y = take_output_slot((n, out_hw[0], out_hw[1], c))
        if y is None:
            y = torch.empty((n, out_hw[0], out_hw[1], c), dtype=torch.float32, device=x.device)
        p = _pool_params(x, k, stride, pad, out_hw, cn._pitch(y))
        nat.check(nat.lib().dtf_avg_pool_fwd(C.byref(p), _p(x), _p(y), nat.stream_ptr()), "dtf_avg_pool_fwd")
        ctx.p = p
        ctx.in_shape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = cn.as_pixel_dense(dy)
        c = ctx.in_shape[-1]
        dx = torch.empty(ctx.in_shape, dtype=torch.float32, device=dy.device)
        nat.check(nat.lib().dtf_avg_pool_bwd(C.byref(ctx.p), _p(dy), cn._pitch(dy), _p(dx), c, nat.stream_ptr()),
                  "dtf_avg_pool_bwd")
        return dx, None, None, None, None


def max_pool(x, k, stride, pad_before, out_hw):
    if x.shape[-1] % 4 != 0:
        raise RuntimeError("native max_pool needs C % 4 == 0")
    return _MaxPoolFn.apply(x, tuple(k), tuple(stride), tuple(pad_before), tuple(out_hw))


def avg_pool(x, k, stride, pad_before, out_hw):
    if x.shape[-1] % 4 != 0:
        raise RuntimeError("native avg_pool needs C % 4 == 0")
    return _AvgPoolFn.apply(x, tuple(k), tuple(stride), tuple(pad_before), tuple(out_hw))


# ---------------------------------------------------------------------------------------------------
# LRN / dropout / loss
# ---------------------------------------------------------------------------------------------------
class _LrnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, radius, bias, alpha, beta):
        x = x.contiguous()
        c = x.shape[-1]
        y = torch.empty_like(x)
        nat.check(nat.lib().dtf_lrn_fwd(_p(x), _p(y), C.c_longlong(x.numel() // c), c, radius, C.c_float(bias),
                                        C.c_float(alpha), C.c_float(beta), nat.stream_ptr()), "dtf_lrn_fwd")
        ctx.save_for_backward(x)
        ctx.cfg = (radius, bias, alpha, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        radius, bias, alpha, beta = ctx.cfg
        c = x.shape[-1]
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        nat.check(nat.lib().dtf_lrn_bwd(_p(x), _p(dy), _p(dx), C.c_longlong(x.numel() // c), c, radius,
                                        C.c_float(bias), C.c_float(alpha), C.c_float(beta), nat.stream_ptr()),
                  "dtf_lrn_bwd")
        return dx, None, None, None, None


def lrn(x, radius, bias, alpha, beta):
    return _LrnFn.apply(x, int(radius), float(bias), float(alpha), float(beta))


_SEEDS = {}
_SALT = [0]


def dropout_seed(device) -> torch.Tensor:
    """Device-resident step counter feeding the dropout generator (bumped by ``advance_dropout_seed``)."""
    key = _dev_index(device)
    t = _SEEDS.get(key)
    if t is None:
        t = _SEEDS[key] = torch.full((1,), 0x5DEECE66D, dtype=torch.int64, device=device)
    return t


def advance_dropout_seed(device) -> None:
    dropout_seed(device).add_(1)
    _SALT[0] = 0


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, keep, salt):
        x = x.contiguous()
        y = torch.empty_like(x)
        seed = dropout_seed(x.device)
        nat.check(nat.lib().dtf_dropout(_p(x), _p(y), C.c_longlong(x.numel()), C.c_float(keep), _p(seed),
                                        C.c_ulonglong(salt), nat.stream_ptr()), "dtf_dropout")
        ctx.cfg = (keep, salt)
        return y

    @staticmethod
    def backward(ctx, dy):
        keep, salt = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        seed = dropout_seed(dy.device)
        nat.check(nat.lib().dtf_dropout(_p(dy), _p(dx), C.c_longlong(dy.numel()), C.c_float(keep), _p(seed),
                                        C.c_ulonglong(salt), nat.stream_ptr()), "dtf_dropout")
        return dx, None, None


def dropout(x, keep_prob):
    _SALT[0] += 1
    return _DropoutFn.apply(x, float(keep_prob), _SALT[0] * 0x9E3779B1)


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, smoothing, weight):
        logits = logits.contiguous()
        n, c = logits.shape
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
        dlog = torch.empty_like(logits)
        nat.check(nat.lib().dtf_softmax_xent(_p(logits), _p(labels), n, c, C.c_float(smoothing), C.c_float(weight),
                                             _p(loss), _p(dlog), None, nat.stream_ptr()), "dtf_softmax_xent")
        ctx.save_for_backward(dlog)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, dloss):
        (dlog,) = ctx.saved_tensors
        return dlog * dloss, None, None, None


def softmax_xent(logits, labels, smoothing, weight):
    return _XentFn.apply(logits, labels.to(torch.int64).contiguous(), float(smoothing), float(weight))


# ---------------------------------------------------------------------------------------------------
# fused optimiser update
# ---------------------------------------------------------------------------------------------------
_HYPER_RING = 4


def hyper_tensor(opt, device) -> torch.Tensor:
    h = getattr(opt, "_hyper", None)
    if h is None:
        h = opt._hyper = torch.zeros(8, dtype=torch.float32, device=device)
        # a ring of pinned staging buffers: the host runs a step ahead of the device, so the buffer a pending
        # copy reads from must not be rewritten by the next step's values
        opt._hyper_ring = [(torch.zeros(8, dtype=torch.float32).pin_memory(), torch.cuda.Event())
                           for _ in range(_HYPER_RING)]
        opt._hyper_i = 0
    return h


def set_hyper(opt, lr: float, grad_scale: float, ema_decay: float) -> torch.Tensor:
    cfg = opt.cfg
    h = hyper_tensor(opt, opt.w.device)
    vals = (float(lr), float(grad_scale), float(cfg.weight_decay), float(cfg.rms_decay), float(cfg.momentum),
            float(cfg.epsilon), float(ema_decay), 0.0)
    if getattr(opt, "_hyper_vals", None) != vals:
        opt._hyper_vals = vals
        host, ev = opt._hyper_ring[opt._hyper_i % _HYPER_RING]
        opt._hyper_i += 1
        ev.synchronize()                   # the copy that last read this slot (4 changes ago) has long finished
        host.copy_(torch.tensor(vals, dtype=torch.float32))
        h.copy_(host, non_blocking=True)
        ev.record()
    return h


def fused_update(opt, g: torch.Tensor, lr: float, grad_scale: float, num_updates, use_ema: bool) -> None:
    from ..optim.schedule import ema_decay as _ema_decay
    d = -1.0
    if opt.ema is not None and use_ema:
        d = _ema_decay(opt.cfg.ema.decay, num_updates if opt.cfg.ema.use_num_updates else None)
    h = set_hyper(opt, lr, grad_scale, d)
    kind = 0 if opt.cfg.kind == "sgd" else 1
    n = opt.w.numel()
    nat.check(nat.lib().dtf_fused_update(_p(opt.w), _p(g), _p(opt.slots.get("RMSProp")), _p(opt.slots.get("RMSProp_1")),
                                         _p(opt.ema), C.c_longlong(n), C.c_longlong(opt.n_decayed), _p(h), kind,
                                         nat.stream_ptr()), "dtf_fused_update")
