"""Autograd wiring of the zoo-only kernels (``csrc/zoo_ops.cu``) and of the transposed convolution, which runs
on the tcgen05 input-gradient kernel of ``csrc/conv_igemm.cu`` / ``conv_wint.cu``.

``ops.functional`` dispatches here for CUDA tensors; every function mirrors its functional counterpart
(NHWC fp32).  Op kinds served ("extra op kinds that appear only here"): depthwise / separable
convolution, ``conv2d_transpose``, ``instance_norm``, ``relu6``, ``leaky_relu``, ``tanh``, ``sigmoid``,
nearest / bilinear resize, zero / reflect padding.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from .. import _native as nat
from . import conv_native as cn

_vp = C.c_void_p


def _p(t: Optional[torch.Tensor]):
    return _vp(t.data_ptr()) if t is not None else None


class DwConvParams(C.Structure):
    """Mirror of ``dtf_dwconv_params``."""
    _fields_ = [(n, C.c_int) for n in ("N", "H", "W", "C", "mult", "R", "S", "sh", "sw", "ph", "pw", "OH", "OW")]


# ---------------------------------------------------------------------------------------------------
# depthwise convolution
# ---------------------------------------------------------------------------------------------------
class _DwConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, out_hw):
        x = x.contiguous()
        w = w.contiguous()
        n, h, wd, c = x.shape
        r, s, c2, mult = w.shape
        assert c2 == c, "depthwise filter is [R, S, C, multiplier]"
        p = DwConvParams(n, h, wd, c, mult, r, s, stride[0], stride[1], pad[0], pad[1], out_hw[0], out_hw[1])
        y = torch.empty((n, out_hw[0], out_hw[1], c * mult), dtype=torch.float32, device=x.device)
        nat.check(nat.lib().dtf_dwconv_fwd(C.byref(p), _p(x), _p(w), _p(bias), _p(y), nat.stream_ptr()),
                  "dtf_dwconv_fwd")
        ctx.save_for_backward(x, w)
        ctx.p = p
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        p = ctx.p
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            nat.check(nat.lib().dtf_dwconv_dgrad(C.byref(p), _p(dy), _p(w), _p(dx), nat.stream_ptr()),
                      "dtf_dwconv_dgrad")
        if ctx.needs_input_grad[1]:
            dw = torch.zeros_like(w)
            nat.check(nat.lib().dtf_dwconv_wgrad(C.byref(p), _p(x), _p(dy), _p(dw), nat.stream_ptr()),
                      "dtf_dwconv_wgrad")
        if ctx.has_bias and ctx.needs_input_grad[2]:
            co = dy.shape[-1]
            db = torch.zeros(co, dtype=torch.float32, device=dy.device)
            nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(db), nat.stream_ptr()),
                      "dtf_colsum")
        return dx, dw, db, None, None, None


def depthwise_conv2d(x, w, bias, stride, pad_before, out_hw):
    return _DwConvFn.apply(x, w, bias, tuple(stride), tuple(pad_before), tuple(out_hw))


# ---------------------------------------------------------------------------------------------------
# activations
# ---------------------------------------------------------------------------------------------------
ACT_MODES = {"relu6": 0, "leaky_relu": 1, "tanh": 2, "sigmoid": 3}


class _ActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mode, alpha):
        x = x.contiguous()
        y = torch.empty_like(x)
        nat.check(nat.lib().dtf_act_fwd(_p(x), _p(y), C.c_longlong(x.numel()), mode, C.c_float(alpha),
                                        nat.stream_ptr()), "dtf_act_fwd")
        ctx.save_for_backward(x if mode < 2 else y)
        ctx.cfg = (mode, alpha)
        return y

    @staticmethod
    def backward(ctx, dy):
        (ref,) = ctx.saved_tensors
        mode, alpha = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        nat.check(nat.lib().dtf_act_bwd(_p(ref), _p(dy), _p(dx), C.c_longlong(dy.numel()), mode, C.c_float(alpha),
                                        nat.stream_ptr()), "dtf_act_bwd")
        return dx, None, None


def activation(x, kind: str, alpha: float = 0.2):
    return _ActFn.apply(x, ACT_MODES[kind], float(alpha))


# ---------------------------------------------------------------------------------------------------
# fake quantisation (--quantize)
# ---------------------------------------------------------------------------------------------------
class _FakeQuantFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, rng, gate, bits, narrow):
        x = x.contiguous()
        y = torch.empty_like(x)
        nat.check(nat.lib().dtf_fake_quant_fwd(_p(x), _p(y), C.c_longlong(x.numel()), _p(rng), _p(gate), int(bits),
                                               int(bool(narrow)), nat.stream_ptr()), "dtf_fake_quant_fwd")
        ctx.save_for_backward(x, rng, gate)
        ctx.cfg = (int(bits), int(bool(narrow)))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, rng, gate = ctx.saved_tensors
        bits, narrow = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        nat.check(nat.lib().dtf_fake_quant_bwd(_p(x), _p(dy), _p(dx), C.c_longlong(dy.numel()), _p(rng), _p(gate),
                                               bits, narrow, nat.stream_ptr()), "dtf_fake_quant_bwd")
        return dx, None, None, None, None


def fake_quant(x, rng, gate, bits: int, narrow: bool):
    return _FakeQuantFn.apply(x, rng, gate, bits, narrow)


def quant_range_update(x, rng, decay: float, mm) -> None:
    """Fold min / max of ``x`` into ``rng`` (two floats) in place; ``decay < 0`` assigns, else moving average.
    ``mm``: two floats of scratch holding {+inf, -inf}; the call hands it back in that state."""
    x = x.detach().contiguous()
    nat.check(nat.lib().dtf_quant_range_update(_p(x), C.c_longlong(x.numel()), _p(mm), _p(rng), C.c_float(decay),
                                               nat.stream_ptr()), "dtf_quant_range_update")


# ---------------------------------------------------------------------------------------------------
# instance norm
# ---------------------------------------------------------------------------------------------------
class _InstNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        x = x.contiguous()
        n, h, w, c = x.shape
        y = torch.empty_like(x)
        sums = torch.empty((n, 2, c), dtype=torch.float32, device=x.device)
        nat.check(nat.lib().dtf_instnorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(sums), n, h * w, c,
                                             C.c_float(eps), nat.stream_ptr()), "dtf_instnorm_fwd")
        ctx.save_for_backward(x, gamma, sums)
        ctx.eps = eps
        ctx.has = (gamma is not None, beta is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, sums = ctx.saved_tensors
        n, h, w, c = x.shape
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        bsums = torch.empty((n, 2, c), dtype=torch.float32, device=x.device)
        nat.check(nat.lib().dtf_instnorm_bwd(_p(x), _p(dy), _p(gamma), _p(sums), _p(bsums), _p(dx), n, h * w, c,
                                             C.c_float(ctx.eps), nat.stream_ptr()), "dtf_instnorm_bwd")
        red = bsums.sum(0) if (ctx.has[0] or ctx.has[1]) else None
        dgamma = red[1] if ctx.has[0] else None
        dbeta = red[0] if ctx.has[1] else None
        return dx, dgamma, dbeta, None


def instance_norm(x, gamma, beta, eps):
    return _InstNormFn.apply(x, gamma, beta, float(eps))


# ---------------------------------------------------------------------------------------------------
# resize / pad
# ---------------------------------------------------------------------------------------------------
class _ResizeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, out_hw, bilinear, align):
        x = x.contiguous()
        n, h, w, c = x.shape
        y = torch.empty((n, out_hw[0], out_hw[1], c), dtype=torch.float32, device=x.device)
        nat.check(nat.lib().dtf_resize_fwd(_p(x), _p(y), n, h, w, c, out_hw[0], out_hw[1], int(bilinear), int(align),
                                           nat.stream_ptr()), "dtf_resize_fwd")
        ctx.cfg = (x.shape, out_hw, bilinear, align)
        return y

    @staticmethod
    def backward(ctx, dy):
        (n, h, w, c), out_hw, bilinear, align = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty((n, h, w, c), dtype=torch.float32, device=dy.device)
        nat.check(nat.lib().dtf_resize_bwd(_p(dy), _p(dx), n, h, w, c, out_hw[0], out_hw[1], int(bilinear),
                                           int(align), nat.stream_ptr()), "dtf_resize_bwd")
        return dx, None, None, None


def resize(x, out_hw, bilinear: bool, align_corners: bool):
    return _ResizeFn.apply(x, (int(out_hw[0]), int(out_hw[1])), bool(bilinear), bool(align_corners))


class _PadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, pads, mode):
        x = x.contiguous()
        n, h, w, c = x.shape
        pt, pb, pl, pr = pads
        y = torch.empty((n, h + pt + pb, w + pl + pr, c), dtype=torch.float32, device=x.device)
        nat.check(nat.lib().dtf_pad_fwd(_p(x), _p(y), n, h, w, c, pt, pb, pl, pr, mode, nat.stream_ptr()),
                  "dtf_pad_fwd")
        ctx.cfg = (x.shape, pads, mode)
        return y

    @staticmethod
    def backward(ctx, dy):
        (n, h, w, c), (pt, pb, pl, pr), mode = ctx.cfg
        if mode == 0:
            return dy[:, pt:pt + h, pl:pl + w, :], None, None
        dy = dy.contiguous()
        dx = torch.empty((n, h, w, c), dtype=torch.float32, device=dy.device)
        nat.check(nat.lib().dtf_pad_bwd(_p(dy), _p(dx), n, h, w, c, pt, pb, pl, pr, mode, nat.stream_ptr()),
                  "dtf_pad_bwd")
        return dx, None, None


def pad(x, pads: Tuple[int, int, int, int], mode: str):
    return _PadFn.apply(x, tuple(int(v) for v in pads), {"constant": 0, "reflect": 1}[mode])


# ---------------------------------------------------------------------------------------------------
# transposed convolution = the input-gradient kernel run forwards
# ---------------------------------------------------------------------------------------------------
class _ConvTransposeFn(torch.autograd.Function):
    """y = conv2d_transpose(x, w).  ``w`` is stored ``[C_out, R, S, C_in]`` — exactly the operand layout the
    dgrad GEMM wants (``conv_native.conv_dgrad``'s ``w_t``), so the forward pass needs no weight shuffle.  The
    input gradient is a plain forward convolution with the weights transposed back to ``[C_in, R, S, C_out]``; the
    weight gradient is the ordinary conv weight gradient with the roles of activation and gradient exchanged."""

    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, out_hw):
        x = x.contiguous()
        y = cn.conv_dgrad(x, w, out_hw, stride, pad)
        if bias is not None:
            y = y + bias
        ctx.save_for_backward(x, w)
        ctx.cfg = (stride, pad, bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        stride, pad, has_bias = ctx.cfg
        dy = dy.contiguous()
        co, r, s, ci = w.shape
        dx = dw = db = None
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
            dyp, cop = dy, co
            if co % 4 != 0:                    # the fprop gather reads 16-byte channel chunks
                dyp = torch.nn.functional.pad(dy, (0, (-co) % 4))
                cop = dyp.shape[-1]
        if ctx.needs_input_grad[0]:
            w_fwd = cn.weight_transpose(w)      # [C_in, R, S, C_out]: OHWI of the convolution being transposed
            if cop != co:
                w_fwd = torch.nn.functional.pad(w_fwd, (0, cop - co)).contiguous()
            dx = cn.conv_fprop(dyp, w_fwd, stride, pad, out_hw=(x.shape[1], x.shape[2]))
        if ctx.needs_input_grad[1]:
            dwf = cn.conv_wgrad(dyp, x, (ci, r, s, cop), stride, pad)       # [C_in, R, S, C_out(+pad)]
            dw = dwf[..., :co].permute(3, 1, 2, 0).contiguous()
        if has_bias and ctx.needs_input_grad[2]:
            db = torch.zeros(co, dtype=torch.float32, device=dy.device)
            nat.check(nat.lib().dtf_colsum(_p(dy), C.c_longlong(dy.numel() // co), co, co, _p(db), nat.stream_ptr()),
                      "dtf_colsum")
        return dx, dw, db, None, None, None


def conv2d_transpose(x, w, bias, stride, pad_before, out_hw):
    return _ConvTransposeFn.apply(x, w, bias, tuple(stride), tuple(pad_before), tuple(out_hw))
