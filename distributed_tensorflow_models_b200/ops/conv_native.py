"""Thin Python launchers for the tcgen05 implicit-GEMM kernels (``csrc/conv_igemm.cu``,
``csrc/conv_wgrad.cu``).  Tensors are NHWC fp32 (the last-dim stride must be 1; the pixel pitch may
exceed the channel count so a layer can read or write a channel slice of a wider buffer — this is how
Inception's concats are eliminated).  Weights are OHWI ``[C_out, R, S, C_in]``.

Autograd wiring lives in ``ops/functional.py``; this module only validates and launches.
"""
from __future__ import annotations

import contextlib
import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from .. import _native as nat


def _pitch(t: torch.Tensor) -> int:
    """Pixel pitch (floats) of an NHWC tensor / view whose pixels are laid out densely."""
    assert t.dim() == 4 and t.stride(3) == 1, "expected NHWC with unit channel stride"
    n, h, w, c = t.shape
    ld = t.stride(2) if w > 1 else (t.stride(1) if h > 1 else (t.stride(0) if n > 1 else c))
    if w > 1 and h > 1:
        assert t.stride(1) == ld * w, "rows must be densely packed"
    if n > 1:
        assert t.stride(0) == ld * w * h, "images must be densely packed"
    return ld


def as_pixel_dense(t: torch.Tensor) -> torch.Tensor:
    """``t`` itself when its pixels are laid out densely (a channel slice of a wider buffer is fine), else a packed
    copy — spatial slices such as ``x[:, 1:, 1:, :]`` have rows that are not back to back."""
    n, h, w, c = t.shape
    if t.stride(3) != 1:
        return t.contiguous()
    ld = t.stride(2) if w > 1 else (t.stride(1) if h > 1 else (t.stride(0) if n > 1 else c))
    if (w > 1 and h > 1 and t.stride(1) != ld * w) or (n > 1 and t.stride(0) != ld * w * h) or ld < c:
        return t.contiguous()
    return t


_gate_provider = [None]


def set_gate_provider(fn) -> None:
    """``fn(w) -> (flags, n, epoch, status, timeout_cycles) | None``: the parameter-server pull gate of the bucket
    that holds the weight tensor ``w`` (parallel/symm.py).  The forward convolution hands it to the kernel, whose
    weight-load warp waits on the flags before its first weight TMA (csrc/ps_gate.cuh)."""
    _gate_provider[0] = fn


def out_size(i: int, k: int, stride: int, pad: int) -> int:
    return (i + 2 * pad - k) // stride + 1


def _conv_fprop_once(x: torch.Tensor, w: torch.Tensor, stride: Tuple[int, int], pad: Tuple[int, int], *,
               bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
               relu: bool = False, in_scale: Optional[torch.Tensor] = None,
               in_shift: Optional[torch.Tensor] = None, in_relu: bool = False,
               stats: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
               out_hw: Optional[Tuple[int, int]] = None, want_act: bool = False, bn_pending=None):
    """y = epilogue(conv(act(x), w)); see ``dtf_conv_gemm`` in include/dtf_b200.h.

    ``pad`` is the padding *before* (top, left); with TensorFlow 'SAME' the padding after may be one
    larger, which is expressed by passing the output size explicitly in ``out_hw``.

    ``want_act``: also return ``act(x)`` (or ``None`` when the kernel chosen for this shape does not produce it) —
    the sliding-window kernels read every input pixel once and can store the transformed value on the way, which
    spares the weight-gradient pass the transform."""
    n, h, wd, c = x.shape
    co, r, s, ci = w.shape
    assert ci == c and w.is_contiguous() and x.dtype == torch.float32 and w.dtype == torch.float32
    if out_hw is None:
        oh, ow = out_size(h, r, stride[0], pad[0]), out_size(wd, s, stride[1], pad[1])
    else:
        oh, ow = out_hw
    if out is None:
        out = torch.empty((n, oh, ow, co), device=x.device, dtype=torch.float32)
    assert out.shape == (n, oh, ow, co)
    p = nat.ConvGemmParams()
    p.src, p.srcH, p.srcW, p.srcC, p.src_ld = x.data_ptr(), h, wd, c, _pitch(x)
    p.N, p.OH, p.OW = n, oh, ow
    p.R, p.S = r, s
    p.a_h, p.b_h, p.c_h, p.div_h = stride[0], -pad[0], 1, 1
    p.a_w, p.b_w, p.c_w, p.div_w = stride[1], -pad[1], 1, 1
    p.wt, p.Nout = w.data_ptr(), co
    p.out, p.out_ld = out.data_ptr(), _pitch(out)
    p.in_scale, p.in_shift, p.in_relu = nat.ptr(in_scale), nat.ptr(in_shift), int(in_relu)
    p.bias, p.residual = nat.ptr(bias), nat.ptr(residual)
    p.res_ld = _pitch(residual) if residual is not None else 0
    p.out_relu = int(relu)
    p.stats = nat.ptr(stats)
    p.bnb_x = p.bnb_scale = p.bnb_shift = None
    p.bnb_ld = 0
    p.act_out = None
    if _gate_provider[0] is not None:
        gate = _gate_provider[0](w)
        if gate is not None:
            p.gate_flags, p.gate_n, p.gate_epoch, p.gate_status, p.gate_timeout_cycles = gate
    if bn_pending is not None and getattr(bn_pending, "pending", None) is not None:
        # deferred batch-norm finalize: the sliding-window kernel derives scale / shift from the statistics itself
        # (and publishes them); any other kernel needs them materialised first
        if nat.lib().dtf_conv_gemm_fuses_finalize(C.byref(p)):
            st, g_, b_, mm, mv, count, eps, decay, upd, unbiased = bn_pending.pending
            bn_pending.pending = None
            p.fin_stats, p.fin_gamma, p.fin_beta = st.data_ptr(), nat.ptr(g_), nat.ptr(b_)
            p.fin_moving_mean, p.fin_moving_var = nat.ptr(mm), nat.ptr(mv)
            p.fin_mean, p.fin_invstd = bn_pending.mean.data_ptr(), bn_pending.invstd.data_ptr()
            p.fin_inv_count = 1.0 / float(count)
            p.fin_unbias = float(count) / float(count - 1) if (unbiased and count > 1) else 1.0
            p.fin_eps, p.fin_decay, p.fin_update_moving = float(eps), float(decay), int(upd)
        else:
            bn_pending.materialize()
    act = None
    if want_act and in_scale is not None and x.is_contiguous() and nat.lib().dtf_conv_gemm_writes_act(C.byref(p)):
        act = torch.empty_like(x)
        p.act_out = act.data_ptr()
    nat.check(nat.lib().dtf_conv_gemm(C.byref(p), nat.stream_ptr()), "dtf_conv_gemm(fprop)")
    return (out, act) if want_act else out


def _conv_dgrad_once(dy: torch.Tensor, w_t: torch.Tensor, in_hw: Tuple[int, int], stride: Tuple[int, int],
                     pad: Tuple[int, int], out: Optional[torch.Tensor] = None, bn_bwd=None,
                     residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = conv_transpose(dy, w).  ``w_t`` is the dgrad operand ``[C_in, R, S, C_out]``
    (``weight_transpose`` of the OHWI weights).

    ``bn_bwd = (x, scale, shift, sums)``: the layer's forward input was ``relu(x*scale+shift)``; the epilogue
    then also accumulates ``sum g`` / ``sum g*x`` (g = dx masked by the ReLU) into ``sums`` (2*C_in floats)."""
    n, p_, q_, co = dy.shape
    ci, r, s, co2 = w_t.shape
    assert co2 == co and w_t.is_contiguous()
    if (stride[0] > 1 or stride[1] > 1) and r * s > 1 and co in (16, 32, 64) and ci in (16, 32, 64):
        # Strided stencil: insert the zeros explicitly (dy at the even positions of a stride-times larger grid) and
        # run the stride-1 sliding-window kernel on that.  The generic path would gather every tap of every pixel
        # and zero-fill the 1 - 1/(sh*sw) of them that do not divide, element by element.
        dyu = torch.zeros((n, p_ * stride[0], q_ * stride[1], co), device=dy.device, dtype=torch.float32)
        dyu[:, ::stride[0], ::stride[1], :] = dy
        return _conv_dgrad_once(dyu, w_t, in_hw, (1, 1), pad, out=out, bn_bwd=bn_bwd, residual=residual)
    if co % 4 != 0:
        # the gather reads 16-byte channel chunks and TMA needs a 16-byte row pitch: zero-pad the
        # (tiny: logits layers only) gradient and weight operand to a multiple of 4 channels
        padc = (-co) % 4
        dy = torch.nn.functional.pad(dy, (0, padc))
        w_t = torch.nn.functional.pad(w_t, (0, padc)).contiguous()
        co += padc
    h, wd = in_hw
    if out is None:
        out = torch.empty((n, h, wd, ci), device=dy.device, dtype=torch.float32)
    p = nat.ConvGemmParams()
    p.src, p.srcH, p.srcW, p.srcC, p.src_ld = dy.data_ptr(), p_, q_, co, _pitch(dy)
    p.N, p.OH, p.OW = n, h, wd
    p.R, p.S = r, s
    p.a_h, p.b_h, p.c_h, p.div_h = 1, pad[0], -1, stride[0]
    p.a_w, p.b_w, p.c_w, p.div_w = 1, pad[1], -1, stride[1]
    p.wt, p.Nout = w_t.data_ptr(), ci
    p.out, p.out_ld = out.data_ptr(), _pitch(out)
    p.in_scale = p.in_shift = None
    p.in_relu = 0
    p.bias = None
    p.residual = nat.ptr(residual)
    p.res_ld = _pitch(residual) if residual is not None else 0
    p.out_relu = 0
    p.stats = None
    p.bnb_x = p.bnb_scale = p.bnb_shift = None
    p.bnb_ld = 0
    p.act_out = None
    if bn_bwd is not None:
        bx, bsc, bsh, sums = bn_bwd
        p.bnb_x, p.bnb_ld = bx.data_ptr(), _pitch(bx)
        p.bnb_scale, p.bnb_shift = bsc.data_ptr(), bsh.data_ptr()
        p.stats = sums.data_ptr()
    nat.check(nat.lib().dtf_conv_gemm(C.byref(p), nat.stream_ptr()), "dtf_conv_gemm(dgrad)")
    return out


class WeightTransposeBatch:
    """dgrad operands ``[C_in,R,S,C_out]`` of every convolution of a step, produced by ONE kernel launch.

    Convolutions register their weights during the forward pass; the first backward convolution that asks for a
    transposed weight triggers the batched launch for all of them.  The record table lives on the device and is
    rebuilt only when the set of registered weights changes, so a captured CUDA graph replays a single launch."""

    def __init__(self):
        self.entries = {}          # data_ptr -> (shape, index)
        self.order = []            # (data_ptr, shape)
        self.table = None
        self.out = None
        self.views = {}
        self.total_blocks = 0
        self.fresh = False         # transposes valid for the current step
        self.dirty = True

    def register(self, w: torch.Tensor) -> None:
        key = (w.data_ptr(), tuple(w.shape))
        if key not in self.entries:
            self.entries[key] = len(self.order)
            self.order.append((key, w.device))
            self.dirty = True
        self.fresh = False         # a forward pass is running: weights may have changed since the last backward

    def _rebuild(self) -> None:
        import struct
        dev = self.order[0][1]
        total = sum(s[0] * s[1] * s[2] * s[3] for (_, s), _ in self.order)
        self.out = torch.empty(total, dtype=torch.float32, device=dev)
        recs, off, blk = [], 0, 0
        self.views = {}
        for (ptr, shape), _ in self.order:
            co, r, s, ci = shape
            view = self.out[off:off + co * r * s * ci].view(ci, r, s, co)
            self.views[(ptr, shape)] = view
            recs.append(struct.pack("<QQiiii", ptr, view.data_ptr(), co, r * s, ci, blk))
            blk += ((ci + 31) // 32) * ((co + 31) // 32) * r * s
            off += co * r * s * ci
        raw = b"".join(recs)
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.total_blocks = blk
        self.dirty = False

    def get(self, w: torch.Tensor) -> torch.Tensor:
        key = (w.data_ptr(), tuple(w.shape))
        if key not in self.entries:
            return weight_transpose(w)
        if self.dirty:
            self._rebuild()
        cur = torch.cuda.current_stream()
        if not self.fresh:
            nat.check(nat.lib().dtf_weight_transpose_batched(C.c_void_p(self.table.data_ptr()), len(self.order),
                                                             self.total_blocks, nat.stream_ptr()),
                      "dtf_weight_transpose_batched")
            self.fresh = True
            # branches of a block run their backward on different streams (ops/streams.py): whoever comes second
            # must order itself after the launch above
            self.done = torch.cuda.Event()
            self.done.record(cur)
            self.done_stream = cur
        elif getattr(self, "done", None) is not None and cur != self.done_stream:
            cur.wait_event(self.done)
        return self.views[key]


_current_wt_batch: Optional[WeightTransposeBatch] = None


def set_wt_batch(batch: Optional[WeightTransposeBatch]) -> None:
    """Install the batch that convolutions register with (the engine owns one per model; ``None`` = one transpose
    kernel per layer)."""
    global _current_wt_batch
    _current_wt_batch = batch


_wt_register_off = [0]


@contextlib.contextmanager
def temporary_weights():
    """Convolutions run inside do not join the step's batched transpose: their weight operand is a temporary (the
    fake-quantised weights of ``slim/quantize.py``) whose address is not a property of the model, and the batch
    is keyed by address.  Their dgrad operand comes from a transpose launch of their own."""
    _wt_register_off[0] += 1
    try:
        yield
    finally:
        _wt_register_off[0] -= 1


def wt_register(w: torch.Tensor) -> None:
    if _current_wt_batch is not None and not _wt_register_off[0]:
        _current_wt_batch.register(w)


def wt_get(w: torch.Tensor) -> torch.Tensor:
    return _current_wt_batch.get(w) if _current_wt_batch is not None else weight_transpose(w)


def weight_transpose(w: torch.Tensor) -> torch.Tensor:
    """OHWI ``[C_out,R,S,C_in]`` -> ``[C_in,R,S,C_out]`` (B operand of the dgrad GEMM)."""
    co, r, s, ci = w.shape
    wt = torch.empty((ci, r, s, co), device=w.device, dtype=torch.float32)
    nat.check(nat.lib().dtf_weight_transpose(C.c_void_p(w.data_ptr()), C.c_void_p(wt.data_ptr()),
                                             co, r * s, ci, nat.stream_ptr()), "dtf_weight_transpose")
    return wt


def _conv_wgrad_once(x: torch.Tensor, dy: torch.Tensor, w_shape: Sequence[int], stride: Tuple[int, int],
                     pad: Tuple[int, int], *, in_scale: Optional[torch.Tensor] = None,
               in_shift: Optional[torch.Tensor] = None, in_relu: bool = False,
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dW (OHWI).  Accumulates into ``out`` when given (it must already hold zeros / a running sum)."""
    n, h, wd, c = x.shape
    co, r, s, ci = w_shape
    assert ci == c
    _, oh, ow, co2 = dy.shape
    assert co2 == co
    if out is None:
        out = torch.zeros((co, r, s, ci), device=x.device, dtype=torch.float32)
    p = nat.ConvWgradParams()
    p.src, p.srcH, p.srcW, p.srcC, p.src_ld = x.data_ptr(), h, wd, c, _pitch(x)
    p.N, p.OH, p.OW = n, oh, ow
    p.R, p.S = r, s
    p.stride_h, p.stride_w, p.pad_h, p.pad_w = stride[0], stride[1], pad[0], pad[1]
    p.dy, p.dy_ld, p.Nout = dy.data_ptr(), _pitch(dy), co
    p.dw = out.data_ptr()
    p.in_scale, p.in_shift, p.in_relu = nat.ptr(in_scale), nat.ptr(in_shift), int(in_relu)
    nat.check(nat.lib().dtf_conv_wgrad(C.byref(p), nat.stream_ptr()), "dtf_conv_wgrad")
    return out


# ---------------------------------------------------------------------------------------------------
# precision: TF32 (default) or 3xTF32
# ---------------------------------------------------------------------------------------------------
# The tensor core multiplies fp32 operands as TF32 (10 mantissa bits) and accumulates in fp32; the reference is fp32
# end to end (inception/slim/variables.py:249).  For parity runs and fp32-sensitive users every operand can be split
# into a part the tensor core represents exactly and the remainder,
#     a = a_hi + a_lo,   a_hi = a with the 13 low mantissa bits cleared,   a_lo = a - a_hi  (exact in fp32)
# and a product is then formed as a_hi*b_hi + a_hi*b_lo + a_lo*b_hi — three passes of the same kernel accumulating into
# one fp32 output (the dropped a_lo*b_lo term is ~2^-22 relative).  That is the error of an fp32 multiply to within a
# bit or two, at three times the tensor-core work.  Fused input transforms (batch-norm apply + ReLU on load) are not
# linear in the operand, so layers that use them keep the single pass; the engine switches that fusion off when
# ``precision='tf32x3'`` is asked for.
_precision = ["tf32"]


def set_precision(name: str) -> None:
    if name not in ("tf32", "tf32x3"):
        raise ValueError("precision must be 'tf32' or 'tf32x3', not %r" % (name,))
    _precision[0] = name


def get_precision() -> str:
    return _precision[0]


def _split_tf32(t: torch.Tensor):
    hi = (t.view(torch.int32) & -8192).view(torch.float32)
    return hi, t - hi


def conv_fprop(x: torch.Tensor, w: torch.Tensor, stride: Tuple[int, int], pad: Tuple[int, int], *,
               bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
               relu: bool = False, in_scale: Optional[torch.Tensor] = None,
               in_shift: Optional[torch.Tensor] = None, in_relu: bool = False,
               stats: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
               out_hw: Optional[Tuple[int, int]] = None, want_act: bool = False, bn_pending=None):
    """y = epilogue(conv(act(x), w)) — see :func:`_conv_fprop_once`; under ``tf32x3`` three accumulating passes."""
    if _precision[0] == "tf32" or in_scale is not None or bn_pending is not None:
        return _conv_fprop_once(x, w, stride, pad, bias=bias, residual=residual, relu=relu, in_scale=in_scale,
                                in_shift=in_shift, in_relu=in_relu, stats=stats, out=out, out_hw=out_hw,
                                want_act=want_act, bn_pending=bn_pending)
    xh, xl = _split_tf32(x)
    wh, wl = _split_tf32(w)
    y = _conv_fprop_once(xh, wh, stride, pad, bias=bias, residual=residual, out=out, out_hw=out_hw)
    y = _conv_fprop_once(xh, wl, stride, pad, residual=y, out=y, out_hw=out_hw)
    y = _conv_fprop_once(xl, wh, stride, pad, residual=y, out=y, out_hw=out_hw, relu=relu, stats=stats)
    return (y, None) if want_act else y


def conv_dgrad(dy: torch.Tensor, w_t: torch.Tensor, in_hw: Tuple[int, int], stride: Tuple[int, int],
               pad: Tuple[int, int], out: Optional[torch.Tensor] = None, bn_bwd=None) -> torch.Tensor:
    """dx = conv_transpose(dy, w) — see :func:`_conv_dgrad_once`; under ``tf32x3`` three accumulating passes."""
    if _precision[0] == "tf32" or bn_bwd is not None:
        return _conv_dgrad_once(dy, w_t, in_hw, stride, pad, out=out, bn_bwd=bn_bwd)
    gh, gl = _split_tf32(dy)
    wh, wl = _split_tf32(w_t)
    dx = _conv_dgrad_once(gh, wh, in_hw, stride, pad, out=out)
    dx = _conv_dgrad_once(gh, wl, in_hw, stride, pad, out=dx, residual=dx)
    return _conv_dgrad_once(gl, wh, in_hw, stride, pad, out=dx, residual=dx)


def conv_wgrad(x: torch.Tensor, dy: torch.Tensor, w_shape: Sequence[int], stride: Tuple[int, int],
               pad: Tuple[int, int], *, in_scale: Optional[torch.Tensor] = None,
               in_shift: Optional[torch.Tensor] = None, in_relu: bool = False,
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dW (OHWI), accumulated into ``out`` when given — see :func:`_conv_wgrad_once`; under ``tf32x3`` three passes
    (the kernel accumulates into dW anyway)."""
    if _precision[0] == "tf32" or in_scale is not None:
        return _conv_wgrad_once(x, dy, w_shape, stride, pad, in_scale=in_scale, in_shift=in_shift, in_relu=in_relu,
                                out=out)
    xh, xl = _split_tf32(x)
    gh, gl = _split_tf32(dy)
    out = _conv_wgrad_once(xh, gh, w_shape, stride, pad, out=out)
    out = _conv_wgrad_once(xh, gl, w_shape, stride, pad, out=out)
    return _conv_wgrad_once(xl, gh, w_shape, stride, pad, out=out)
