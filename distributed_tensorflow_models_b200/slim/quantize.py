"""Quantisation-aware training and evaluation (``--quantize``).

Counterpart of the ``tf.contrib.quantize.create_training_graph(quant_delay=...)`` /
``create_eval_graph()`` rewrites that the reference's MobileNet scripts apply to the finished graph
(vgg/nets/mobilenet_v1_train.py:66-74,135-139, mobilenet_v1_eval.py:126-127).  There is no graph to rewrite here,
so the layer functions of ``slim.ops`` ask this module what to do while they run; a network needs no change.
With quantisation switched on (``store.quant`` is a :class:`QuantState`), every convolution, depthwise
convolution and fully connected layer becomes

    s      = gamma / sqrt(moving_variance + eps)                (batch-norm fold; 1 without a normaliser)
    w_q    = fake_quant(w * s, [min(w*s), max(w*s)], 8 bits, narrow range)        "weights_quant"
    z      = conv(x, w_q)
    y      = activation(batch_norm(z / s))   or   activation(z + b)
    y_q    = fake_quant(y, moving-average range, 8 bits)                           "act_quant"

* Weights are quantised *after* the batch-norm scale has been folded in, with the moving variance, so the
  numbers that get rounded are the ones an integer engine would hold; dividing ``z`` by ``s`` afterwards hands
  batch norm the plain convolution output again, so training still normalises with batch statistics.  In
  evaluation mode ``batch_norm(z / s)`` is exactly ``z + (beta - s * moving_mean)``, the folded inference form.
  (The rewrite in TensorFlow takes its batch statistics from a second, unquantised convolution; here they come
  from the quantised one — one convolution per layer instead of two.)
* The weight range is the current min / max of the folded weights (``LastValueQuantize``); the activation range
  is an exponential moving average (decay 0.999) of the batch min / max widened to contain zero
  (``MovingAvgQuantize``), starting from [-6, 6].  Ranges move only in training passes with autograd enabled.
* ``quant_delay``: until the global step reaches it the fake-quant ops are the identity while the ranges already
  track the data.  The switch is a float in device memory (``QuantState.gate``) that the trainer sets before each
  step, so a captured CUDA graph stays valid when the delay runs out.
* Residual additions (MobileNet-v2 blocks, the slim ResNet units) quantise their output too (:func:`after_add`,
  range ``<unit>/add_quant``).  Not quantised: the network input (the converter's job in TensorFlow as well).

Variables: one non-trainable 2-vector ``<layer>/weights_quant/min_max`` and ``<layer>/act_quant/min_max`` per
layer (``depthwise_weights_quant`` / ``pointwise_weights_quant`` for a separable convolution that has both).
Checkpoints hold them as TensorFlow does, as two scalars ``.../min`` and ``.../max`` (``ckpt/store.py``).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, Optional

import torch

from ..ops import functional as Fn
from . import variables as V

MINMAX_LAYOUT = "minmax"


@dataclass
class QuantState:
    """Attach to a store (``store.quant = QuantState(...)``) to switch quantisation on for its networks."""
    quant_delay: int = 0
    weight_bits: int = 8
    activation_bits: int = 8
    ema_decay: float = 0.999
    init_min: float = -6.0
    init_max: float = 6.0
    gate: Optional[torch.Tensor] = None          # scalar float tensor; None = always on
    is_training: bool = True                     # mode of the current pass for layers without a normaliser
    _scratch: Dict[str, torch.Tensor] = field(default_factory=dict)

    def set_step(self, global_step: int) -> None:
        """Open the gate once ``global_step`` has reached ``quant_delay`` (call before each training step)."""
        if self.gate is not None:
            self.gate.fill_(1.0 if global_step >= self.quant_delay else 0.0)

    def scratch(self, name: str, device: torch.device) -> Optional[torch.Tensor]:
        if device.type != "cuda":
            return None
        t = self._scratch.get(name)
        if t is None or t.device != device:
            t = torch.tensor([float("inf"), float("-inf")], dtype=torch.float32, device=device)
            self._scratch[name] = t
        return t


def current() -> Optional[QuantState]:
    try:
        return getattr(V.current_store(), "quant", None)
    except RuntimeError:
        return None


def _range_variable(q: QuantState, name: str) -> torch.Tensor:
    def init(shape, gen, fan_in=None, fan_out=None):
        return torch.tensor([q.init_min, q.init_max], dtype=torch.float32)
    return V.variable(name + "/min_max", (2,), initializer=init, trainable=False, tf_layout=MINMAX_LAYOUT,
                      tf_shape=(2,))


def _may_update(training: bool) -> bool:
    return training and torch.is_grad_enabled()


def quantize_weights(q: QuantState, w: torch.Tensor, training: bool, name: str = "weights_quant") -> torch.Tensor:
    """Last-value range + narrow-range fake quantisation (the variable lands in the current scope)."""
    rng = _range_variable(q, name)
    if _may_update(training):
        full = V.current_store().full_name(name)
        Fn.quant_range_update(w, rng, -1.0, q.scratch(full, w.device))
    return Fn.fake_quant(w, rng, q.gate if training else None, q.weight_bits, narrow_range=True)


def quantize_activation(q: QuantState, y: torch.Tensor, training: bool, name: str = "act_quant") -> torch.Tensor:
    """Moving-average range + fake quantisation of a layer output."""
    rng = _range_variable(q, name)
    if _may_update(training):
        full = V.current_store().full_name(name)
        Fn.quant_range_update(y, rng, q.ema_decay, q.scratch(full, y.device))
    return Fn.fake_quant(y, rng, q.gate if training else None, q.activation_bits, narrow_range=False)


def after_add(y: torch.Tensor, name: str = "add_quant") -> torch.Tensor:
    """Output of a residual addition (after its activation, where it has one): quantised like a layer output when
    quantisation is on, returned as it is otherwise.  Models call it where TensorFlow's rewriter finds a bypass."""
    q = current()
    if q is None:
        return y
    return quantize_activation(q, y, q.is_training, name)


def _safe(s: torch.Tensor) -> torch.Tensor:
    """Keep a fold multiplier away from zero so that the division after the convolution stays finite."""
    tiny = 1e-12
    return torch.where(s.abs() < tiny, torch.full_like(s, tiny), s)


def _run_with_temporary_weights(conv: Callable[[torch.Tensor], torch.Tensor], w_q: torch.Tensor) -> torch.Tensor:
    if w_q.is_cuda:
        from ..ops import conv_native
        with conv_native.temporary_weights():
            return conv(w_q)
    return conv(w_q)


def folded_layer(q: QuantState, conv: Callable[[torch.Tensor], torch.Tensor], w: torch.Tensor,
                 bias: Optional[torch.Tensor], c: int, out_axis_shape, normalizer_params: Optional[dict],
                 activation_fn: Optional[Callable], trainable: bool = True,
                 weights_quant_name: str = "weights_quant", quantize_output: bool = True) -> torch.Tensor:
    """One quantised layer, called inside the layer's variable scope.

    ``conv(w_q)`` runs the layer's linear op with the quantised weights; ``c`` is its number of output channels
    and ``out_axis_shape`` the shape that broadcasts a per-output-channel vector against ``w`` (``(c, 1, 1, 1)``
    for OHWI weights, ``(1, 1, C, multiplier)`` for a depthwise filter);
    ``normalizer_params`` is None for a layer with a bias, else the batch-norm arguments."""
    store = V.current_store()
    if normalizer_params is None:
        training = q.is_training
        y = _run_with_temporary_weights(conv, quantize_weights(q, w, training, weights_quant_name))
        if bias is not None:
            y = y + bias
    else:
        bn = dict(normalizer_params)
        training = bool(bn.pop("is_training", True))
        bn.pop("activation", None)
        names = bn.get("names", ("beta", "gamma", "moving_mean", "moving_variance"))
        eps = bn.get("epsilon", 0.001)
        with store.scope(bn.get("scope"), "BatchNorm"):
            beta = V.variable(names[0], (c,), initializer=V.zeros_initializer(), trainable=trainable) \
                if bn.get("center", True) else None
            gamma = V.variable(names[1], (c,), initializer=V.ones_initializer(), trainable=trainable) \
                if bn.get("scale", False) else None
            colls = [bn.get("moving_vars", "moving_vars"), V.MOVING_AVERAGE_VARIABLES]
            mm = V.variable(names[2], (c,), initializer=V.zeros_initializer(), trainable=False, collections=colls)
            mv = V.variable(names[3], (c,), initializer=V.ones_initializer(), trainable=False, collections=colls)
        s = torch.rsqrt(mv + eps)
        if gamma is not None:
            s = s * gamma
        s = _safe(s)
        w_q = quantize_weights(q, w * s.reshape(out_axis_shape), training, weights_quant_name)
        z = _run_with_temporary_weights(conv, w_q)
        u = z / s
        x4 = u if u.dim() == 4 else u.reshape(u.shape[0], 1, 1, c)
        y = Fn.batch_norm(x4, gamma, beta, mm, mv, training=training, decay=bn.get("decay", 0.999), eps=eps,
                          relu=False, update_moving=bn.get("update_moving", True),
                          unbiased_moving_var=bn.get("unbiased_moving_var", False)).reshape(u.shape)
    if activation_fn is not None:
        y = activation_fn(y)
    if quantize_output:
        y = quantize_activation(q, y, training)
    return y
