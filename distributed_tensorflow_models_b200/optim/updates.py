"""Parameter-server update rules on a flat fp32 shard.

This is the math the reference runs on its PS CPUs, restated as in-place updates of
contiguous shard views so that one fused kernel can do reduce + update per shard:

* SGD (``tf.train.GradientDescentOptimizer``, cnn/cifar10_cnn_bsp.py:56):  ``w -= lr * (g + wd*w)``
* RMSProp in its TF1 form (``tf.train.RMSPropOptimizer(lr, 0.9, momentum=0.9, epsilon=1.0)``,
  inception/imagenet_inception_bsp.py:134-136): ``ms = d*ms + (1-d)*g^2 ; mom = m*mom + lr*g/sqrt(ms+eps) ;
  w -= mom`` with ``ms`` initialised to **ones** and epsilon *inside* the square root
* EMA shadows (``tf.train.ExponentialMovingAverage``): ``s -= (1-d)*(s-w)``, d = min(decay,(1+n)/(10+n))
* gradient scale ``batch/base_batch`` (the per-step dynamic batch hook, cnn/cifar10_cnn_bsp.py:66)

``g`` already contains the mean over workers (BSP) or one worker's gradient (ASP/SSP).  The L2 terms of the
reference losses enter as ``wd*w`` for the elements below ``wd_boundary`` (the layout sorts decayed
tensors first), which equals the gradient of ``wd*sum(w^2)/2``.

The torch implementation below is the oracle and the CPU path; on CUDA ``ShardOptimizer.apply`` calls the
fused native kernel (``dtf_fused_update`` / the push kernels).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from .schedule import ema_decay


@dataclass
class EmaConfig:
    decay: float = 0.9999
    use_num_updates: bool = True


@dataclass
class OptimizerConfig:
    kind: str = "sgd"                 # 'sgd' | 'rmsprop'
    rms_decay: float = 0.9
    momentum: float = 0.9
    epsilon: float = 1.0
    weight_decay: float = 0.0         # value applied to elements [0, wd_boundary)
    ema: Optional[EmaConfig] = None


class ShardOptimizer:
    """State + update for the slice ``[lo, hi)`` of the flat parameter vector owned by one PS rank."""

    def __init__(self, cfg: OptimizerConfig, master: torch.Tensor, lo: int, hi: int, wd_boundary: int):
        self.cfg = cfg
        self.lo, self.hi = lo, hi
        self.w = master[lo:hi]
        n = hi - lo
        self.n_decayed = max(0, min(hi, wd_boundary) - lo)
        self.slots: Dict[str, torch.Tensor] = {}
        if cfg.kind == "rmsprop":
            self.slots["RMSProp"] = torch.ones(n, dtype=master.dtype, device=master.device)      # ms
            self.slots["RMSProp_1"] = torch.zeros(n, dtype=master.dtype, device=master.device)   # mom
        elif cfg.kind != "sgd":
            raise ValueError(cfg.kind)
        self.ema: Optional[torch.Tensor] = None
        if cfg.ema is not None:
            self.ema = self.w.detach().clone()

    # -- reference math (torch ops; also the CPU path) -----------------------------------------
    @torch.no_grad()
    def apply_reference(self, g: torch.Tensor, lr: float, grad_scale: float = 1.0,
                        num_updates: Optional[int] = None, use_ema: bool = True) -> None:
        w = self.w
        g = g * grad_scale if grad_scale != 1.0 else g.clone()
        if self.cfg.weight_decay and self.n_decayed > 0:
            g[: self.n_decayed].add_(w[: self.n_decayed], alpha=self.cfg.weight_decay)
        if self.cfg.kind == "sgd":
            w.add_(g, alpha=-lr)
        else:
            ms, mom = self.slots["RMSProp"], self.slots["RMSProp_1"]
            ms.mul_(self.cfg.rms_decay).addcmul_(g, g, value=1.0 - self.cfg.rms_decay)
            mom.mul_(self.cfg.momentum).add_(g / torch.sqrt(ms + self.cfg.epsilon), alpha=lr)
            w.sub_(mom)
        if self.ema is not None and use_ema:
            d = ema_decay(self.cfg.ema.decay, num_updates if self.cfg.ema.use_num_updates else None)
            self.ema.sub_((self.ema - w) * (1.0 - d))

    @torch.no_grad()
    def apply(self, g: torch.Tensor, lr: float, grad_scale: float = 1.0, num_updates: Optional[int] = None,
              use_ema: bool = True) -> None:
        if self.w.is_cuda:
            from ..ops import native_ops
            native_ops.fused_update(self, g, lr, grad_scale, num_updates, use_ema)
        else:
            self.apply_reference(g, lr, grad_scale, num_updates, use_ema)

    def state_tensors(self) -> Dict[str, torch.Tensor]:
        out = dict(self.slots)
        if self.ema is not None:
            out["ExponentialMovingAverage"] = self.ema
        return out
