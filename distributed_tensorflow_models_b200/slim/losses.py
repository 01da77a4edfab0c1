"""Regularisers and losses (role of the reference's ``inception/slim/losses.py``).

Regulariser objects carry their weights; attaching one to a variable records the weight on the variable so
that the parameter-server update can apply ``wd * w`` inside the fused push kernel instead of
materialising a loss tensor.  Calling the object on a tensor still returns the loss value
exactly as the reference's closures do (losses.py:37-99), which is what the oracle tests use.

``cross_entropy_loss`` follows losses.py:142-174: label smoothing ``y(1-eps)+eps/C``, result
``weight * mean(xent)``, appended to the current store's loss list (the ``_losses`` collection).
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import functional as Fn
from . import variables as V


class _Regularizer:
    def __init__(self, l1: float = 0.0, l2: float = 0.0, scope: Optional[str] = None):
        self.l1 = float(l1)
        self.l2 = float(l2)
        self.scope = scope

    def __call__(self, tensor: torch.Tensor) -> torch.Tensor:
        out = torch.zeros((), dtype=tensor.dtype, device=tensor.device)
        if self.l1:
            out = out + self.l1 * tensor.abs().sum()
        if self.l2:
            out = out + self.l2 * 0.5 * tensor.pow(2).sum()     # tf.nn.l2_loss = sum(t^2)/2
        return out


def l1_regularizer(weight: float = 1.0, scope: Optional[str] = None) -> _Regularizer:
    return _Regularizer(l1=weight, scope=scope or "L1Regularizer")


def l2_regularizer(weight: float = 1.0, scope: Optional[str] = None) -> _Regularizer:
    return _Regularizer(l2=weight, scope=scope or "L2Regularizer")


def l1_l2_regularizer(weight_1: float = 1.0, weight_2: float = 1.0, scope: Optional[str] = None) -> _Regularizer:
    return _Regularizer(l1=weight_1, l2=weight_2, scope=scope or "L1L2Regularizer")


def _collect(loss: torch.Tensor) -> torch.Tensor:
    try:
        V.current_store().losses.append(loss)
    except RuntimeError:
        pass
    return loss


def l1_loss(tensor: torch.Tensor, weight: float = 1.0, scope: Optional[str] = None) -> torch.Tensor:
    return _collect(weight * tensor.abs().sum())


def l2_loss(tensor: torch.Tensor, weight: float = 1.0, scope: Optional[str] = None) -> torch.Tensor:
    return _collect(weight * 0.5 * tensor.pow(2).sum())


def cross_entropy_loss(logits: torch.Tensor, one_hot_labels: torch.Tensor, label_smoothing: float = 0.0,
                       weight: float = 1.0, scope: Optional[str] = None) -> torch.Tensor:
    """``one_hot_labels``: dense [N, C] targets, or class indices [N] (fast fused path)."""
    if one_hot_labels.dim() == 2 and one_hot_labels.shape != logits.shape:
        raise ValueError("logits and one_hot_labels must have the same shape")
    return _collect(Fn.softmax_cross_entropy(logits, one_hot_labels, label_smoothing, weight))
