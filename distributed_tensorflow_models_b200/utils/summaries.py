"""The summaries the reference's model files define for TensorBoard: per-activation histogram and
sparsity (``_activation_summary``, cnn/cifar10.py:104-121; ``_activation_summaries`` over the end points,
inception/inception_model.py:139-158), the moving average of every loss next to its raw value
(``_add_loss_summaries``, cnn/cifar10.py:309-335; the chief's ``loss_averages`` in
inception/imagenet_inception_bsp.py:116-120), and histograms of every trainable variable and of its gradient
(``cifar10.train``, cnn/cifar10.py:376-382).  The reference's trainers run with ``summary_op=None``, so it never
writes them during training; here they are opt-in (``--summarize_activations``) and land in the chief's event file.
"""
from __future__ import annotations

import re
from typing import Dict, Optional, Tuple

import torch

from ..ops import functional as Fn

TOWER_NAME = "tower"            # vestige of the multi-GPU-tower tutorial (cnn/cifar10.py:99): stripped from names


def _clean(name: str) -> str:
    return re.sub("%s_[0-9]*/" % TOWER_NAME, "", name)


def activation_summaries(end_points: Dict[str, torch.Tensor]) -> Tuple[Dict[str, float], Dict[str, torch.Tensor]]:
    """-> (scalars ``<name>/sparsity``, histograms ``<name>/activations``) for every end point."""
    scalars, histos = {}, {}
    for name, x in end_points.items():
        if not isinstance(x, torch.Tensor) or not x.is_floating_point():
            continue
        n = _clean(name)
        scalars[n + "/sparsity"] = float(Fn.zero_fraction(x.detach()))
        histos[n + "/activations"] = x.detach()
    return scalars, histos


def variable_summaries(store, layout=None) -> Dict[str, torch.Tensor]:
    """Histograms ``<var>`` of every trainable variable and ``<var>/gradients`` of its current gradient."""
    histos = {}
    for v in store.trainable_variables():
        histos[v.name] = v.tensor.detach()
        g = v.tensor.grad
        if g is not None:
            histos[v.name + "/gradients"] = g.detach()
    return histos


class LossAverages:
    """``tf.train.ExponentialMovingAverage(0.9)`` over the losses: ``<name> (raw)`` is the value of the step,
    ``<name>`` its moving average (zero-debiasing is not applied, as in the reference)."""

    def __init__(self, decay: float = 0.9):
        self.decay = decay
        self.avg: Dict[str, float] = {}

    def update(self, **losses: float) -> Dict[str, float]:
        out = {}
        for name, value in losses.items():
            value = float(value)
            prev = self.avg.get(name)
            # TensorFlow initialises the shadow variable with the first value it sees
            self.avg[name] = value if prev is None else prev - (1.0 - self.decay) * (prev - value)
            out[name + " (raw)"] = value
            out[name] = self.avg[name]
        return out
