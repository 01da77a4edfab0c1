"""The plain CIFAR-10 CNN of the TF tutorial as shipped in the reference (``cnn/cifar10.py:203-282``).

24x24x3 -> conv1 5x5x64 +bias ReLU -> maxpool 3x3/2 SAME -> LRN -> conv2 5x5x64 +bias(0.1) ReLU -> LRN ->
maxpool 3x3/2 SAME -> flatten 2304 -> local3 384 (wd .004) -> local4 192 (wd .004) -> softmax_linear 10.
Initialisers are the reference's, including its sigma=1e-4 conv init.  Variable names:
``conv1/{weights,biases}`` ... ``softmax_linear/{weights,biases}``.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from ..slim import losses
from ..slim import variables as V

NUM_CLASSES = 10
IMAGE_SIZE = 24

# (scope, kind, shape-args, weight stddev, weight decay, bias init)
_LAYERS = (
    ("conv1", "conv", 64, 1e-4, 0.0, 0.0),
    ("conv2", "conv", 64, 1e-4, 0.0, 0.1),
    ("local3", "fc", 384, 0.04, 0.004, 0.1),
    ("local4", "fc", 192, 0.04, 0.004, 0.1),
    ("softmax_linear", "fc", NUM_CLASSES, 1 / 192.0, 0.0, 0.0),
)


def _layer_vars(store, scope, kind, cin, cout, stddev, wd, bias0):
    with store.scope(scope):
        if kind == "conv":
            w = store.variable("weights", (cout, 5, 5, cin), V.truncated_normal_initializer(stddev), weight_decay=wd,
                               tf_layout="hwio", tf_shape=(5, 5, cin, cout))
        else:
            w = store.variable("weights", (cout, cin), V.truncated_normal_initializer(stddev), weight_decay=wd,
                               tf_layout="io", tf_shape=(cin, cout))
        b = store.variable("biases", (cout,), V.constant_initializer(bias0))
    return w, b


def inference(images: torch.Tensor, batch_size=None) -> torch.Tensor:
    """images NHWC [N,24,24,3] -> logits [N,10] (reference cnn/cifar10.py:203-282)."""
    store = V.current_store()
    spec = {s[0]: s for s in _LAYERS}
    w, b = _layer_vars(store, *spec["conv1"][:2], images.shape[-1], *spec["conv1"][2:])
    x = Fn.conv2d(images, w, b, 1, "SAME", relu=True)
    end_points = {"conv1": x}          # what the reference hands to _activation_summary (cnn/cifar10.py:226-280)
    x = Fn.max_pool(x, 3, 2, "SAME")
    x = Fn.lrn(x, 4, 1.0, 0.001 / 9.0, 0.75)
    w, b = _layer_vars(store, *spec["conv2"][:2], 64, *spec["conv2"][2:])
    x = Fn.conv2d(x, w, b, 1, "SAME", relu=True)
    end_points["conv2"] = x
    x = Fn.lrn(x, 4, 1.0, 0.001 / 9.0, 0.75)
    x = Fn.max_pool(x, 3, 2, "SAME")
    x = Fn.flatten(x)
    for name in ("local3", "local4"):
        w, b = _layer_vars(store, *spec[name][:2], x.shape[-1], *spec[name][2:])
        x = end_points[name] = Fn.linear(x, w, b, relu=True)
    w, b = _layer_vars(store, *spec["softmax_linear"][:2], x.shape[-1], *spec["softmax_linear"][2:])
    logits = end_points["softmax_linear"] = Fn.linear(x, w, b)
    store.end_points = end_points
    return logits


def loss(logits: torch.Tensor, labels: torch.Tensor, batch_size=None) -> torch.Tensor:
    """mean sparse softmax cross-entropy (+ the weight-decay terms unless the optimiser applies them
    fused) — reference cnn/cifar10.py:285-306."""
    return losses.cross_entropy_loss(logits, labels.long())
