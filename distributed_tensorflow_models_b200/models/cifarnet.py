"""CifarNet (TF-slim) as the reference trains it (``cifarnet/cifarnet.py:28-95``).

32x32x3 -> conv1 5x5x64 -> maxpool 2/2 -> LRN -> conv2 5x5x64 -> LRN -> maxpool 2/2 -> flatten 4096 ->
fc3 384 -> dropout .5 -> fc4 192 -> logits 10 -> softmax.  The reference trainer does not apply
``cifarnet_arg_scope``, so layers get slim defaults (Xavier weights, zero biases, ReLU) except
the logits layer, which sets its own initialisers.  The function returns **softmax probabilities**
(cifarnet.py:95) — the trainer then feeds them to a softmax cross-entropy (quirk Q4); ``return_logits``
selects the sane behaviour.  Names: ``CifarNet/{conv1,conv2,fc3,fc4,logits}/{weights,biases}``.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from ..slim import ops
from ..slim import variables as V

default_image_size = 32


def cifarnet(images: torch.Tensor, num_classes: int = 10, is_training: bool = False,
             dropout_keep_prob: float = 0.5, prediction_fn=ops.softmax, scope: str = "CifarNet",
             return_logits: bool = False):
    store = V.current_store()
    end_points = {}
    with store.scope(scope):
        net = end_points["conv1"] = ops.convolution2d(images, 64, [5, 5], scope="conv1")
        net = end_points["pool1"] = ops.max_pool2d(net, [2, 2], 2, scope="pool1")
        net = Fn.lrn(net, 4, 1.0, 0.001 / 9.0, 0.75)
        net = end_points["conv2"] = ops.convolution2d(net, 64, [5, 5], scope="conv2")
        net = Fn.lrn(net, 4, 1.0, 0.001 / 9.0, 0.75)
        net = end_points["pool2"] = ops.max_pool2d(net, [2, 2], 2, scope="pool2")
        net = end_points["Flatten"] = ops.flatten(net)
        net = end_points["fc3"] = ops.fully_connected(net, 384, scope="fc3")
        net = end_points["dropout3"] = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout3")
        net = end_points["fc4"] = ops.fully_connected(net, 192, scope="fc4")
        logits = ops.fully_connected(net, num_classes, activation_fn=None,
                                     weights_initializer=V.truncated_normal_initializer(1 / 192.0),
                                     biases_initializer=V.zeros_initializer(), scope="logits")
        end_points["Logits"] = logits
        end_points["Predictions"] = prediction_fn(logits)
    store.end_points = end_points
    return logits if return_logits else end_points["Predictions"]


cifarnet.default_image_size = 32
