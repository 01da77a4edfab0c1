"""AlexNet-v2 in the CIFAR-adapted form of the reference (``alexnet/alexnet.py:56-141``).

For ``num_classes == 10`` conv1 runs with stride 1 (32 -> 22, alexnet.py:102-103), otherwise stride 4.
conv1 11x11 VALID 64 -> maxpool 3/2 -> conv2 5x5 192 -> maxpool -> conv3/4/5 3x3 384/384/256 -> maxpool ->
fc6 = 5x5 conv 4096 (SAME, on a 1x1 map) -> dropout -> fc7 1x1 4096 -> dropout -> fc8 1x1 classes -> squeeze.
The trainer does not apply ``alexnet_v2_arg_scope``: Xavier weights / zero biases / VALID pools
except fc6-fc8, which use sigma=0.005 weights and 0.1 biases (fc8: zero bias).
Names: ``alexnet_v2/{conv1..conv5,fc6,fc7,fc8}/{weights,biases}``.
"""
from __future__ import annotations

import torch

from ..slim import ops, scopes
from ..slim import variables as V

default_image_size = 224


def alexnet_v2(inputs: torch.Tensor, num_classes: int = 1000, is_training: bool = True,
               dropout_keep_prob: float = 0.5, spatial_squeeze: bool = True, scope: str = "alexnet_v2"):
    store = V.current_store()
    end_points = {}
    convs = (("conv2", 192, 5, "pool2"), ("conv3", 384, 3, None), ("conv4", 384, 3, None), ("conv5", 256, 3, "pool5"))
    with store.scope(scope):
        stride1 = 1 if num_classes == 10 else 4
        net = end_points["conv1"] = ops.convolution2d(inputs, 64, [11, 11], stride1, padding="VALID", scope="conv1")
        net = end_points["pool1"] = ops.max_pool2d(net, [3, 3], 2, scope="pool1")
        for name, ch, k, pool in convs:
            net = end_points[name] = ops.convolution2d(net, ch, [k, k], scope=name)
            if pool:
                net = end_points[pool] = ops.max_pool2d(net, [3, 3], 2, scope=pool)
        with scopes.arg_scope([ops.convolution2d], weights_initializer=V.truncated_normal_initializer(0.005),
                              biases_initializer=V.constant_initializer(0.1)):
            net = end_points["fc6"] = ops.convolution2d(net, 4096, [5, 5], padding="SAME", scope="fc6")
            net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout6")
            net = end_points["fc7"] = ops.convolution2d(net, 4096, [1, 1], scope="fc7")
            net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout7")
            net = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None,
                                    biases_initializer=V.zeros_initializer(), scope="fc8")
        if spatial_squeeze:
            net = net.reshape(net.shape[0], net.shape[-1]) if net.shape[1] == 1 and net.shape[2] == 1 else net
        end_points["fc8"] = net
    store.end_points = end_points
    return net, end_points


alexnet_v2.default_image_size = 224
