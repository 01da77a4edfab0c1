"""VGG-A / VGG-16 / VGG-19 (TF-slim) with the reference's local edit (``vgg/nets/vgg.py:144-222``):
fc6 is a 7x7 convolution with padding forced to 'same' (vgg.py:202) so 32x32 CIFAR inputs, which reach
fc6 as a 1x1 map, survive.  All convs 3x3 SAME + ReLU, maxpool 2x2/2, fc7 1x1 4096, fc8 1x1 classes.
``vgg_arg_scope``: ReLU, zero biases, l2(weight_decay) (vgg.py:49-66); the factory calls it with
``weight_decay=0.0`` and ``is_training=False`` (nets_factory.py:103,139) so dropout is inactive in the
reference's training runs.
Names: ``vgg_16/conv{i}/conv{i}_{j}/{weights,biases}``, ``vgg_16/fc{6,7,8}/{weights,biases}``.
"""
from __future__ import annotations

import contextlib

import torch

from ..slim import losses, ops, scopes
from ..slim import variables as V

_CFG = {
    "vgg_a": (1, 1, 2, 2, 2),
    "vgg_16": (2, 2, 3, 3, 3),
    "vgg_19": (2, 2, 4, 4, 4),
}
_WIDTH = (64, 128, 256, 512, 512)


@contextlib.contextmanager
def vgg_arg_scope(weight_decay: float = 0.0005):
    reg = losses.l2_regularizer(weight_decay) if weight_decay else None
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected], activation_fn=ops.relu,
                          weights_regularizer=reg, biases_initializer=V.zeros_initializer()):
        with scopes.arg_scope([ops.convolution2d], padding="SAME") as sc:
            yield sc


def _vgg(name: str, inputs, num_classes, is_training, dropout_keep_prob, spatial_squeeze, fc_conv_padding,
         scope=None):
    store = V.current_store()
    end_points = {}
    with store.scope(scope or name):
        net = inputs
        for i, (reps, ch) in enumerate(zip(_CFG[name], _WIDTH), start=1):
            net = ops.repeat(net, reps, ops.convolution2d, ch, [3, 3], scope=f"conv{i}")
            end_points[f"conv{i}"] = net
            net = end_points[f"pool{i}"] = ops.max_pool2d(net, [2, 2], 2, scope=f"pool{i}")
        net = end_points["fc6"] = ops.convolution2d(net, 4096, [7, 7], padding=fc_conv_padding, scope="fc6")
        net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout6")
        net = end_points["fc7"] = ops.convolution2d(net, 4096, [1, 1], scope="fc7")
        net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout7")
        net = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, scope="fc8")
        if spatial_squeeze and net.shape[1] == 1 and net.shape[2] == 1:
            net = net.reshape(net.shape[0], net.shape[-1])
        end_points["fc8"] = net
    store.end_points = end_points
    return net, end_points


def vgg_16(inputs: torch.Tensor, num_classes: int = 1000, is_training: bool = True, dropout_keep_prob: float = 0.5,
           spatial_squeeze: bool = True, scope: str = "vgg_16", fc_conv_padding: str = "SAME"):
    return _vgg("vgg_16", inputs, num_classes, is_training, dropout_keep_prob, spatial_squeeze,
                fc_conv_padding, scope)


def vgg_a(inputs, num_classes=1000, is_training=True, dropout_keep_prob=0.5, spatial_squeeze=True,
          scope="vgg_a", fc_conv_padding="SAME"):
    return _vgg("vgg_a", inputs, num_classes, is_training, dropout_keep_prob, spatial_squeeze, fc_conv_padding, scope)


def vgg_19(inputs, num_classes=1000, is_training=True, dropout_keep_prob=0.5, spatial_squeeze=True,
           scope="vgg_19", fc_conv_padding="SAME"):
    return _vgg("vgg_19", inputs, num_classes, is_training, dropout_keep_prob, spatial_squeeze, fc_conv_padding, scope)


vgg_16.default_image_size = vgg_a.default_image_size = vgg_19.default_image_size = 224
