"""Small members of the vendored model zoo: LeNet and OverFeat.

The reference reaches them only through ``nets_factory`` (``vgg/nets/nets_factory.py:39-100``); they are built here
from the same layer library as the trained models, with the zoo's layer specifications:

* ``lenet``  (``vgg/nets/lenet.py:60-77``): conv 5x5/32 -> max-pool 2/2 -> conv 5x5/64 -> max-pool 2/2 -> flatten ->
  fc 1024 -> dropout -> fc classes.  Arg scope: truncated-normal(0.1) weights, L2 ``weight_decay``, ReLU.
  Names ``LeNet/{conv1,conv2,fc3,fc4}/{weights,biases}``; 28x28 inputs.
* ``overfeat`` (``vgg/nets/overfeat.py:93-124``): conv 11x11/4 VALID 64 -> pool 2 -> conv 5x5 VALID 256 -> pool 2 ->
  conv 3x3 512 / 1024 / 1024 -> pool 2 -> fc6 as a 6x6 VALID conv 3072 -> dropout -> fc7 1x1 4096 -> dropout ->
  fc8 1x1 classes; the fc layers use sigma = 0.005 weights and 0.1 biases.  231x231 inputs.
"""
from __future__ import annotations

import contextlib

from ..slim import losses, ops, scopes
from ..slim import variables as V


@contextlib.contextmanager
def lenet_arg_scope(weight_decay: float = 0.0):
    reg = losses.l2_regularizer(weight_decay) if weight_decay else None
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected], weights_regularizer=reg,
                          weights_initializer=V.truncated_normal_initializer(0.1), activation_fn=ops.relu) as sc:
        yield sc


def lenet(images, num_classes: int = 10, is_training: bool = False, dropout_keep_prob: float = 0.5,
          scope: str = "LeNet"):
    store = V.current_store()
    end_points = {}
    with store.scope(scope):
        net = end_points["conv1"] = ops.convolution2d(images, 32, [5, 5], scope="conv1")
        net = end_points["pool1"] = ops.max_pool2d(net, [2, 2], 2, scope="pool1")
        net = end_points["conv2"] = ops.convolution2d(net, 64, [5, 5], scope="conv2")
        net = end_points["pool2"] = ops.max_pool2d(net, [2, 2], 2, scope="pool2")
        net = end_points["Flatten"] = ops.flatten(net)
        net = end_points["fc3"] = ops.fully_connected(net, 1024, scope="fc3")
        if not num_classes:
            return net, end_points
        net = end_points["dropout3"] = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout3")
        logits = end_points["Logits"] = ops.fully_connected(net, num_classes, activation_fn=None, scope="fc4")
    end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


lenet.default_image_size = 28


@contextlib.contextmanager
def overfeat_arg_scope(weight_decay: float = 0.0005):
    reg = losses.l2_regularizer(weight_decay) if weight_decay else None
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected], activation_fn=ops.relu,
                          weights_regularizer=reg, biases_initializer=V.zeros_initializer()):
        with scopes.arg_scope([ops.convolution2d], padding="SAME"):
            with scopes.arg_scope([ops.max_pool2d], padding="VALID") as sc:
                yield sc


def overfeat(inputs, num_classes: int = 1000, is_training: bool = True, dropout_keep_prob: float = 0.5,
             spatial_squeeze: bool = True, scope: str = "overfeat"):
    store = V.current_store()
    end_points = {}
    with store.scope(scope):
        net = end_points["conv1"] = ops.convolution2d(inputs, 64, [11, 11], 4, padding="VALID", scope="conv1")
        net = end_points["pool1"] = ops.max_pool2d(net, [2, 2], scope="pool1")
        net = end_points["conv2"] = ops.convolution2d(net, 256, [5, 5], padding="VALID", scope="conv2")
        net = end_points["pool2"] = ops.max_pool2d(net, [2, 2], scope="pool2")
        net = end_points["conv3"] = ops.convolution2d(net, 512, [3, 3], scope="conv3")
        net = end_points["conv4"] = ops.convolution2d(net, 1024, [3, 3], scope="conv4")
        net = end_points["conv5"] = ops.convolution2d(net, 1024, [3, 3], scope="conv5")
        net = end_points["pool5"] = ops.max_pool2d(net, [2, 2], scope="pool5")
        with scopes.arg_scope([ops.convolution2d], weights_initializer=V.truncated_normal_initializer(0.005),
                              biases_initializer=V.constant_initializer(0.1)):
            net = end_points["fc6"] = ops.convolution2d(net, 3072, [6, 6], padding="VALID", scope="fc6")
            net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout6")
            net = end_points["fc7"] = ops.convolution2d(net, 4096, [1, 1], scope="fc7")
            if num_classes:
                net = ops.dropout(net, dropout_keep_prob, is_training=is_training, scope="dropout7")
                net = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None,
                                        biases_initializer=V.zeros_initializer(), scope="fc8")
        if spatial_squeeze and net.shape[1] == 1 and net.shape[2] == 1:
            net = net.reshape(net.shape[0], net.shape[-1])
        end_points["fc8"] = net
    store.end_points = end_points
    return net, end_points


overfeat.default_image_size = 231
