"""MobileNet v1 and v2 from the vendored zoo, on this framework's layer library.

* ``mobilenet_v1`` (+ ``_075`` / ``_050`` / ``_025``): reference ``vgg/nets/mobilenet_v1.py:108-140`` (layer table),
  ``:165-310`` (base), ``:313-390`` (head), ``:425-472`` (arg scope).  A 3x3/2 convolution, then thirteen
  depthwise-separable pairs written as two layers each — ``Conv2d_<i>_depthwise`` (``depthwise_weights`` +
  BatchNorm) and ``Conv2d_<i>_pointwise`` (1x1 ``weights`` + BatchNorm) — every one followed by ReLU6; head =
  global average pool (kernel shrunk to the feature map) -> dropout(0.999) -> ``Logits/Conv2d_1c_1x1``.
* ``mobilenet_v2``: reference ``vgg/nets/mobilenet/mobilenet_v2.py:40-85`` (table), ``mobilenet/conv_blocks.py:163-312``
  (the inverted-residual ``expanded_conv``: 1x1 ``expand`` x6 -> 3x3 ``depthwise`` -> linear 1x1 ``project``, identity
  shortcut when stride 1 and depths match), ``mobilenet/mobilenet.py:148-296`` (builder), ``:304-386`` (head),
  ``:415-468`` (training scope).

The depthwise stages run on ``csrc/zoo_ops.cu`` (bandwidth kernels), the 1x1 stages on the tcgen05 GEMM path.
"""
from __future__ import annotations

import contextlib
import functools
from typing import Dict, Optional

from ..slim import losses, ops, quantize, scopes
from ..slim import variables as V

# (kind, kernel, stride, depth): the v1 layer table
_V1_DEFS = [("conv", 3, 2, 32)] + [("sep", 3, s, d) for s, d in (
    (1, 64), (2, 128), (1, 128), (2, 256), (1, 256), (2, 512), (1, 512), (1, 512), (1, 512), (1, 512), (1, 512),
    (2, 1024), (1, 1024))]


@contextlib.contextmanager
def mobilenet_v1_arg_scope(is_training: Optional[bool] = True, weight_decay: float = 0.00004, stddev: float = 0.09,
                           regularize_depthwise: bool = False, batch_norm_decay: float = 0.9997,
                           batch_norm_epsilon: float = 0.001):
    bn = {"center": True, "scale": True, "decay": batch_norm_decay, "epsilon": batch_norm_epsilon}
    if is_training is not None:
        bn["is_training"] = is_training
    reg = losses.l2_regularizer(weight_decay) if weight_decay else None
    with scopes.arg_scope([ops.convolution2d, ops.separable_convolution2d],
                          weights_initializer=V.truncated_normal_initializer(stddev), activation_fn=ops.relu6,
                          normalizer_fn=ops.batch_norm):
        with scopes.arg_scope([ops.batch_norm], **bn):
            with scopes.arg_scope([ops.convolution2d], weights_regularizer=reg):
                with scopes.arg_scope([ops.separable_convolution2d],
                                      weights_regularizer=reg if regularize_depthwise else None) as sc:
                    yield sc


def mobilenet_v1_base(inputs, final_endpoint: str = "Conv2d_13_pointwise", min_depth: int = 8,
                      depth_multiplier: float = 1.0, output_stride: Optional[int] = None, scope: str = "MobilenetV1"):
    """-> (net, end_points) up to ``final_endpoint``.  ``output_stride`` (8 / 16 / 32) switches the remaining
    strided layers to dilation once the running stride reaches it."""
    if depth_multiplier <= 0:
        raise ValueError("depth_multiplier is not greater than zero.")
    if output_stride is not None and output_stride not in (8, 16, 32):
        raise ValueError("Only allowed output_stride values are 8, 16, 32.")
    depth = lambda d: max(int(d * depth_multiplier), min_depth)
    store = V.current_store()
    end_points: Dict[str, object] = {}
    net = inputs
    current_stride, rate = 1, 1
    with store.scope(scope):
        for i, (kind, k, stride, d) in enumerate(_V1_DEFS):
            if output_stride is not None and current_stride == output_stride:
                layer_stride, layer_rate = 1, rate
                rate *= stride
            else:
                layer_stride, layer_rate = stride, 1
                current_stride *= stride
            base = f"Conv2d_{i}"
            if kind == "conv":
                net = end_points[base] = ops.convolution2d(net, depth(d), [k, k], stride=stride, padding="SAME",
                                                           scope=base)
                if base == final_endpoint:
                    return net, end_points
                continue
            name = base + "_depthwise"
            net = end_points[name] = ops.separable_convolution2d(net, None, [k, k], depth_multiplier=1,
                                                                 stride=layer_stride, rate=layer_rate, padding="SAME",
                                                                 scope=name)
            if name == final_endpoint:
                return net, end_points
            name = base + "_pointwise"
            net = end_points[name] = ops.convolution2d(net, depth(d), [1, 1], stride=1, padding="SAME", scope=name)
            if name == final_endpoint:
                return net, end_points
    raise ValueError("Unknown final endpoint %s" % final_endpoint)


def _reduced_kernel(net, k):
    return [min(net.shape[1], k[0]), min(net.shape[2], k[1])]


def mobilenet_v1(inputs, num_classes: int = 1000, dropout_keep_prob: float = 0.999, is_training: bool = True,
                 min_depth: int = 8, depth_multiplier: float = 1.0, spatial_squeeze: bool = True,
                 global_pool: bool = False, scope: str = "MobilenetV1"):
    if inputs.dim() != 4:
        raise ValueError("Invalid input tensor rank, expected 4, was: %d" % inputs.dim())
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        net, end_points = mobilenet_v1_base(inputs, min_depth=min_depth, depth_multiplier=depth_multiplier,
                                            scope=scope)
        with store.scope(scope + "/Logits"):
            if global_pool:
                net = net.mean((1, 2), keepdim=True)
            else:
                net = ops.avg_pool2d(net, _reduced_kernel(net, [7, 7]), stride=1, padding="VALID", scope="AvgPool_1a")
            end_points["AvgPool_1a" if not global_pool else "global_pool"] = net
            if not num_classes:
                return net, end_points
            net = ops.dropout(net, dropout_keep_prob, scope="Dropout_1b")
            logits = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                       scope="Conv2d_1c_1x1")
            if spatial_squeeze:
                logits = logits.reshape(logits.shape[0], logits.shape[-1]) if logits.shape[1:3] == (1, 1) else logits
        end_points["Logits"] = logits
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


mobilenet_v1.default_image_size = 224


def _partial(fn, **kw):
    p = functools.partial(fn, **kw)
    functools.update_wrapper(p, fn)
    return p


mobilenet_v1_075 = _partial(mobilenet_v1, depth_multiplier=0.75)
mobilenet_v1_050 = _partial(mobilenet_v1, depth_multiplier=0.50)
mobilenet_v1_025 = _partial(mobilenet_v1, depth_multiplier=0.25)


# ---------------------------------------------------------------------------------------------------------
# v2
# ---------------------------------------------------------------------------------------------------------
def _make_divisible(v: float, divisor: int, min_value: Optional[int] = None) -> int:
    """Round to the nearest multiple of ``divisor`` without dropping more than 10 % (conv_blocks.py:50-57)."""
    min_value = divisor if min_value is None else min_value
    new_v = max(min_value, int(v + divisor / 2) // divisor * divisor)
    if new_v < 0.9 * v:
        new_v += divisor
    return new_v


# (num_outputs, stride, expansion factor) of the seventeen inverted-residual blocks
_V2_BLOCKS = [(16, 1, 1), (24, 2, 6), (24, 1, 6), (32, 2, 6), (32, 1, 6), (32, 1, 6), (64, 2, 6), (64, 1, 6),
              (64, 1, 6), (64, 1, 6), (96, 1, 6), (96, 1, 6), (96, 1, 6), (160, 2, 6), (160, 1, 6), (160, 1, 6),
              (320, 1, 6)]


def expanded_conv(net, num_outputs: int, stride: int = 1, expansion: int = 6, kernel_size=(3, 3), rate: int = 1,
                  residual: bool = True, divisible_by: int = 8, scope: Optional[str] = None,
                  end_points: Optional[dict] = None):
    """Inverted residual with linear bottleneck: ``expand`` (1x1, ReLU6, only when it widens) -> ``depthwise``
    (3x3, ReLU6) -> ``project`` (1x1, no activation), + input when shapes allow."""
    store = V.current_store()
    with store.scope(scope, "expanded_conv"):
        inp = net
        prev = net.shape[-1]
        inner = _make_divisible(prev * expansion, divisible_by if expansion != 1 else 1)
        if inner > prev:
            net = ops.convolution2d(net, inner, [1, 1], scope="expand")
        if end_points is not None:
            end_points["expansion_output"] = net
        net = ops.separable_convolution2d(net, None, list(kernel_size), depth_multiplier=1, stride=stride, rate=rate,
                                          scope="depthwise")
        if end_points is not None:
            end_points["depthwise_output"] = net
        net = ops.convolution2d(net, num_outputs, [1, 1], activation_fn=None, scope="project")
        if residual and stride == 1 and net.shape[-1] == inp.shape[-1]:
            net = quantize.after_add(net + inp)
        return net


@contextlib.contextmanager
def mobilenet_v2_training_scope(is_training: Optional[bool] = True, weight_decay: float = 0.00004,
                                stddev: float = 0.09, dropout_keep_prob: Optional[float] = 0.8,
                                bn_decay: float = 0.997):
    """Truncated-normal(0.09) weights, L2 on conv weights but not on depthwise filters, BN decay 0.997,
    dropout keep 0.8 (mobilenet/mobilenet.py:415-468)."""
    bn = {"decay": bn_decay}
    if is_training is not None:
        bn["is_training"] = is_training
    drop = {}
    if is_training is not None:
        drop["is_training"] = is_training
    if dropout_keep_prob is not None:
        drop["keep_prob"] = dropout_keep_prob
    init = V.truncated_normal_initializer(stddev) if stddev >= 0 else V.xavier_initializer()
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected, ops.separable_convolution2d],
                          weights_initializer=init):
        with scopes.arg_scope([ops.batch_norm], **bn):
            with scopes.arg_scope([ops.dropout], **drop):
                with scopes.arg_scope([ops.convolution2d], weights_regularizer=losses.l2_regularizer(weight_decay)):
                    with scopes.arg_scope([ops.separable_convolution2d], weights_regularizer=None) as sc:
                        yield sc


def mobilenet_v2_base(inputs, depth_multiplier: float = 1.0, min_depth: Optional[int] = None,
                      divisible_by: Optional[int] = None, final_endpoint: Optional[str] = None,
                      finegrain_classification_mode: bool = False, scope: str = "MobilenetV2"):
    """-> (net, end_points); end points ``layer_1 .. layer_19`` as in the reference builder."""
    if depth_multiplier <= 0:
        raise ValueError("depth_multiplier is not greater than zero.")
    div = divisible_by or 8
    mind = min_depth or 8
    depth = lambda d: _make_divisible(d * depth_multiplier, div, mind)
    store = V.current_store()
    end_points: Dict[str, object] = {}
    with store.scope(scope):
        # architecture-defining defaults (V2_DEF['defaults']): BN with gamma, BN + ReLU6 after every convolution
        with scopes.arg_scope([ops.batch_norm], center=True, scale=True):
            with scopes.arg_scope([ops.convolution2d, ops.separable_convolution2d], normalizer_fn=ops.batch_norm,
                                  activation_fn=ops.relu6, padding="SAME"):
                net = end_points["layer_1"] = ops.convolution2d(inputs, depth(32), [3, 3], stride=2, scope="Conv")
                if final_endpoint == "layer_1":
                    return net, end_points
                for i, (d, stride, t) in enumerate(_V2_BLOCKS):
                    name = "expanded_conv" if i == 0 else f"expanded_conv_{i}"
                    sub: dict = {}
                    net = expanded_conv(net, depth(d), stride=stride, expansion=t, scope=name, end_points=sub)
                    end_points[f"layer_{i + 2}"] = net
                    for k, v in sub.items():
                        end_points[f"layer_{i + 2}/{k}"] = v
                    if final_endpoint == f"layer_{i + 2}":
                        return net, end_points
                last = 1280 if (finegrain_classification_mode and depth_multiplier < 1.0) else depth(1280)
                net = end_points["layer_19"] = ops.convolution2d(net, last, [1, 1], scope="Conv_1")
    return net, end_points


def mobilenet_v2(inputs, num_classes: int = 1001, depth_multiplier: float = 1.0,
                 is_training: Optional[bool] = None,
                 dropout_keep_prob: Optional[float] = None, min_depth: Optional[int] = None,
                 divisible_by: Optional[int] = None, finegrain_classification_mode: bool = False,
                 scope: str = "MobilenetV2"):
    """Inference mode unless wrapped in ``mobilenet_v2_training_scope`` / ``is_training=True``."""
    if inputs.dim() != 4:
        raise ValueError("Expected rank 4 input, was: %d" % inputs.dim())
    store = V.current_store()
    if is_training is None:      # inherit from an enclosing training scope, else inference
        is_training = scopes.current_arg_scope().get(ops.batch_norm._dtf_arg_scope_key, {}).get("is_training", False)
    train = is_training
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=bool(train)):
        net, end_points = mobilenet_v2_base(inputs, depth_multiplier, min_depth, divisible_by,
                                            finegrain_classification_mode=finegrain_classification_mode, scope=scope)
        with store.scope(scope + "/Logits"):
            net = end_points["global_pool"] = net.mean((1, 2), keepdim=True) if (net.shape[1] * net.shape[2] == 1) \
                else ops.avg_pool2d(net, [net.shape[1], net.shape[2]], stride=1, padding="VALID")
            if not num_classes:
                return net, end_points
            if dropout_keep_prob is not None:
                net = ops.dropout(net, dropout_keep_prob, scope="Dropout")
            else:
                net = ops.dropout(net, scope="Dropout")
            logits = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                       biases_initializer=V.zeros_initializer(), scope="Conv2d_1c_1x1")
            logits = logits.reshape(logits.shape[0], logits.shape[-1])
        end_points["Logits"] = logits
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


mobilenet_v2.default_image_size = 224
