"""Layer library with the reference slim call signatures, executing eagerly on the current
:class:`~.variables.VariableStore`.

Counterparts in the reference (``inception/slim/ops.py``): ``batch_norm`` :46-135, ``conv2d`` :170-249,
``fc`` :253-320, ``one_hot_encoding`` :323-341, ``max_pool`` :345-373, ``avg_pool`` :377-403,
``dropout`` :407-424, ``flatten`` :427-446, ``repeat_op`` :449-476.  Variable names produced:
``<scope|Conv[_k]>/weights``, ``/biases``, ``/BatchNorm/{beta,gamma,moving_mean,moving_variance}``,
FC default scope ``FC``.  conv2d uses a bias only when no batch-norm follows (ops.py:234-246).

The same functions also serve the ``tf.contrib.slim`` spellings used by the CIFAR models
(``slim.conv2d(..., weights_initializer=, biases_initializer=, activation_fn=)``) through the thin
aliases at the bottom of the file.

Everything computes through ``ops.functional`` (NHWC; reference or native sm_100a kernels).  When the
store is in *fused-BN* mode (``store.fuse_bn = True``, CUDA only) ``conv2d`` with ``batch_norm_params``
hands the conv+BN+ReLU triple to ``native_ops.conv_bn_relu`` so statistics come out of the conv epilogue.
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence, Union

import torch

from ..ops import functional as Fn
from . import losses
from . import quantize as Q
from . import variables as V
from .scopes import add_arg_scope, current_arg_scope

UPDATE_OPS_COLLECTION = V.UPDATE_OPS_COLLECTION

relu = Fn.relu


def _two(v) -> Sequence[int]:
    if isinstance(v, int):
        return [v, v]
    if len(v) == 1:
        return [int(v[0]), int(v[0])]
    return [int(v[0]), int(v[1])]


def _is_relu(activation) -> bool:
    return activation is relu or activation is torch.relu or activation is torch.nn.functional.relu


@add_arg_scope
def batch_norm(inputs: torch.Tensor, decay: float = 0.999, center: bool = True, scale: bool = False,
               epsilon: float = 0.001, moving_vars: str = "moving_vars", activation: Optional[Callable] = None,
               is_training: bool = True, trainable: bool = True, restore: bool = True,
               scope: Optional[str] = None, reuse=None, update_moving: bool = True,
               unbiased_moving_var: bool = False, names=("beta", "gamma", "moving_mean", "moving_variance")):
    """Batch normalisation over all but the last axis (reference ops.py:46-135)."""
    store = V.current_store()
    c = inputs.shape[-1]
    with store.scope(scope, "BatchNorm"):
        beta = gamma = None
        if center:
            beta = V.variable(names[0], (c,), initializer=V.zeros_initializer(), trainable=trainable, restore=restore)
        if scale:
            gamma = V.variable(names[1], (c,), initializer=V.ones_initializer(), trainable=trainable, restore=restore)
        colls = [moving_vars, V.MOVING_AVERAGE_VARIABLES]
        mm = V.variable(names[2], (c,), initializer=V.zeros_initializer(), trainable=False, restore=restore,
                        collections=colls)
        mv = V.variable(names[3], (c,), initializer=V.ones_initializer(), trainable=False, restore=restore,
                        collections=colls)
    x = inputs if inputs.dim() == 4 else inputs.reshape(inputs.shape[0], 1, 1, c)
    fuse_relu = _is_relu(activation)
    y = Fn.batch_norm(x, gamma, beta, mm, mv, training=is_training, decay=decay, eps=epsilon, relu=fuse_relu,
                      update_moving=update_moving, unbiased_moving_var=unbiased_moving_var)
    y = y.reshape(inputs.shape)
    if activation is not None and not fuse_relu:
        y = activation(y)
    return y


@add_arg_scope
def conv2d(inputs: torch.Tensor, num_filters_out: int, kernel_size, stride=1, padding="SAME",
           activation: Optional[Callable] = relu, stddev: float = 0.01, bias: float = 0.0,
           weight_decay: float = 0.0, batch_norm_params: Optional[dict] = None, is_training: bool = True,
           trainable: bool = True, restore: bool = True, scope: Optional[str] = None, reuse=None,
           weights_initializer: Optional[Callable] = None, biases_initializer: Optional[Callable] = None,
           use_bias: bool = True, weights_name: str = "weights", biases_name: str = "biases"):
    """2-D convolution (+BN | +bias) (+activation); NHWC in, NHWC out (reference ops.py:170-249)."""
    store = V.current_store()
    kh, kw = _two(kernel_size)
    sh, sw = _two(stride)
    cin = inputs.shape[-1]
    with store.scope(scope, "Conv"):
        w_init = weights_initializer or V.truncated_normal_initializer(stddev=stddev)
        reg = losses.l2_regularizer(weight_decay) if weight_decay and weight_decay > 0 else None
        w = V.variable(weights_name, (num_filters_out, kh, kw, cin), initializer=w_init, regularizer=reg,
                       trainable=trainable, restore=restore, tf_layout="hwio",
                       tf_shape=(kh, kw, cin, num_filters_out), fan_in=kh * kw * cin,
                       fan_out=kh * kw * num_filters_out)
        fuse_relu = _is_relu(activation)
        quant = Q.current()
        if quant is not None:                  # --quantize: folded batch norm + fake-quantised weights / output
            bn, b = None, None
            if batch_norm_params is not None:
                bn = dict(batch_norm_params, is_training=is_training)
            elif use_bias:
                b = V.variable(biases_name, (num_filters_out,),
                               initializer=biases_initializer or V.constant_initializer(bias), trainable=trainable,
                               restore=restore)
            return Q.folded_layer(quant, lambda wq: Fn.conv2d(inputs, wq, None, (sh, sw), padding), w, b,
                                  num_filters_out, (num_filters_out, 1, 1, 1), bn, activation, trainable)
        if batch_norm_params is not None:
            if getattr(store, "fuse_bn", False) and inputs.is_cuda and Fn.use_native(inputs) and is_training:
                from ..ops import native_ops
                bn = dict(batch_norm_params)
                with store.scope(bn.pop("scope", None), "BatchNorm"):
                    c = num_filters_out
                    beta = V.variable("beta", (c,), initializer=V.zeros_initializer(), trainable=trainable,
                                      restore=restore) if bn.get("center", True) else None
                    gamma = V.variable("gamma", (c,), initializer=V.ones_initializer(), trainable=trainable,
                                       restore=restore) if bn.get("scale", False) else None
                    colls = [bn.get("moving_vars", "moving_vars"), V.MOVING_AVERAGE_VARIABLES]
                    mm = V.variable("moving_mean", (c,), initializer=V.zeros_initializer(), trainable=False,
                                    restore=restore, collections=colls)
                    mv = V.variable("moving_variance", (c,), initializer=V.ones_initializer(), trainable=False,
                                    restore=restore, collections=colls)
                out = native_ops.conv_bn_act(inputs, w, (sh, sw), padding, gamma, beta, mm, mv,
                                             decay=bn.get("decay", 0.999), eps=bn.get("epsilon", 0.001),
                                             relu=fuse_relu)
                if activation is not None and not fuse_relu:
                    out = activation(out)
                return out
            y = Fn.conv2d(inputs, w, None, (sh, sw), padding)
            y = batch_norm(y, is_training=is_training, trainable=trainable, restore=restore,
                           activation=activation, **batch_norm_params)
            return y
        b = None
        if use_bias:
            b_init = biases_initializer or V.constant_initializer(bias)
            b = V.variable(biases_name, (num_filters_out,), initializer=b_init, trainable=trainable, restore=restore)
        y = Fn.conv2d(inputs, w, b, (sh, sw), padding, relu=fuse_relu)
        if activation is not None and not fuse_relu:
            y = activation(y)
        return y


@add_arg_scope
def fc(inputs: torch.Tensor, num_units_out: int, activation: Optional[Callable] = relu, stddev: float = 0.01,
       bias: float = 0.0, weight_decay: float = 0.0, batch_norm_params: Optional[dict] = None,
       is_training: bool = True, trainable: bool = True, restore: bool = True, scope: Optional[str] = None,
       reuse=None, weights_initializer: Optional[Callable] = None, biases_initializer: Optional[Callable] = None,
       weights_name: str = "weights", biases_name: str = "biases"):
    """Fully connected layer on [N, in] (reference ops.py:253-320)."""
    store = V.current_store()
    nin = inputs.shape[-1]
    with store.scope(scope, "FC"):
        w_init = weights_initializer or V.truncated_normal_initializer(stddev=stddev)
        reg = losses.l2_regularizer(weight_decay) if weight_decay and weight_decay > 0 else None
        w = V.variable(weights_name, (num_units_out, nin), initializer=w_init, regularizer=reg, trainable=trainable,
                       restore=restore, tf_layout="io", tf_shape=(nin, num_units_out), fan_in=nin,
                       fan_out=num_units_out)
        fuse_relu = _is_relu(activation)
        quant = Q.current()
        if quant is not None and batch_norm_params is None:
            b = V.variable(biases_name, (num_units_out,), initializer=biases_initializer or V.constant_initializer(bias),
                           trainable=trainable, restore=restore)
            return Q.folded_layer(quant, lambda wq: Fn.linear(inputs, wq, None), w, b, num_units_out,
                                  (num_units_out, 1), None, activation, trainable)
        if batch_norm_params is not None:
            y = Fn.linear(inputs, w, None)
            return batch_norm(y, is_training=is_training, trainable=trainable, restore=restore,
                              activation=activation, **batch_norm_params)
        b_init = biases_initializer or V.constant_initializer(bias)
        b = V.variable(biases_name, (num_units_out,), initializer=b_init, trainable=trainable, restore=restore)
        y = Fn.linear(inputs, w, b, relu=fuse_relu)
        if activation is not None and not fuse_relu:
            y = activation(y)
        return y


def one_hot_encoding(labels: torch.Tensor, num_classes: int, scope: Optional[str] = None) -> torch.Tensor:
    """Dense [N, num_classes] float one-hot (reference ops.py:323-341)."""
    return Fn.one_hot(labels, num_classes)


@add_arg_scope
def max_pool(inputs: torch.Tensor, kernel_size, stride=2, padding="VALID", scope: Optional[str] = None):
    """reference ops.py:345-373 (default stride 2, VALID)."""
    return Fn.max_pool(inputs, _two(kernel_size), _two(stride), padding)


@add_arg_scope
def avg_pool(inputs: torch.Tensor, kernel_size, stride=2, padding="VALID", scope: Optional[str] = None):
    """reference ops.py:377-403."""
    return Fn.avg_pool(inputs, _two(kernel_size), _two(stride), padding)


@add_arg_scope
def dropout(inputs: torch.Tensor, keep_prob: float = 0.5, is_training: bool = True, scope: Optional[str] = None):
    """Identity unless training and keep_prob < 1 (reference ops.py:407-424)."""
    if is_training and keep_prob > 0:
        return Fn.dropout(inputs, keep_prob, True)
    return inputs


def flatten(inputs: torch.Tensor, scope: Optional[str] = None) -> torch.Tensor:
    """[N, ...] -> [N, k] (reference ops.py:427-446)."""
    if inputs.dim() < 2:
        raise ValueError("Inputs must be have a least 2 dimensions")
    return Fn.flatten(inputs)


def repeat_op(repetitions: int, inputs, op: Callable, *args, **kwargs):
    """Apply ``op`` ``repetitions`` times, scopes ``<scope>_1 … _n`` (reference ops.py:449-476)."""
    store = V.current_store()
    scope = kwargs.pop("scope", None)
    with store.scope(scope, "RepeatOp"):
        tower = inputs
        for _ in range(repetitions):
            tower = op(tower, *args, **kwargs)
        return tower


def repeat(inputs, repetitions: int, layer: Callable, *args, **kwargs):
    """``tf.contrib.slim.repeat``: scopes ``<scope>/<scope>_1 …`` (vgg/nets/vgg.py:190 → conv1/conv1_1)."""
    store = V.current_store()
    scope = kwargs.pop("scope", None)
    with store.scope(scope, "Repeat"):
        out = inputs
        base = scope or getattr(layer, "__name__", None) or getattr(getattr(layer, "func", None), "__name__", "repeat")
        for i in range(repetitions):
            out = layer(out, *args, scope=f"{base}_{i + 1}", **kwargs)
        return out


# ---- tf.contrib.slim / tf.layers spellings used by the CIFAR models ---------------------------------
@add_arg_scope
def max_pool2d(inputs, kernel_size, stride=2, padding="VALID", scope=None):
    """slim.max_pool2d (cifarnet/cifarnet.py:66, alexnet/alexnet.py:106, vgg/nets/vgg.py:191)."""
    return Fn.max_pool(inputs, _two(kernel_size), _two(stride), padding)


@add_arg_scope
def avg_pool2d(inputs, kernel_size, stride=2, padding="VALID", scope=None):
    return Fn.avg_pool(inputs, _two(kernel_size), _two(stride), padding)


@add_arg_scope
def fully_connected(inputs, num_outputs, activation_fn=relu, weights_initializer=None, biases_initializer=None,
                    weights_regularizer=None, scope=None, trainable=True):
    """slim.fully_connected: Xavier weights / zero biases by default (cifarnet/cifarnet.py:76-89)."""
    wd = getattr(weights_regularizer, "l2", 0.0) if weights_regularizer is not None else 0.0
    return fc(inputs, num_outputs, activation=activation_fn, weight_decay=wd,
              weights_initializer=weights_initializer or V.xavier_initializer(),
              biases_initializer=biases_initializer or V.zeros_initializer(), scope=scope, trainable=trainable)


def _normalize_and_activate(y, normalizer_fn, normalizer_params, activation_fn, trainable=True):
    """tf.contrib.layers post-processing order: normaliser (which replaces the bias), then activation."""
    if normalizer_fn is not None:
        params = dict(normalizer_params or {})
        if normalizer_fn is batch_norm and _is_relu(activation_fn):
            return batch_norm(y, activation=activation_fn, trainable=trainable, **params)
        y = normalizer_fn(y, **params)
    if activation_fn is not None:
        y = activation_fn(y)
    return y


def _merged_bn_params(normalizer_fn, normalizer_params) -> Optional[dict]:
    """Batch-norm arguments in force for a layer: the enclosing ``arg_scope([batch_norm], ...)`` defaults, then the
    layer's own ``normalizer_params``; None when the layer has no normaliser."""
    if normalizer_fn is None:
        return None
    params = dict(current_arg_scope().get(batch_norm._dtf_arg_scope_key, {}))
    params.update(normalizer_params or {})
    return params


@add_arg_scope
def convolution2d(inputs, num_outputs, kernel_size, stride=1, padding="SAME", activation_fn=relu,
                  weights_initializer=None, biases_initializer=None, weights_regularizer=None, scope=None,
                  trainable=True, normalizer_fn=None, normalizer_params=None, rate=1, use_bias=True):
    """slim.conv2d in its tf.contrib form (cifarnet/cifarnet.py:64, alexnet/alexnet.py:103, vgg.py:190; with
    ``normalizer_fn`` / ``rate`` throughout the zoo, e.g. vgg/nets/resnet_utils.py:123, mobilenet_v1.py:255)."""
    wd = getattr(weights_regularizer, "l2", 0.0) if weights_regularizer is not None else 0.0
    w_init = weights_initializer or V.xavier_initializer()
    quant = Q.current()
    if quant is not None and (normalizer_fn is None or normalizer_fn is batch_norm):
        store = V.current_store()
        kh, kw = _two(kernel_size)
        cin = inputs.shape[-1]
        with store.scope(scope, "Conv"):
            w = V.variable("weights", (num_outputs, kh, kw, cin), initializer=w_init,
                           regularizer=losses.l2_regularizer(wd) if wd > 0 else None, trainable=trainable,
                           tf_layout="hwio", tf_shape=(kh, kw, cin, num_outputs), fan_in=kh * kw * cin,
                           fan_out=kh * kw * num_outputs)
            b = None
            if normalizer_fn is None and use_bias:
                b = V.variable("biases", (num_outputs,), initializer=biases_initializer or V.zeros_initializer(),
                               trainable=trainable)
            if rate != 1:
                run = lambda wq: Fn.atrous_conv2d(inputs, wq, None, rate, padding)
            else:
                run = lambda wq: Fn.conv2d(inputs, wq, None, _two(stride), padding)
            return Q.folded_layer(quant, run, w, b, num_outputs, (num_outputs, 1, 1, 1),
                                  _merged_bn_params(normalizer_fn, normalizer_params), activation_fn, trainable)
    if (normalizer_fn is None or normalizer_fn is batch_norm) and rate == 1:
        bn_params, is_tr = None, True
        if normalizer_fn is not None:
            # defaults that an enclosing arg_scope([batch_norm], ...) put in force, then the explicit parameters
            bn_params = dict(current_arg_scope().get(batch_norm._dtf_arg_scope_key, {}))
            bn_params.update(normalizer_params or {})
            is_tr = bn_params.pop("is_training", True)
            bn_params.pop("activation", None)
        return conv2d(inputs, num_outputs, kernel_size, stride=stride, padding=padding, activation=activation_fn,
                      weight_decay=wd, weights_initializer=w_init,
                      biases_initializer=biases_initializer or V.zeros_initializer(), scope=scope,
                      trainable=trainable, is_training=is_tr, batch_norm_params=bn_params, use_bias=use_bias)
    store = V.current_store()
    kh, kw = _two(kernel_size)
    cin = inputs.shape[-1]
    with store.scope(scope, "Conv"):
        reg = losses.l2_regularizer(wd) if wd > 0 else None
        w = V.variable("weights", (num_outputs, kh, kw, cin), initializer=w_init, regularizer=reg,
                       trainable=trainable, tf_layout="hwio", tf_shape=(kh, kw, cin, num_outputs),
                       fan_in=kh * kw * cin, fan_out=kh * kw * num_outputs)
        b = None
        if normalizer_fn is None and use_bias:
            b = V.variable("biases", (num_outputs,), initializer=biases_initializer or V.zeros_initializer(),
                           trainable=trainable)
        if rate != 1:
            y = Fn.atrous_conv2d(inputs, w, b, rate, padding)
        else:
            y = Fn.conv2d(inputs, w, b, _two(stride), padding)
        return _normalize_and_activate(y, normalizer_fn, normalizer_params, activation_fn, trainable)


@add_arg_scope
def separable_convolution2d(inputs, num_outputs, kernel_size, depth_multiplier=1, stride=1, padding="SAME", rate=1,
                            activation_fn=relu, normalizer_fn=None, normalizer_params=None,
                            weights_initializer=None, weights_regularizer=None, biases_initializer=None, scope=None,
                            trainable=True, use_bias=True):
    """slim.separable_conv2d: a depthwise convolution (``depthwise_weights`` [kh, kw, C, multiplier]) followed,
    when ``num_outputs`` is not None, by a 1x1 convolution (``pointwise_weights``); the stride belongs to the
    depthwise stage; bias / normaliser / activation come once, at the end
    (vgg/nets/mobilenet_v1.py:260-277, mobilenet/conv_blocks.py:120-140, nasnet/nasnet_utils.py:206-240)."""
    store = V.current_store()
    kh, kw = _two(kernel_size)
    cin = inputs.shape[-1]
    wd = getattr(weights_regularizer, "l2", 0.0) if weights_regularizer is not None else 0.0
    reg = losses.l2_regularizer(wd) if wd > 0 else None
    w_init = weights_initializer or V.xavier_initializer()
    with store.scope(scope, "SeparableConv2d"):
        dw = V.variable("depthwise_weights", (kh, kw, cin, depth_multiplier), initializer=w_init, regularizer=reg,
                        trainable=trainable, fan_in=kh * kw * cin, fan_out=kh * kw * depth_multiplier)
        cmid = cin * depth_multiplier
        cout = cmid if num_outputs is None else num_outputs
        pw = None
        if num_outputs is not None:
            pw = V.variable("pointwise_weights", (cout, 1, 1, cmid), initializer=w_init, regularizer=reg,
                            trainable=trainable, tf_layout="hwio", tf_shape=(1, 1, cmid, cout), fan_in=cmid,
                            fan_out=cout)
        b = None
        if normalizer_fn is None and use_bias:
            b = V.variable("biases", (cout,), initializer=biases_initializer or V.zeros_initializer(),
                           trainable=trainable)
        quant = Q.current()
        if quant is not None and (normalizer_fn is None or normalizer_fn is batch_norm):
            bn = _merged_bn_params(normalizer_fn, normalizer_params)
            run_dw = lambda wq: Fn.depthwise_conv2d(inputs, wq, None, _two(stride), padding, rate=rate)
            if pw is None:
                return Q.folded_layer(quant, run_dw, dw, b, cmid, (1, 1, cin, depth_multiplier), bn, activation_fn,
                                      trainable)
            # depthwise stage: weights only (no normaliser / activation of its own), then the quantised 1x1 layer
            y = run_dw(Q.quantize_weights(quant, dw, quant.is_training if bn is None else bn.get("is_training", True),
                                          "depthwise_weights_quant"))
            run_pw = lambda wq: Fn.conv2d(y, wq, None, (1, 1), "VALID")
            return Q.folded_layer(quant, run_pw, pw, b, cout, (cout, 1, 1, 1), bn, activation_fn, trainable,
                                  weights_quant_name="pointwise_weights_quant")
        y = Fn.depthwise_conv2d(inputs, dw, b if pw is None else None, _two(stride), padding, rate=rate)
        if pw is not None:
            y = Fn.conv2d(y, pw, b, (1, 1), "VALID")
        return _normalize_and_activate(y, normalizer_fn, normalizer_params, activation_fn, trainable)


@add_arg_scope
def convolution2d_transpose(inputs, num_outputs, kernel_size, stride=1, padding="SAME", activation_fn=relu,
                            normalizer_fn=None, normalizer_params=None, weights_initializer=None,
                            weights_regularizer=None, biases_initializer=None, scope=None, trainable=True):
    """slim.conv2d_transpose; ``weights`` is [kh, kw, C_out, C_in] in TensorFlow's layout
    (vgg/nets/dcgan.py:160-181, pix2pix.py:90, cyclegan.py:110)."""
    store = V.current_store()
    kh, kw = _two(kernel_size)
    cin = inputs.shape[-1]
    wd = getattr(weights_regularizer, "l2", 0.0) if weights_regularizer is not None else 0.0
    with store.scope(scope, "Conv2d_transpose"):
        w = V.variable("weights", (num_outputs, kh, kw, cin), initializer=weights_initializer or V.xavier_initializer(),
                       regularizer=losses.l2_regularizer(wd) if wd > 0 else None, trainable=trainable,
                       tf_layout="hwoi", tf_shape=(kh, kw, num_outputs, cin), fan_in=kh * kw * cin,
                       fan_out=kh * kw * num_outputs)
        b = None
        if normalizer_fn is None:
            b = V.variable("biases", (num_outputs,), initializer=biases_initializer or V.zeros_initializer(),
                           trainable=trainable)
        y = Fn.conv2d_transpose(inputs, w, b, _two(stride), padding)
        return _normalize_and_activate(y, normalizer_fn, normalizer_params, activation_fn, trainable)


@add_arg_scope
def instance_norm(inputs, center=True, scale=True, epsilon=1e-6, activation_fn=None, trainable=True, scope=None):
    """tf.contrib.layers.instance_norm: variables ``InstanceNorm/{beta,gamma}`` (vgg/nets/cyclegan.py:47-60)."""
    store = V.current_store()
    c = inputs.shape[-1]
    with store.scope(scope, "InstanceNorm"):
        beta = V.variable("beta", (c,), initializer=V.zeros_initializer(), trainable=trainable) if center else None
        gamma = V.variable("gamma", (c,), initializer=V.ones_initializer(), trainable=trainable) if scale else None
    y = Fn.instance_norm(inputs, gamma, beta, epsilon)
    return activation_fn(y) if activation_fn is not None else y


relu6 = Fn.relu6
leaky_relu = Fn.leaky_relu
tanh = Fn.tanh
sigmoid = Fn.sigmoid
separable_conv2d = separable_convolution2d
conv2d_transpose = convolution2d_transpose


def softmax(logits: torch.Tensor, scope=None) -> torch.Tensor:
    return torch.softmax(logits, dim=-1)
