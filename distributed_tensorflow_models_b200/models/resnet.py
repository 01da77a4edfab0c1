"""Pre-activation ResNet (v2) for CIFAR-10 (6n+2 layers) and ImageNet (18/34/50/101/152/200).

Behavioural counterpart of ``resnet/resnet_model.py`` (CIFAR generator :217-282, ImageNet generator
:285-349, size table :352-370, units :94-175, strided-conv padding rule :53-91, BN constants :36-50).

Design: a residual unit is described by its list of (kernel, filters, stride) convolutions — the basic unit
is ``[(3,f,s),(3,f,1)]``, the bottleneck ``[(1,f,1),(3,f,s),(1,4f,1)]`` — and one routine executes any
such list as  BN-ReLU -> [1x1 projection of the activated input] -> conv, (BN-ReLU -> conv)*, + shortcut.
On the native backend every "BN-ReLU -> conv" pair is a single tensor-core kernel: the batch statistics
are produced by the epilogue of whichever kernel wrote the tensor, and normalisation + ReLU are applied
while the next convolution gathers its A operand (``native_ops.bn_relu_conv``), so normalised
activations are never written to HBM.

Variable names are the tf.layers auto-names in creation order, un-prefixed (the reference's scope bug,
SURVEY Q2): ``conv2d[_k]/kernel``, ``batch_normalization[_k]/{gamma,beta,moving_mean,moving_variance}``,
``dense/{kernel,bias}``.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch

from ..ops import functional as Fn
from ..slim import variables as V

BATCH_NORM_DECAY = 0.997
BATCH_NORM_EPSILON = 1e-5

ConvSpec = Tuple[int, int, int]   # kernel, filters, stride


def _bn_vars(store: V.VariableStore, c: int):
    with store.scope(store.unique_name("batch_normalization")):
        gamma = store.variable("gamma", (c,), V.ones_initializer())
        beta = store.variable("beta", (c,), V.zeros_initializer())
        mm = store.variable("moving_mean", (c,), V.zeros_initializer(), trainable=False,
                            collections=[V.MOVING_AVERAGE_VARIABLES])
        mv = store.variable("moving_variance", (c,), V.ones_initializer(), trainable=False,
                            collections=[V.MOVING_AVERAGE_VARIABLES])
    return gamma, beta, mm, mv


def _conv_var(store: V.VariableStore, cin: int, k: int, filters: int) -> torch.Tensor:
    with store.scope(store.unique_name("conv2d")):
        return store.variable("kernel", (filters, k, k, cin), V.variance_scaling_initializer(),
                              tf_layout="hwio", tf_shape=(k, k, cin, filters), fan_in=k * k * cin,
                              fan_out=k * k * filters)


def _fixed_padding(k: int, stride: int):
    """Strided convs pad explicitly by (k-1)//2 / k//2 and run VALID; stride-1 convs are SAME
    (resnet_model.py:53-91).  For odd k both give symmetric (k-1)/2 padding."""
    if stride > 1:
        total = k - 1
        beg = total // 2
        return ((beg, total - beg), (beg, total - beg))
    return "SAME"


class _Ctx:
    """Per-forward settings shared by the unit routine."""
    def __init__(self, store, is_training, update_moving):
        self.store = store
        self.is_training = is_training
        self.update_moving = update_moving
        self.native = False


def _bn_relu(ctx: _Ctx, x: torch.Tensor) -> torch.Tensor:
    gamma, beta, mm, mv = _bn_vars(ctx.store, x.shape[-1])
    return Fn.batch_norm(x, gamma, beta, mm, mv, training=ctx.is_training, decay=BATCH_NORM_DECAY,
                         eps=BATCH_NORM_EPSILON, relu=True, update_moving=ctx.update_moving,
                         unbiased_moving_var=True)


def _conv(ctx: _Ctx, x: torch.Tensor, k: int, filters: int, stride: int) -> torch.Tensor:
    w = _conv_var(ctx.store, x.shape[-1], k, filters)
    return Fn.conv2d(x, w, None, stride, _fixed_padding(k, stride))


def residual_unit(ctx: _Ctx, x: torch.Tensor, convs: Sequence[ConvSpec], project: Optional[ConvSpec]) -> torch.Tensor:
    """One pre-activation unit (building_block :94-130 / bottleneck_block :133-175)."""
    if ctx.native:
        # x is a (tensor, channel sums) pair: statistics travel with the tensor from the kernel that wrote it
        from ..ops import native_ops
        return native_ops.resnet_unit(ctx, x, convs, project, _bn_vars, _conv_var, _fixed_padding,
                                      BATCH_NORM_DECAY, BATCH_NORM_EPSILON)
    shortcut = x
    y = _bn_relu(ctx, x)
    if project is not None:
        shortcut = _conv(ctx, y, *project)
    for i, (k, f, s) in enumerate(convs):
        if i > 0:
            y = _bn_relu(ctx, y)
        y = _conv(ctx, y, k, f, s)
    return y + shortcut


def _stage(ctx: _Ctx, x, filters: int, bottleneck: bool, blocks: int, stride: int):
    """block_layer (:178-214): the first unit projects the shortcut (always — also in stage 1)."""
    out_f = 4 * filters if bottleneck else filters

    def unit_convs(s):
        return [(1, filters, 1), (3, filters, s), (1, out_f, 1)] if bottleneck else [(3, filters, s), (3, filters, 1)]

    x = residual_unit(ctx, x, unit_convs(stride), (1, out_f, stride))
    for _ in range(1, blocks):
        x = residual_unit(ctx, x, unit_convs(1), None)
    return x


def _head(ctx: _Ctx, x, num_classes: int):
    if ctx.native:
        from ..ops import native_ops
        x, st = x
        gamma, beta, mm, mv = _bn_vars(ctx.store, x.shape[-1])
        y = native_ops.batch_norm(x, gamma, beta, mm, mv, True, BATCH_NORM_DECAY, BATCH_NORM_EPSILON, True,
                                  ctx.update_moving, True, stats=st)
    else:
        y = _bn_relu(ctx, x)
    y = Fn.global_avg_pool(y)
    store = ctx.store
    with store.scope(store.unique_name("dense")):
        nin = y.shape[-1]
        w = store.variable("kernel", (num_classes, nin), V.xavier_initializer(), tf_layout="io",
                           tf_shape=(nin, num_classes), fan_in=nin, fan_out=num_classes)
        b = store.variable("bias", (num_classes,), V.zeros_initializer())
    return Fn.linear(y, w, b)


def cifar10_resnet_v2_generator(resnet_size: int, num_classes: int, data_format=None) -> Callable:
    """model(inputs NHWC, is_training) -> logits (reference :217-282)."""
    if resnet_size % 6 != 2:
        raise ValueError("resnet_size must be 6n + 2:", resnet_size)
    num_blocks = (resnet_size - 2) // 6

    def model(inputs: torch.Tensor, is_training: bool, update_moving: bool = True) -> torch.Tensor:
        store = V.current_store()
        ctx = _Ctx(store, is_training, update_moving)
        ctx.native = bool(getattr(store, "fuse_bn", False)) and inputs.is_cuda and Fn.use_native(inputs) and is_training
        if ctx.native:
            from ..ops import native_ops
            w0 = _conv_var(store, inputs.shape[-1], 3, 16)
            st0 = native_ops._zeros(32, inputs.device)
            x = (native_ops.conv2d(inputs, w0, None, (1, 1), (1, 1), (inputs.shape[1], inputs.shape[2]), False,
                                   stats=st0), st0)
        else:
            x = _conv(ctx, inputs, 3, 16, 1)
        for filters, stride in ((16, 1), (32, 2), (64, 2)):
            x = _stage(ctx, x, filters, False, num_blocks, stride)
        return _head(ctx, x, num_classes)

    model.default_image_size = 32
    return model


_IMAGENET_SIZES = {
    18: (False, [2, 2, 2, 2]), 34: (False, [3, 4, 6, 3]), 50: (True, [3, 4, 6, 3]),
    101: (True, [3, 4, 23, 3]), 152: (True, [3, 8, 36, 3]), 200: (True, [3, 24, 36, 3]),
}


def imagenet_resnet_v2_generator(bottleneck: bool, layers: Sequence[int], num_classes: int, data_format=None):
    """model(inputs NHWC 224x224, is_training) -> logits (reference :285-349)."""
    def model(inputs: torch.Tensor, is_training: bool, update_moving: bool = True) -> torch.Tensor:
        store = V.current_store()
        ctx = _Ctx(store, is_training, update_moving)
        x = _conv(ctx, inputs, 7, 64, 2)
        x = Fn.max_pool(x, 3, 2, "SAME")
        for filters, blocks, stride in zip((64, 128, 256, 512), layers, (1, 2, 2, 2)):
            x = _stage(ctx, x, filters, bottleneck, blocks, stride)
        return _head(ctx, x, num_classes)

    model.default_image_size = 224
    return model


def resnet_v2(resnet_size: int, num_classes: int, data_format=None):
    """Size table (reference :352-370)."""
    if resnet_size not in _IMAGENET_SIZES:
        raise ValueError("Not a valid resnet_size:", resnet_size)
    bottleneck, layers = _IMAGENET_SIZES[resnet_size]
    return imagenet_resnet_v2_generator(bottleneck, layers, num_classes, data_format)
