"""The zoo's bottleneck ResNets: ``resnet_v1_{50,101,152,200}`` and ``resnet_v2_{50,101,152,200}``.

Behaviour follows the reference's ``vgg/nets/resnet_utils.py`` (``Block`` :47-58, ``subsample`` :61-75, ``conv2d_same``
:78-123, ``stack_blocks_dense`` :126-218, ``resnet_arg_scope`` :221-272), ``resnet_v1.py`` (``bottleneck`` :76-134,
``resnet_v1`` :137-246, depth tables :268-375) and ``resnet_v2.py`` (pre-activation ``bottleneck`` :63-122, ``resnet_v2``
:125-233, tables :255-337):

* a block is ``num_units`` bottleneck units; the block's stride sits in its **last** unit (so the dense-prediction
  mode can swap it for dilation), block 4 has stride 1;
* strided 3x3 / 7x7 convolutions pad explicitly by ``k-1`` (split floor / ceil) and run VALID, so the output does
  not depend on whether the input extent is even;
* identity shortcuts are sub-sampled with a 1x1 max-pool of stride ``s``;
* v1 unit: conv-BN-ReLU x2, conv-BN, add, ReLU.  v2 unit: shared BN-ReLU ``preact``; the projection shortcut and the
  last 1x1 carry biases instead of BN; ``postnorm`` BN-ReLU after the last block; the root convolution has a bias
  and no BN;
* ``output_stride`` replaces strides by atrous rates once the running stride reaches the target.

Variable names: ``resnet_v1_50/conv1/{weights,BatchNorm/...}``, ``resnet_v1_50/block<b>/unit_<u>/bottleneck_v1/
{shortcut,conv1,conv2,conv3}/...``, ``resnet_v1_50/logits/{weights,biases}`` (v2: ``bottleneck_v2``, ``preact``,
``postnorm``).  (The CIFAR ResNet the reference actually trains lives in ``models/resnet.py``.)
"""
from __future__ import annotations

import contextlib
from typing import Dict, List, NamedTuple, Optional

from ..ops import functional as Fn
from ..slim import losses, ops, quantize, scopes
from ..slim import variables as V


class Block(NamedTuple):
    scope: str
    version: int               # 1 or 2
    units: List[dict]          # dicts of depth, depth_bottleneck, stride


@contextlib.contextmanager
def resnet_arg_scope(weight_decay: float = 0.0001, batch_norm_decay: float = 0.997, batch_norm_epsilon: float = 1e-5,
                     batch_norm_scale: bool = True, activation_fn=ops.relu, use_batch_norm: bool = True):
    bn = {"decay": batch_norm_decay, "epsilon": batch_norm_epsilon, "scale": batch_norm_scale}
    with scopes.arg_scope([ops.convolution2d], weights_regularizer=losses.l2_regularizer(weight_decay),
                          weights_initializer=V.variance_scaling_initializer(), activation_fn=activation_fn,
                          normalizer_fn=ops.batch_norm if use_batch_norm else None, normalizer_params=bn):
        with scopes.arg_scope([ops.batch_norm], **bn):
            with scopes.arg_scope([ops.max_pool2d], padding="SAME") as sc:
                yield sc


def subsample(inputs, factor: int, scope: Optional[str] = None):
    return inputs if factor == 1 else ops.max_pool2d(inputs, [1, 1], stride=factor, scope=scope)


def conv2d_same(inputs, num_outputs: int, kernel_size: int, stride: int, rate: int = 1, scope: Optional[str] = None,
                **kw):
    """'SAME' for stride 1; otherwise explicit symmetric-ish padding + 'VALID' (resnet_utils.py:78-123)."""
    if stride == 1:
        return ops.convolution2d(inputs, num_outputs, kernel_size, stride=1, rate=rate, padding="SAME", scope=scope, **kw)
    k_eff = kernel_size + (kernel_size - 1) * (rate - 1)
    total = k_eff - 1
    beg = total // 2
    inputs = Fn.pad_spatial(inputs, (beg, total - beg, beg, total - beg))
    return ops.convolution2d(inputs, num_outputs, kernel_size, stride=stride, rate=rate, padding="VALID", scope=scope,
                             **kw)


def bottleneck_v1(inputs, depth: int, depth_bottleneck: int, stride: int, rate: int = 1, scope: Optional[str] = None):
    store = V.current_store()
    with store.scope(scope, "bottleneck_v1"):
        if depth == inputs.shape[-1]:
            shortcut = subsample(inputs, stride, "shortcut")
        else:
            shortcut = ops.convolution2d(inputs, depth, [1, 1], stride=stride, activation_fn=None, scope="shortcut")
        residual = ops.convolution2d(inputs, depth_bottleneck, [1, 1], stride=1, scope="conv1")
        residual = conv2d_same(residual, depth_bottleneck, 3, stride, rate=rate, scope="conv2")
        residual = ops.convolution2d(residual, depth, [1, 1], stride=1, activation_fn=None, scope="conv3")
        return quantize.after_add(Fn.relu(shortcut + residual))


def bottleneck_v2(inputs, depth: int, depth_bottleneck: int, stride: int, rate: int = 1, scope: Optional[str] = None):
    store = V.current_store()
    with store.scope(scope, "bottleneck_v2"):
        preact = ops.batch_norm(inputs, activation=ops.relu, scope="preact")
        if depth == inputs.shape[-1]:
            shortcut = subsample(inputs, stride, "shortcut")
        else:
            shortcut = ops.convolution2d(preact, depth, [1, 1], stride=stride, normalizer_fn=None, activation_fn=None,
                                         scope="shortcut")
        residual = ops.convolution2d(preact, depth_bottleneck, [1, 1], stride=1, scope="conv1")
        residual = conv2d_same(residual, depth_bottleneck, 3, stride, rate=rate, scope="conv2")
        residual = ops.convolution2d(residual, depth, [1, 1], stride=1, normalizer_fn=None, activation_fn=None,
                                     scope="conv3")
        return quantize.after_add(shortcut + residual)


def stack_blocks_dense(net, blocks: List[Block], output_stride: Optional[int], end_points: dict, prefix: str):
    store = V.current_store()
    current_stride, rate = 1, 1
    for block in blocks:
        unit_fn = bottleneck_v1 if block.version == 1 else bottleneck_v2
        kind = "bottleneck_v1" if block.version == 1 else "bottleneck_v2"
        with store.scope(block.scope):
            for i, unit in enumerate(block.units):
                if output_stride is not None and current_stride > output_stride:
                    raise ValueError("The target output_stride cannot be reached.")
                with store.scope(f"unit_{i + 1}"):
                    if output_stride is not None and current_stride == output_stride:
                        net = unit_fn(net, unit["depth"], unit["depth_bottleneck"], 1, rate=rate, scope=kind)
                        rate *= unit["stride"]
                    else:
                        net = unit_fn(net, unit["depth"], unit["depth_bottleneck"], unit["stride"], rate=1, scope=kind)
                        current_stride *= unit["stride"]
                end_points[f"{prefix}/{block.scope}/unit_{i + 1}/{kind}"] = net
        end_points[f"{prefix}/{block.scope}"] = net
    if output_stride is not None and current_stride != output_stride:
        raise ValueError("The target output_stride cannot be reached.")
    return net


def _block(scope: str, version: int, base_depth: int, num_units: int, stride: int) -> Block:
    unit = {"depth": base_depth * 4, "depth_bottleneck": base_depth, "stride": 1}
    return Block(scope, version, [dict(unit) for _ in range(num_units - 1)] + [dict(unit, stride=stride)])


def _resnet(version: int, inputs, blocks: List[Block], num_classes: Optional[int], is_training: Optional[bool],
            global_pool: bool, output_stride: Optional[int], include_root_block: bool, spatial_squeeze: bool,
            scope: str):
    store = V.current_store()
    end_points: Dict[str, object] = {}
    bn_scope = scopes.arg_scope([ops.batch_norm], is_training=is_training) if is_training is not None \
        else contextlib.nullcontext()
    with store.scope(scope), bn_scope:
        net = inputs
        if include_root_block:
            if output_stride is not None:
                if output_stride % 4 != 0:
                    raise ValueError("The output_stride needs to be a multiple of 4.")
                output_stride //= 4
            if version == 1:
                net = conv2d_same(net, 64, 7, stride=2, scope="conv1")
            else:       # v2: plain convolution with bias; the first unit's preact normalises
                net = conv2d_same(net, 64, 7, stride=2, scope="conv1", activation_fn=None, normalizer_fn=None)
            end_points[f"{scope}/conv1"] = net
            net = ops.max_pool2d(net, [3, 3], stride=2, scope="pool1")
        net = stack_blocks_dense(net, blocks, output_stride, end_points, scope)
        if version == 2:
            net = ops.batch_norm(net, activation=ops.relu, scope="postnorm")
        if global_pool:
            net = end_points["global_pool"] = net.mean((1, 2), keepdim=True)
        if num_classes:
            net = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None, scope="logits")
            end_points[f"{scope}/logits"] = net
            if spatial_squeeze and net.shape[1] == 1 and net.shape[2] == 1:
                net = net.reshape(net.shape[0], net.shape[-1])
                end_points[f"{scope}/spatial_squeeze"] = net
            end_points["predictions"] = ops.softmax(net)
    store.end_points = end_points
    return net, end_points


_DEPTHS = {50: (3, 4, 6, 3), 101: (3, 4, 23, 3), 152: (3, 8, 36, 3), 200: (3, 24, 36, 3)}


def _make(version: int, size: int):
    name = f"resnet_v{version}_{size}"

    def net_fn(inputs, num_classes=None, is_training=True, global_pool=True, output_stride=None,
               spatial_squeeze=True, include_root_block=True, scope=name):
        n = _DEPTHS[size]
        blocks = [_block("block1", version, 64, n[0], 2), _block("block2", version, 128, n[1], 2),
                  _block("block3", version, 256, n[2], 2), _block("block4", version, 512, n[3], 1)]
        return _resnet(version, inputs, blocks, num_classes, is_training, global_pool, output_stride,
                       include_root_block, spatial_squeeze, scope)

    net_fn.__name__ = name
    net_fn.default_image_size = 224
    return net_fn


resnet_v1_50, resnet_v1_101, resnet_v1_152, resnet_v1_200 = (_make(1, s) for s in (50, 101, 152, 200))
resnet_v2_50, resnet_v2_101, resnet_v2_152, resnet_v2_200 = (_make(2, s) for s in (50, 101, 152, 200))
