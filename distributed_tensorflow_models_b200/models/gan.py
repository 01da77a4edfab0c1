"""The zoo's image-to-image / generative networks: DCGAN, pix2pix and the CycleGAN ResNet generator.

They are the only users of transposed convolution, instance normalisation, leaky-ReLU, tanh, nearest / bilinear
up-sampling and reflect padding, which here run on ``csrc/zoo_ops.cu`` and on the tcgen05 input-gradient kernel
(``ops/zoo_native.py``).  Behaviour per the reference:

* ``dcgan.discriminator`` / ``dcgan.generator`` — ``vgg/nets/dcgan.py:45-113`` / ``:119-202``: log2(size) stride-2
  4x4 convolutions with leaky-ReLU (BN from the second on) and a 1x1 logit; a 4x4 VALID transposed convolution from
  the noise vector, then stride-2 4x4 transposed convolutions with BN + ReLU, the last one linear, and a 1x1 head.
* ``pix2pix_generator`` / ``pix2pix_discriminator`` — ``vgg/nets/pix2pix.py:110-223`` / ``:226-292``: U-Net of
  4x4/2 convolutions (instance norm, leaky-ReLU) with skip concatenations and ``upsample`` (``:62-95``) by
  nearest-neighbour resize + 4x4 convolution or by transposed convolution; PatchGAN discriminator with reflect padding.
* ``cyclegan_generator_resnet`` — ``vgg/nets/cyclegan.py:117-273``: 7x7 reflect-padded stem, two stride-2
  encoders, ``num_resnet_blocks`` residual blocks, two up-sampling decoders (``cyclegan_upsample`` ``:66-114``),
  7x7 head, ``tanh(logits) + slope * logits``.
"""
from __future__ import annotations

import contextlib
import math
from typing import Dict, List, NamedTuple, Optional, Sequence

import torch

from ..ops import functional as Fn
from ..slim import losses, ops, scopes
from ..slim import variables as V


def _is_pow2(n: int) -> bool:
    return n > 0 and (n & (n - 1)) == 0


# ---- DCGAN -------------------------------------------------------------------------------------------------
def dcgan_discriminator(inputs, depth: int = 64, is_training: bool = True, scope: str = "Discriminator"):
    """-> (logits [N, 1], end_points).  ``inputs`` must be square with a power-of-two side."""
    if inputs.dim() != 4:
        raise ValueError("expected a rank-4 image tensor")
    if inputs.shape[1] != inputs.shape[2]:
        raise ValueError("Input tensor does not have equal width and height: %s" % (tuple(inputs.shape[1:3]),))
    size = inputs.shape[1]
    if not _is_pow2(size):
        raise ValueError("Input tensor `width` is not a power of 2: %d" % size)
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with store.scope(scope), scopes.arg_scope([ops.batch_norm], is_training=is_training), \
            scopes.arg_scope([ops.convolution2d], stride=2, kernel_size=4, activation_fn=ops.leaky_relu):
        net = inputs
        for i in range(int(math.log2(size))):
            name = "conv%i" % (i + 1)
            net = end_points[name] = ops.convolution2d(net, depth * 2 ** i,
                                                       normalizer_fn=None if i == 0 else ops.batch_norm, scope=name)
        logits = ops.convolution2d(net, 1, kernel_size=1, stride=1, padding="VALID", normalizer_fn=None,
                                   activation_fn=None)
        logits = end_points["logits"] = logits.reshape(-1, 1)
    return logits, end_points


def dcgan_generator(inputs, depth: int = 64, final_size: int = 32, num_outputs: int = 3, is_training: bool = True,
                    scope: str = "Generator"):
    """``inputs`` [N, noise] -> (logits [N, final_size, final_size, num_outputs], end_points)."""
    if inputs.dim() != 2:
        raise ValueError("expected a rank-2 noise tensor")
    if not _is_pow2(final_size):
        raise ValueError("`final_size` (%i) must be a power of 2." % final_size)
    if final_size < 8:
        raise ValueError("`final_size` (%i) must be greater than 8." % final_size)
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    num_layers = int(math.log2(final_size)) - 1
    with store.scope(scope), scopes.arg_scope([ops.batch_norm], is_training=is_training), \
            scopes.arg_scope([ops.convolution2d_transpose], normalizer_fn=ops.batch_norm, stride=2, kernel_size=4):
        net = inputs.reshape(inputs.shape[0], 1, 1, inputs.shape[1])
        net = end_points["deconv1"] = ops.convolution2d_transpose(net, depth * 2 ** (num_layers - 1), stride=1,
                                                                  padding="VALID", scope="deconv1")
        for i in range(2, num_layers):
            name = "deconv%i" % i
            net = end_points[name] = ops.convolution2d_transpose(net, depth * 2 ** (num_layers - i), scope=name)
        name = "deconv%i" % num_layers
        net = end_points[name] = ops.convolution2d_transpose(net, depth, normalizer_fn=None, activation_fn=None,
                                                             scope=name)
        logits = end_points["logits"] = ops.convolution2d(net, num_outputs, kernel_size=1, stride=1, padding="VALID",
                                                          normalizer_fn=None, activation_fn=None, scope="logits")
    assert tuple(logits.shape[1:]) == (final_size, final_size, num_outputs)
    return logits, end_points


# ---- pix2pix -----------------------------------------------------------------------------------------------
@contextlib.contextmanager
def pix2pix_arg_scope():
    """Instance norm (centre + scale, eps 1e-5) and N(0, 0.02) weights on conv2d / conv2d_transpose."""
    params = {"center": True, "scale": True, "epsilon": 0.00001}
    with scopes.arg_scope([ops.convolution2d, ops.convolution2d_transpose], normalizer_fn=ops.instance_norm,
                          normalizer_params=params, weights_initializer=V.random_normal_initializer(0, 0.02)) as sc:
        yield sc


def pix2pix_upsample(net, num_outputs: int, kernel_size: Sequence[int], method: str = "nn_upsample_conv"):
    if method == "nn_upsample_conv":
        net = Fn.resize_nearest(net, (kernel_size[0] * net.shape[1], kernel_size[1] * net.shape[2]))
        return ops.convolution2d(net, num_outputs, [4, 4], activation_fn=None)
    if method == "conv2d_transpose":
        return ops.convolution2d_transpose(net, num_outputs, [4, 4], stride=list(kernel_size), activation_fn=None)
    raise ValueError("Unknown method: [%s]" % method)


class Pix2PixBlock(NamedTuple):
    num_filters: int
    decoder_keep_prob: float


def pix2pix_default_blocks() -> List[Pix2PixBlock]:
    return [Pix2PixBlock(64, 0.5), Pix2PixBlock(128, 0.5), Pix2PixBlock(256, 0.5)] + [Pix2PixBlock(512, 0)] * 4


def pix2pix_generator(net, num_outputs: int, blocks: Optional[List[Pix2PixBlock]] = None,
                      upsample_method: str = "nn_upsample_conv", is_training: bool = False):
    """U-Net generator; the decoder's dropout is always active (it is pix2pix's noise source)."""
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    blocks = blocks or pix2pix_default_blocks()
    out_shape = (net.shape[0], net.shape[1], net.shape[2], num_outputs)
    skips = []
    with store.scope("encoder"), scopes.arg_scope([ops.convolution2d], kernel_size=[4, 4], stride=2,
                                                  activation_fn=ops.leaky_relu):
        for i, block in enumerate(blocks):
            if i == 0:
                net = ops.convolution2d(net, block.num_filters, normalizer_fn=None)
            elif i < len(blocks) - 1:
                net = ops.convolution2d(net, block.num_filters)
            else:
                net = ops.convolution2d(net, block.num_filters, activation_fn=None, normalizer_fn=None)
            skips.append(net)
            end_points["encoder%d" % i] = net
    with store.scope("decoder"), scopes.arg_scope([ops.dropout], is_training=True):
        for i, block in enumerate(reversed(blocks)):
            if i > 0:
                net = torch.cat([net, skips[-i - 1]], dim=3)
            net = Fn.relu(net)
            net = pix2pix_upsample(net, block.num_filters, [2, 2], upsample_method)
            if block.decoder_keep_prob > 0:
                net = ops.dropout(net, keep_prob=block.decoder_keep_prob)
            end_points["decoder%d" % i] = net
    with store.scope("output"):
        logits = ops.convolution2d(net, num_outputs, [4, 4], activation_fn=None, normalizer_fn=None)
        logits = end_points["logits"] = logits.reshape(out_shape)
        end_points["predictions"] = Fn.tanh(logits)
    return logits, end_points


def pix2pix_discriminator(net, num_filters: Sequence[int], padding: int = 2, is_training: bool = False):
    """PatchGAN: 4x4/2 VALID convolutions on reflect-padded maps -> (logits, end_points)."""
    end_points: Dict[str, torch.Tensor] = {}
    n = len(num_filters)

    def padded(t):
        return Fn.pad_spatial(t, (padding,) * 4, "reflect") if padding else t

    with scopes.arg_scope([ops.convolution2d], kernel_size=[4, 4], stride=2, padding="VALID",
                          activation_fn=ops.leaky_relu):
        net = end_points["conv0"] = ops.convolution2d(padded(net), num_filters[0], normalizer_fn=None, scope="conv0")
        for i in range(1, n - 1):
            net = end_points["conv%d" % i] = ops.convolution2d(padded(net), num_filters[i], scope="conv%d" % i)
        net = end_points["conv%d" % (n - 1)] = ops.convolution2d(padded(net), num_filters[-1], stride=1,
                                                                 scope="conv%d" % (n - 1))
        logits = end_points["logits"] = ops.convolution2d(padded(net), 1, stride=1, activation_fn=None,
                                                          normalizer_fn=None, scope="conv%d" % n)
        end_points["predictions"] = Fn.sigmoid(logits)
    return logits, end_points


# ---- CycleGAN ----------------------------------------------------------------------------------------------
@contextlib.contextmanager
def cyclegan_arg_scope(instance_norm_center: bool = True, instance_norm_scale: bool = True,
                       instance_norm_epsilon: float = 0.001, weights_init_stddev: float = 0.02,
                       weight_decay: float = 0.0):
    params = {"center": instance_norm_center, "scale": instance_norm_scale, "epsilon": instance_norm_epsilon}
    reg = losses.l2_regularizer(weight_decay) if weight_decay and weight_decay > 0.0 else None
    with scopes.arg_scope([ops.convolution2d], normalizer_fn=ops.instance_norm, normalizer_params=params,
                          weights_initializer=V.random_normal_initializer(0, weights_init_stddev),
                          weights_regularizer=reg) as sc:
        yield sc


def cyclegan_upsample(net, num_outputs: int, stride: Sequence[int], method: str = "conv2d_transpose"):
    """x``stride`` up-sampling: resize (nearest | bilinear) + reflect-pad + 3x3 VALID conv, or a 3x3 VALID
    transposed convolution whose first row / column is dropped so that it lines up with a 'SAME' convolution."""
    store = V.current_store()
    with store.scope("upconv"):
        size = (stride[0] * net.shape[1], stride[1] * net.shape[2])
        if method == "nn_upsample_conv":
            net = Fn.pad_spatial(Fn.resize_nearest(net, size), (1, 1, 1, 1), "reflect")
            return ops.convolution2d(net, num_outputs, kernel_size=[3, 3], padding="VALID")
        if method == "bilinear_upsample_conv":
            net = Fn.pad_spatial(Fn.resize_bilinear(net, size), (1, 1, 1, 1), "reflect")
            return ops.convolution2d(net, num_outputs, kernel_size=[3, 3], padding="VALID")
        if method == "conv2d_transpose":
            net = ops.convolution2d_transpose(net, num_outputs, kernel_size=[3, 3], stride=list(stride),
                                              padding="VALID")
            return net[:, 1:, 1:, :]
        raise ValueError("Unknown method: [%s]" % method)


def cyclegan_generator_resnet(images, arg_scope_fn=cyclegan_arg_scope, num_resnet_blocks: int = 6,
                              num_filters: int = 64, upsample_fn=cyclegan_upsample, kernel_size=3,
                              num_outputs: int = 3, tanh_linear_slope: float = 0.0, is_training: bool = False):
    """-> (predictions, end_points); height and width must be multiples of 4."""
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    h, w = images.shape[1], images.shape[2]
    if h % 4 != 0:
        raise ValueError("The input height must be a multiple of 4.")
    if w % 4 != 0:
        raise ValueError("The input width must be a multiple of 4.")
    kh, kw = (kernel_size, kernel_size) if isinstance(kernel_size, int) else kernel_size
    pads = ((kh - 1) // 2, kh // 2, (kw - 1) // 2, kw // 2)
    reflect = lambda t, p: Fn.pad_spatial(t, p, "reflect")
    with arg_scope_fn():
        with store.scope("input"):
            net = end_points["encoder_0"] = ops.convolution2d(reflect(images, (3, 3, 3, 3)), num_filters,
                                                              kernel_size=[7, 7], padding="VALID")
        with store.scope("encoder"), scopes.arg_scope([ops.convolution2d], kernel_size=[kh, kw], stride=2,
                                                      activation_fn=ops.relu, padding="VALID"):
            net = end_points["encoder_1"] = ops.convolution2d(reflect(net, pads), num_filters * 2)
            net = end_points["encoder_2"] = ops.convolution2d(reflect(net, pads), num_filters * 4)
        with store.scope("residual_blocks"), scopes.arg_scope([ops.convolution2d], kernel_size=[kh, kw], stride=1,
                                                              activation_fn=ops.relu, padding="VALID"):
            for b in range(num_resnet_blocks):
                with store.scope("block_%d" % b):
                    res = ops.convolution2d(reflect(net, pads), num_filters * 4)
                    res = ops.convolution2d(reflect(res, pads), num_filters * 4, activation_fn=None)
                    net = end_points["resnet_block_%d" % b] = net + res
        with store.scope("decoder"), scopes.arg_scope([ops.convolution2d], kernel_size=[kh, kw], stride=1,
                                                      activation_fn=ops.relu):
            with store.scope("decoder1"):
                net = end_points["decoder1"] = upsample_fn(net, num_outputs=num_filters * 2, stride=[2, 2])
            with store.scope("decoder2"):
                net = end_points["decoder2"] = upsample_fn(net, num_outputs=num_filters, stride=[2, 2])
        with store.scope("output"):
            logits = ops.convolution2d(reflect(net, (3, 3, 3, 3)), num_outputs, [7, 7], activation_fn=None,
                                       normalizer_fn=None, padding="VALID")
            logits = end_points["logits"] = logits.reshape(images.shape[0], h, w, num_outputs)
            end_points["predictions"] = Fn.tanh(logits) + logits * tanh_linear_slope
    return end_points["predictions"], end_points
