"""Model registry — the ``nets_factory`` of the framework (reference ``vgg/nets/nets_factory.py:39-145``:
name -> constructor, matching arg_scope, ``default_image_size``), covering the six networks the reference
trains plus the ImageNet ResNets and VGG variants it ships.

``get_network_fn(name, num_classes, weight_decay, is_training)`` returns ``fn(images) -> (logits,
end_points)`` exactly like the reference factory (including its ``is_training=False`` default that makes
VGG train without dropout).  ``MODEL_SPECS`` additionally records what the trainers need:
input geometry, class count and default batch size per worker.
"""
from __future__ import annotations

import contextlib
import functools
from dataclasses import dataclass
from typing import Callable, Dict, Optional

from . import (alexnet, cifarnet, inception_v3, inception_zoo, mobilenet, nasnet, plain_cnn, resnet, resnet_slim, vgg,
               zoo_small)


@dataclass(frozen=True)
class ModelSpec:
    name: str
    image_size: int          # network input H = W
    source_size: int         # dataset image size before cropping
    channels: int
    num_classes: int
    batch_size: int          # reference default per-worker batch
    dataset: str             # 'cifar10' | 'imagenet'


MODEL_SPECS: Dict[str, ModelSpec] = {
    "cnn": ModelSpec("cnn", 24, 32, 3, 10, 512, "cifar10"),
    "cifarnet": ModelSpec("cifarnet", 32, 32, 3, 10, 512, "cifar10"),
    "alexnet": ModelSpec("alexnet", 32, 32, 3, 10, 128, "cifar10"),
    "vgg": ModelSpec("vgg", 32, 32, 3, 10, 72, "cifar10"),
    "resnet": ModelSpec("resnet", 32, 32, 3, 10, 128, "cifar10"),
    "inception": ModelSpec("inception", 299, 299, 3, 1001, 32, "imagenet"),
}


def is_zoo_model(name: str) -> bool:
    """A registry network that is not one of the six trained families (those keep their own trainer presets)."""
    return name not in MODEL_SPECS and name in networks_map


def spec_for(name: str, image_size: Optional[int] = None, num_classes: Optional[int] = None) -> ModelSpec:
    """``MODEL_SPECS[name]`` for the six trained models; for any other registry network an ImageNet-shaped spec at
    the network's ``default_image_size`` (``vgg/nets/mobilenet_v1_train.py:36-40``: 1001 classes, batch 64)."""
    if name in MODEL_SPECS:
        return MODEL_SPECS[name]
    if name not in networks_map:
        raise ValueError("Name of network unknown %s" % name)
    size = image_size or getattr(networks_map[name], "default_image_size", default_image_sizes.get(name, 224))
    small = size <= 32
    return ModelSpec(name, size, size, 3, num_classes or (10 if small else 1001), 64,
                     "cifar10" if small else "imagenet")


@contextlib.contextmanager
def _null_scope(weight_decay: float = 0.0):
    yield None


def _resnet_cifar(images, num_classes=10, is_training=False, resnet_size=32, **_):
    net = resnet.cifar10_resnet_v2_generator(resnet_size, num_classes)
    return net(images, is_training), {}


def _resnet_imagenet(size):
    def fn(images, num_classes=1001, is_training=False, **_):
        return resnet.resnet_v2(size, num_classes)(images, is_training), {}
    fn.default_image_size = 224
    return fn


def _plain_cnn(images, num_classes=10, is_training=False, **_):
    return plain_cnn.inference(images), {}


def _cifarnet(images, num_classes=10, is_training=False, **kw):
    out = cifarnet.cifarnet(images, num_classes, is_training=is_training, **kw)
    from ..slim import variables as V
    return out, V.current_store().end_points


def _inception(images, num_classes=1001, is_training=False, **_):
    logits, aux = inception_v3.inference(images, num_classes, for_training=is_training)
    return (logits, aux), {"aux_logits": aux}


networks_map: Dict[str, Callable] = {
    "cnn": _plain_cnn,
    "cifarnet": _cifarnet,
    "alexnet_v2": alexnet.alexnet_v2,
    "vgg_a": vgg.vgg_a,
    "vgg_16": vgg.vgg_16,
    "vgg_19": vgg.vgg_19,
    "resnet_cifar": _resnet_cifar,
    "resnet_v2_18": _resnet_imagenet(18),
    "resnet_v2_34": _resnet_imagenet(34),
    "resnet_v2_50": _resnet_imagenet(50),
    "resnet_v2_101": _resnet_imagenet(101),
    "resnet_v2_152": _resnet_imagenet(152),
    "resnet_v2_200": _resnet_imagenet(200),
    "inception_v3": _inception,
    "lenet": zoo_small.lenet,
    "overfeat": zoo_small.overfeat,
    # the rest of the vendored zoo (vgg/nets/nets_factory.py:39-68); the slim bottleneck ResNet-v2s are registered
    # as resnet_v2_*_slim because the plain names above are the reference's own resnet_model.py ImageNet nets
    "inception_v1": inception_zoo.inception_v1,
    "inception_v2": inception_zoo.inception_v2,
    "inception_v3_slim": inception_zoo.inception_v3,
    "inception_v4": inception_zoo.inception_v4,
    "inception_resnet_v2": inception_zoo.inception_resnet_v2,
    "resnet_v1_50": resnet_slim.resnet_v1_50,
    "resnet_v1_101": resnet_slim.resnet_v1_101,
    "resnet_v1_152": resnet_slim.resnet_v1_152,
    "resnet_v1_200": resnet_slim.resnet_v1_200,
    "resnet_v2_50_slim": resnet_slim.resnet_v2_50,
    "resnet_v2_101_slim": resnet_slim.resnet_v2_101,
    "resnet_v2_152_slim": resnet_slim.resnet_v2_152,
    "resnet_v2_200_slim": resnet_slim.resnet_v2_200,
    "mobilenet_v1": mobilenet.mobilenet_v1,
    "mobilenet_v1_075": mobilenet.mobilenet_v1_075,
    "mobilenet_v1_050": mobilenet.mobilenet_v1_050,
    "mobilenet_v1_025": mobilenet.mobilenet_v1_025,
    "mobilenet_v2": mobilenet.mobilenet_v2,
    "nasnet_cifar": nasnet.build_nasnet_cifar,
    "nasnet_mobile": nasnet.build_nasnet_mobile,
    "nasnet_large": nasnet.build_nasnet_large,
    "pnasnet_large": nasnet.build_pnasnet_large,
}

arg_scopes_map: Dict[str, Callable] = {
    "vgg_a": vgg.vgg_arg_scope, "vgg_16": vgg.vgg_arg_scope, "vgg_19": vgg.vgg_arg_scope,
    "lenet": zoo_small.lenet_arg_scope, "overfeat": zoo_small.overfeat_arg_scope,
    "inception_v1": inception_zoo.inception_v3_arg_scope, "inception_v2": inception_zoo.inception_v3_arg_scope,
    "inception_v3_slim": inception_zoo.inception_v3_arg_scope, "inception_v4": inception_zoo.inception_v4_arg_scope,
    "inception_resnet_v2": inception_zoo.inception_resnet_v2_arg_scope,
    **{f"resnet_v1_{n}": resnet_slim.resnet_arg_scope for n in (50, 101, 152, 200)},
    **{f"resnet_v2_{n}_slim": resnet_slim.resnet_arg_scope for n in (50, 101, 152, 200)},
    **{k: mobilenet.mobilenet_v1_arg_scope for k in ("mobilenet_v1", "mobilenet_v1_075", "mobilenet_v1_050",
                                                      "mobilenet_v1_025")},
    "mobilenet_v2": mobilenet.mobilenet_v2_training_scope,
    "nasnet_cifar": nasnet.nasnet_cifar_arg_scope, "nasnet_mobile": nasnet.nasnet_mobile_arg_scope,
    "nasnet_large": nasnet.nasnet_large_arg_scope, "pnasnet_large": nasnet.pnasnet_large_arg_scope,
}

default_image_sizes = {"cnn": 24, "cifarnet": 32, "alexnet_v2": 224, "vgg_a": 224, "vgg_16": 224, "vgg_19": 224,
                       "resnet_cifar": 32, "inception_v3": 299, "lenet": 28, "overfeat": 231}


def get_network_fn(name: str, num_classes: int, weight_decay: Optional[float] = 0.0,
                   is_training: bool = False) -> Callable:
    """-> ``network_fn(images, **kwargs) -> (logits, end_points)`` with ``default_image_size``."""
    if name not in networks_map:
        raise ValueError("Name of network unknown %s" % name)
    func = networks_map[name]
    scope_fn = arg_scopes_map.get(name, _null_scope)

    @functools.wraps(func)
    def network_fn(images, **kwargs):
        with (scope_fn() if weight_decay is None else scope_fn(weight_decay=weight_decay)):
            return func(images, num_classes=num_classes, is_training=is_training, **kwargs)

    network_fn.default_image_size = getattr(func, "default_image_size", default_image_sizes.get(name, 224))
    return network_fn
