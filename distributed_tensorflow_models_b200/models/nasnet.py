"""NASNet-A (CIFAR / mobile / large) and PNASNet-5 large from the vendored zoo.

Reference behaviour: ``vgg/nets/nasnet/nasnet_utils.py`` (``calc_reduction_layers`` :45-52, ``factorized_reduction``
:88-135, ``drop_path`` :138-150, stacked separable convs :185-215, the cell machinery :253-460, the NASNet-A normal /
reduction cell genotypes :463-497), ``nasnet.py`` (configs :36-96, arg scopes :104-228, aux head :231-255, stems
:258-296, ``_build_nasnet_base`` :433-526) and ``pnasnet.py`` (config :33-46, base :55-123, PNASNet cell :180-197).

A searched cell is a *genotype*: ten operation names, ten hidden-state indices and a used-mask.  A cell call
 1. brings the previous cell's output to the current geometry (1x1 conv, or a factorized reduction when the spatial
    size differs), squeezes the current input through ReLU-1x1-BN, giving hidden states 0 and 1;
 2. runs five ``comb_iter_<k>`` steps, each adding the results of a ``left`` and a ``right`` operation applied to two
    chosen hidden states (``separable_KxK_n`` = n x [ReLU, depthwise KxK + pointwise, BN]; ``avg/max_pool_3x3``;
    ``none``), stride 2 only when the operand is one of the two cell inputs; every op except ``none`` is followed by
    scheduled drop-path;
 3. concatenates the hidden states nobody consumed (after a factorized reduction where their shape lags behind).

The depthwise stages run on ``csrc/zoo_ops.cu`` (3x3 / 5x5 / 7x7 register-window kernels), pointwise 1x1s on the
tcgen05 GEMM path.
"""
from __future__ import annotations

import contextlib
import copy
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch

from ..ops import functional as Fn
from ..slim import losses, ops, scopes
from ..slim import variables as V


@dataclass
class NasConfig:
    stem_multiplier: float
    dense_dropout_keep_prob: float
    num_cells: int
    filter_scaling_rate: float
    num_conv_filters: int
    drop_path_keep_prob: float
    use_aux_head: int = 1
    num_reduction_layers: int = 2
    skip_reduction_layer_input: int = 0
    total_training_steps: int = 250000


def cifar_config() -> NasConfig:
    return NasConfig(3.0, 1.0, 18, 2.0, 32, 0.6, total_training_steps=937500)


def mobile_imagenet_config() -> NasConfig:
    return NasConfig(1.0, 0.5, 12, 2.0, 44, 1.0)


def large_imagenet_config() -> NasConfig:
    return NasConfig(3.0, 0.5, 18, 2.0, 168, 0.7, skip_reduction_layer_input=1)


def pnasnet_large_config() -> NasConfig:
    return NasConfig(3.0, 0.5, 12, 2.0, 216, 0.6)


@dataclass(frozen=True)
class Genotype:
    operations: Sequence[str]
    used_hiddenstates: Sequence[int]
    hiddenstate_indices: Sequence[int]


NASNET_A_NORMAL = Genotype(
    ("separable_5x5_2", "separable_3x3_2", "separable_5x5_2", "separable_3x3_2", "avg_pool_3x3", "none",
     "avg_pool_3x3", "avg_pool_3x3", "separable_3x3_2", "none"),
    (1, 0, 0, 0, 0, 0, 0), (0, 1, 1, 1, 0, 1, 1, 1, 0, 0))
NASNET_A_REDUCTION = Genotype(
    ("separable_5x5_2", "separable_7x7_2", "max_pool_3x3", "separable_7x7_2", "avg_pool_3x3", "separable_5x5_2",
     "none", "avg_pool_3x3", "separable_3x3_2", "max_pool_3x3"),
    (1, 1, 1, 0, 0, 0, 0), (0, 1, 0, 1, 0, 1, 3, 2, 2, 0))
PNASNET_5 = Genotype(
    ("separable_5x5_2", "max_pool_3x3", "separable_7x7_2", "max_pool_3x3", "separable_5x5_2", "separable_3x3_2",
     "separable_3x3_2", "max_pool_3x3", "separable_3x3_2", "none"),
    (1, 1, 0, 0, 0, 0, 0), (1, 1, 0, 0, 0, 0, 4, 0, 1, 0))


def calc_reduction_layers(num_cells: int, num_reduction_layers: int) -> List[int]:
    return [int(float(k) / (num_reduction_layers + 1) * num_cells) for k in range(1, num_reduction_layers + 1)]


@contextlib.contextmanager
def _nasnet_arg_scope(weight_decay: float, batch_norm_decay: float, batch_norm_epsilon: float):
    """L2 + fan-out variance scaling on every weight; convolutions are bare (no bias, no activation: the cells
    apply ReLU *before* and BN *after* explicitly); BN with gamma."""
    reg = losses.l2_regularizer(weight_decay)
    init = V.contrib_variance_scaling_initializer(2.0, "fan_out")
    with scopes.arg_scope([ops.fully_connected, ops.convolution2d, ops.separable_convolution2d],
                          weights_regularizer=reg, weights_initializer=init):
        with scopes.arg_scope([ops.fully_connected], activation_fn=None, scope="FC"):
            with scopes.arg_scope([ops.convolution2d, ops.separable_convolution2d], activation_fn=None,
                                  use_bias=False):
                with scopes.arg_scope([ops.batch_norm], decay=batch_norm_decay, epsilon=batch_norm_epsilon,
                                      scale=True) as sc:
                    yield sc


def nasnet_cifar_arg_scope(weight_decay: float = 5e-4, batch_norm_decay: float = 0.9, batch_norm_epsilon: float = 1e-5):
    return _nasnet_arg_scope(weight_decay, batch_norm_decay, batch_norm_epsilon)


def nasnet_mobile_arg_scope(weight_decay: float = 4e-5, batch_norm_decay: float = 0.9997,
                            batch_norm_epsilon: float = 1e-3):
    return _nasnet_arg_scope(weight_decay, batch_norm_decay, batch_norm_epsilon)


def nasnet_large_arg_scope(weight_decay: float = 5e-5, batch_norm_decay: float = 0.9997,
                           batch_norm_epsilon: float = 1e-3):
    return _nasnet_arg_scope(weight_decay, batch_norm_decay, batch_norm_epsilon)


def pnasnet_large_arg_scope(weight_decay: float = 4e-5, batch_norm_decay: float = 0.9997,
                            batch_norm_epsilon: float = 0.001):
    return _nasnet_arg_scope(weight_decay, batch_norm_decay, batch_norm_epsilon)


def factorized_reduction(net, output_filters: int, stride: int):
    """Halve the resolution without discarding the odd pixels: two 1x1 convolutions read the even and the odd
    (shifted by one) sampling lattices and their outputs are concatenated."""
    assert output_filters % 2 == 0, "Need even number of filters when using this factorized reduction."
    if stride == 1:
        net = ops.convolution2d(net, output_filters, 1, scope="path_conv")
        return ops.batch_norm(net, scope="path_bn")
    path1 = ops.avg_pool2d(net, [1, 1], stride=stride, padding="VALID")
    path1 = ops.convolution2d(path1, output_filters // 2, 1, scope="path1_conv")
    shifted = Fn.pad_spatial(net, (0, 1, 0, 1))[:, 1:, 1:, :]
    path2 = ops.avg_pool2d(shifted, [1, 1], stride=stride, padding="VALID")
    path2 = ops.convolution2d(path2, output_filters // 2, 1, scope="path2_conv")
    return ops.batch_norm(torch.cat([path1, path2], dim=3), scope="final_path_bn")


def _parse_separable(op: str):
    parts = op.split("_")
    layers = 1 if "x" in parts[-1] else int(parts[-1])
    k = int(op.split("x")[0][-1])
    return layers, k


def _stacked_separable(net, stride: int, op: str, filters: int):
    layers, k = _parse_separable(op)
    for n in range(1, layers + 1):
        net = Fn.relu(net)
        net = ops.separable_convolution2d(net, filters, k, depth_multiplier=1, stride=stride, padding="SAME",
                                          scope=f"separable_{k}x{k}_{n}")
        net = ops.batch_norm(net, scope=f"bn_sep_{k}x{k}_{n}")
        stride = 1
    return net


class Cell:
    """One searched cell; ``__call__`` builds it under ``scope``."""

    def __init__(self, genotype: Genotype, num_conv_filters: int, drop_path_keep_prob: float, total_num_cells: int,
                 total_training_steps: int, is_training: bool, current_step: float = 0.0):
        self.g = genotype
        self.num_conv_filters = num_conv_filters
        self.drop_path_keep_prob = drop_path_keep_prob
        self.total_num_cells = total_num_cells
        self.total_training_steps = total_training_steps
        self.is_training = is_training
        self.current_step = current_step

    # -- pieces ----------------------------------------------------------------------------------
    def _reduce_prev(self, prev, cur, filters: int):
        if prev is None:
            return cur
        if cur.shape[2] != prev.shape[2]:
            return factorized_reduction(Fn.relu(prev), filters, stride=2)
        if filters != prev.shape[-1]:
            prev = ops.convolution2d(Fn.relu(prev), filters, 1, scope="prev_1x1")
            return ops.batch_norm(prev, scope="prev_bn")
        return prev

    def _drop_path(self, net, cell_num: int):
        keep = self.drop_path_keep_prob
        if keep >= 1.0 or not self.is_training:
            return net
        assert cell_num != -1
        keep = 1 - (cell_num + 1) / float(self.total_num_cells) * (1 - keep)            # deeper cells drop more
        ratio = min(1.0, float(self.current_step) / self.total_training_steps)          # linear ramp over training
        keep = 1 - ratio * (1 - keep)
        return Fn.drop_path(net, keep, True)

    def _apply(self, net, op: str, stride: int, from_input: bool, filters: int, cell_num: int):
        if stride > 1 and not from_input:
            stride = 1
        cin = net.shape[-1]
        if "separable" in op:
            net = _stacked_separable(net, stride, op, filters)
        elif op == "none":
            if stride > 1 or cin != filters:
                net = ops.convolution2d(Fn.relu(net), filters, 1, stride=stride, scope="1x1")
                net = ops.batch_norm(net, scope="bn_1")
        elif "pool" in op:
            kind, k = op.split("_")[0], int(op.split("_")[-1].split("x")[0])
            pool = ops.avg_pool2d if kind == "avg" else ops.max_pool2d
            if kind not in ("avg", "max"):
                raise NotImplementedError("Unimplemented pooling type: %s" % kind)
            net = pool(net, k, stride=stride, padding="SAME")
            if cin != filters:
                net = ops.convolution2d(net, filters, 1, stride=1, scope="1x1")
                net = ops.batch_norm(net, scope="bn_1")
        else:
            raise ValueError("Unimplemented operation %s" % op)
        return self._drop_path(net, cell_num) if op != "none" else net

    def _combine_unused(self, states):
        used = self.g.used_hiddenstates
        assert len(used) == len(states)
        final_h, final_c = states[-1].shape[2], states[-1].shape[-1]
        store = V.current_store()
        for idx, u in enumerate(used):
            h, c = states[idx].shape[2], states[idx].shape[-1]
            if (final_h != h or final_c != c) and not u:
                with store.scope(f"reduction_{idx}"):
                    states[idx] = factorized_reduction(states[idx], final_c, 2 if final_h != h else 1)
        return torch.cat([s for s, u in zip(states, used) if not u], dim=3)

    # -- the cell --------------------------------------------------------------------------------
    def __call__(self, net, scope: str, filter_scaling: float = 1, stride: int = 1, prev_layer=None,
                 cell_num: int = -1):
        filters = int(self.num_conv_filters * filter_scaling)
        store = V.current_store()
        with store.scope(scope):
            prev = self._reduce_prev(prev_layer, net, filters)
            net = ops.convolution2d(Fn.relu(net), filters, 1, scope="1x1")
            net = ops.batch_norm(net, scope="beginning_bn")
            states = [net, prev]
            for it in range(5):
                li, ri = self.g.hiddenstate_indices[2 * it], self.g.hiddenstate_indices[2 * it + 1]
                with store.scope(f"comb_iter_{it}"):
                    with store.scope("left"):
                        h1 = self._apply(states[li], self.g.operations[2 * it], stride, li < 2, filters, cell_num)
                    with store.scope("right"):
                        h2 = self._apply(states[ri], self.g.operations[2 * it + 1], stride, ri < 2, filters, cell_num)
                    states.append(h1 + h2)
            with store.scope("cell_output"):
                return self._combine_unused(states)


def _aux_head(net, end_points: dict, num_classes: int, scope: str):
    store = V.current_store()
    with store.scope(scope + "/aux_logits"):
        aux = ops.avg_pool2d(net, [5, 5], stride=3, padding="VALID")
        aux = Fn.relu(ops.batch_norm(ops.convolution2d(aux, 128, [1, 1], scope="proj"), scope="aux_bn0"))
        aux = ops.convolution2d(aux, 768, [aux.shape[1], aux.shape[2]], padding="VALID")
        aux = Fn.relu(ops.batch_norm(aux, scope="aux_bn1"))
        end_points["AuxLogits"] = ops.fully_connected(ops.flatten(aux), num_classes)


def _imagenet_stem(inputs, cfg: NasConfig, stem_cell: Cell):
    net = ops.convolution2d(inputs, int(32 * cfg.stem_multiplier), [3, 3], stride=2, padding="VALID", scope="conv0")
    net = ops.batch_norm(net, scope="conv0_bn")
    outputs = [None, net]
    scaling = 1.0 / (cfg.filter_scaling_rate ** 2)
    for n in range(2):
        net = stem_cell(net, scope=f"cell_stem_{n}", filter_scaling=scaling, stride=2, prev_layer=outputs[-2],
                        cell_num=n)
        outputs.append(net)
        scaling *= cfg.filter_scaling_rate
    return net, outputs


def _cifar_stem(inputs, cfg: NasConfig):
    net = ops.convolution2d(inputs, int(cfg.num_conv_filters * cfg.stem_multiplier), 3, scope="l1_stem_3x3")
    net = ops.batch_norm(net, scope="l1_stem_bn")
    return net, [None, net]


def _head(net, end_points: dict, num_classes, cfg: NasConfig, final_endpoint):
    store = V.current_store()
    with store.scope("final_layer"):
        net = end_points["global_pool"] = Fn.relu(net).mean((1, 2))
        if final_endpoint == "global_pool" or not num_classes:
            return net, end_points
        net = ops.dropout(net, cfg.dense_dropout_keep_prob, scope="dropout")
        logits = end_points["Logits"] = ops.fully_connected(net, num_classes)
        end_points["Predictions"] = ops.softmax(logits)
    return logits, end_points


def _build(images, num_classes, is_training: bool, cfg: NasConfig, stem_type: str, normal: Genotype,
           reduction: Optional[Genotype], final_endpoint: Optional[str], current_step: float):
    """``reduction is None`` = PNASNet: one cell type, run with stride 2 at the reduction positions."""
    cfg = copy.deepcopy(cfg)
    if not is_training:
        cfg.drop_path_keep_prob = 1.0
    pnas = reduction is None
    total = cfg.num_cells + 2 + (2 if (stem_type == "imagenet" and not pnas) else 0)
    mk = lambda g: Cell(g, cfg.num_conv_filters, cfg.drop_path_keep_prob, total, cfg.total_training_steps, is_training,
                        current_step)
    normal_cell = mk(normal)
    reduction_cell = None if pnas else mk(reduction)
    end_points: Dict[str, torch.Tensor] = {}
    red_idx = calc_reduction_layers(cfg.num_cells, cfg.num_reduction_layers)
    with scopes.arg_scope([ops.dropout, ops.batch_norm], is_training=is_training):
        if stem_type == "imagenet":
            net, outputs = _imagenet_stem(images, cfg, normal_cell if pnas else reduction_cell)
        elif stem_type == "cifar":
            net, outputs = _cifar_stem(images, cfg)
        else:
            raise ValueError("Unknown stem_type: %s" % stem_type)
        end_points["Stem"] = net
        if final_endpoint == "Stem":
            return net, end_points
        aux_cells = [red_idx[1] - 1] if len(red_idx) >= 2 else []
        scaling = 1.0
        true_num = 2 if stem_type == "imagenet" else 0
        for n in range(cfg.num_cells):
            stride = 1
            prev = outputs[-2]
            if n in red_idx:
                scaling *= cfg.filter_scaling_rate
                if pnas:
                    stride = 2
                else:
                    k = red_idx.index(n)
                    net = reduction_cell(net, scope=f"reduction_cell_{k}", filter_scaling=scaling, stride=2,
                                         prev_layer=outputs[-2], cell_num=true_num)
                    end_points[f"Reduction_Cell_{k}"] = net
                    if final_endpoint == f"Reduction_Cell_{k}":
                        return net, end_points
                    true_num += 1
                    outputs.append(net)
                    if not cfg.skip_reduction_layer_input:
                        prev = outputs[-2]
            net = normal_cell(net, scope=f"cell_{n}", filter_scaling=scaling, stride=stride, prev_layer=prev,
                              cell_num=true_num)
            end_points[f"Cell_{n}"] = net
            if final_endpoint == f"Cell_{n}":
                return net, end_points
            true_num += 1
            if cfg.use_aux_head and n in aux_cells and num_classes and is_training:
                _aux_head(Fn.relu(net), end_points, num_classes, scope=f"aux_{n}")
            outputs.append(net)
        out, end_points = _head(net, end_points, num_classes, cfg, final_endpoint)
    V.current_store().end_points = end_points
    return out, end_points


def build_nasnet_cifar(images, num_classes, is_training: bool = True, config: Optional[NasConfig] = None,
                       current_step: float = 0.0):
    return _build(images, num_classes, is_training, config or cifar_config(), "cifar", NASNET_A_NORMAL,
                  NASNET_A_REDUCTION, None, current_step)


def build_nasnet_mobile(images, num_classes, is_training: bool = True, final_endpoint: Optional[str] = None,
                        config: Optional[NasConfig] = None, current_step: float = 0.0):
    return _build(images, num_classes, is_training, config or mobile_imagenet_config(), "imagenet", NASNET_A_NORMAL,
                  NASNET_A_REDUCTION, final_endpoint, current_step)


def build_nasnet_large(images, num_classes, is_training: bool = True, final_endpoint: Optional[str] = None,
                       config: Optional[NasConfig] = None, current_step: float = 0.0):
    return _build(images, num_classes, is_training, config or large_imagenet_config(), "imagenet", NASNET_A_NORMAL,
                  NASNET_A_REDUCTION, final_endpoint, current_step)


def build_pnasnet_large(images, num_classes, is_training: bool = True, final_endpoint: Optional[str] = None,
                        config: Optional[NasConfig] = None, current_step: float = 0.0):
    return _build(images, num_classes, is_training, config or pnasnet_large_config(), "imagenet", PNASNET_5, None,
                  final_endpoint, current_step)


build_nasnet_cifar.default_image_size = 32
build_nasnet_mobile.default_image_size = 224
build_nasnet_large.default_image_size = 331
build_pnasnet_large.default_image_size = 331
