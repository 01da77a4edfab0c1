"""The zoo's Inception family: ``inception_v1``, ``inception_v2``, ``inception_v3`` (the ``nets``
edition, as locally edited by the reference for small inputs), ``inception_v4`` and ``inception_resnet_v2``.

Every network is written as *data*: a list of ``(end_point, spec)`` rows where a spec is either one layer or a
``Mixed`` block — a tuple of branches, each a list of layers — and one interpreter walks the tables.  The layer
constructors below are the whole grammar:

    C(depth, k, name, stride=1, padding=None, init=None)    convolution (+BN +ReLU from the arg scope)
    MP(k, name, stride=1, padding=None) / AP(...)           max / average pool
    FORK(layer_a, layer_b)                                  both on the same input, channel-concatenated

Reference counterparts: ``vgg/nets/inception_v1.py:28-256`` + head ``:259-330``; ``inception_v2.py:29-449`` +
``:452-560``; ``inception_v3.py:29-423`` + ``:437-560`` (with the reference's own SAME-padding edits ``:100-102,
218-229,356-369,506-538`` and its irregular scope names kept: ``Conv_1_0c_5x5``, ``Conv2d_1a_1x1``...);
``inception_v4.py:34-258`` + ``:261-334``; ``inception_resnet_v2.py:33-293`` + ``:296-371``; arg scopes
``inception_utils.py:30-75`` and ``inception_resnet_v2.py:374-397``.  (The Inception-v3 the reference *trains* is the
older in-tree slim model in ``models/inception_v3.py``.)
"""
from __future__ import annotations

import contextlib
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from ..ops import streams
from ..slim import losses, ops, scopes
from ..slim import variables as V


# ---- grammar -------------------------------------------------------------------------------------------
def C(depth, k, name, stride=1, padding=None, init=None):
    k = (k, k) if isinstance(k, int) else tuple(k)
    return ("conv", depth, k, name, stride, padding, init)


def MP(k, name, stride=1, padding=None):
    return ("max", None, (k, k) if isinstance(k, int) else tuple(k), name, stride, padding, None)


def AP(k, name, stride=1, padding=None):
    return ("avg", None, (k, k) if isinstance(k, int) else tuple(k), name, stride, padding, None)


def FORK(*layers):
    return ("fork", layers)


def _run_layer(net, spec, depth: Callable[[int], int], padding: str):
    if spec[0] == "fork":
        return torch.cat([_run_layer(net, s, depth, padding) for s in spec[1]], dim=3)
    kind, d, k, name, stride, pad, init = spec
    pad = pad or padding
    if kind == "conv":
        kw = {"weights_initializer": V.truncated_normal_initializer(init)} if init is not None else {}
        return ops.convolution2d(net, depth(d), list(k), stride=stride, padding=pad, scope=name, **kw)
    if kind == "max":
        return ops.max_pool2d(net, list(k), stride=stride, padding=pad, scope=name)
    return ops.avg_pool2d(net, list(k), stride=stride, padding=pad, scope=name)


def _run_mixed(net, branches, depth, padding: str, scope: str):
    store = V.current_store()

    def _branch(i, branch):
        def run():
            with store.scope(f"Branch_{i}"):
                b = net
                for spec in branch:
                    b = _run_layer(b, spec, depth, padding)
                return b
        return run

    with store.scope(scope):
        # independent branches: side streams on CUDA (ops/streams.py), plain sequence otherwise
        outs = streams.run_branches([_branch(i, br) for i, br in enumerate(branches)], net)
    return torch.cat(outs, dim=3)


def _run_table(net, table, depth, padding: str, final_endpoint: Optional[str], end_points: dict):
    for name, spec in table:
        if isinstance(spec, tuple) and spec and isinstance(spec[0], list):
            net = _run_mixed(net, spec, depth, padding, name)
        else:
            net = _run_layer(net, spec, depth, padding)
        end_points[name] = net
        if name == final_endpoint:
            return net, True
    return net, False


@contextlib.contextmanager
def inception_arg_scope(weight_decay: float = 0.00004, use_batch_norm: bool = True, batch_norm_decay: float = 0.9997,
                        batch_norm_epsilon: float = 0.001, activation_fn=ops.relu):
    """L2 on conv / fc weights; conv = variance-scaling init + BN(decay, eps; centre only) + ReLU."""
    reg = losses.l2_regularizer(weight_decay)
    bn = {"decay": batch_norm_decay, "epsilon": batch_norm_epsilon}
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected], weights_regularizer=reg):
        with scopes.arg_scope([ops.convolution2d], weights_initializer=V.variance_scaling_initializer(),
                              activation_fn=activation_fn, normalizer_fn=ops.batch_norm if use_batch_norm else None,
                              normalizer_params=bn if use_batch_norm else None) as sc:
            yield sc


inception_v3_arg_scope = inception_arg_scope
inception_v4_arg_scope = inception_arg_scope


@contextlib.contextmanager
def inception_resnet_v2_arg_scope(weight_decay: float = 0.00004, batch_norm_decay: float = 0.9997,
                                  batch_norm_epsilon: float = 0.001, activation_fn=ops.relu):
    reg = losses.l2_regularizer(weight_decay)
    with scopes.arg_scope([ops.convolution2d, ops.fully_connected], weights_regularizer=reg):
        with scopes.arg_scope([ops.convolution2d], activation_fn=activation_fn, normalizer_fn=ops.batch_norm,
                              normalizer_params={"decay": batch_norm_decay, "epsilon": batch_norm_epsilon}) as sc:
            yield sc


def _squeeze(logits, spatial_squeeze: bool):
    if spatial_squeeze and logits.dim() == 4 and logits.shape[1] == 1 and logits.shape[2] == 1:
        return logits.reshape(logits.shape[0], logits.shape[-1])
    return logits


def _small(net, k):
    """Pooling / "fc" kernels shrink to the feature map for small inputs."""
    return [min(net.shape[1], k[0]), min(net.shape[2], k[1])]


# ---- Inception v1 (GoogLeNet) ----------------------------------------------------------------------------
def _v1_mixed(b0, b1r, b1, b2r, b2, b3, quirk=False):
    return ([C(b0, 1, "Conv2d_0a_1x1")],
            [C(b1r, 1, "Conv2d_0a_1x1"), C(b1, 3, "Conv2d_0b_3x3")],
            [C(b2r, 1, "Conv2d_0a_1x1"), C(b2, 3, "Conv2d_0a_3x3" if quirk else "Conv2d_0b_3x3")],
            [MP(3, "MaxPool_0a_3x3"), C(b3, 1, "Conv2d_0b_1x1")])


_V1 = [
    ("Conv2d_1a_7x7", C(64, 7, "Conv2d_1a_7x7", 2)), ("MaxPool_2a_3x3", MP(3, "MaxPool_2a_3x3", 2)),
    ("Conv2d_2b_1x1", C(64, 1, "Conv2d_2b_1x1")), ("Conv2d_2c_3x3", C(192, 3, "Conv2d_2c_3x3")),
    ("MaxPool_3a_3x3", MP(3, "MaxPool_3a_3x3", 2)),
    ("Mixed_3b", _v1_mixed(64, 96, 128, 16, 32, 32)), ("Mixed_3c", _v1_mixed(128, 128, 192, 32, 96, 64)),
    ("MaxPool_4a_3x3", MP(3, "MaxPool_4a_3x3", 2)),
    ("Mixed_4b", _v1_mixed(192, 96, 208, 16, 48, 64)), ("Mixed_4c", _v1_mixed(160, 112, 224, 24, 64, 64)),
    ("Mixed_4d", _v1_mixed(128, 128, 256, 24, 64, 64)), ("Mixed_4e", _v1_mixed(112, 144, 288, 32, 64, 64)),
    ("Mixed_4f", _v1_mixed(256, 160, 320, 32, 128, 128)),
    ("MaxPool_5a_2x2", MP(2, "MaxPool_5a_2x2", 2)),
    ("Mixed_5b", _v1_mixed(256, 160, 320, 32, 128, 128, quirk=True)),     # reference names it Conv2d_0a_3x3
    ("Mixed_5c", _v1_mixed(384, 192, 384, 48, 128, 128)),
]


def inception_v1_base(inputs, final_endpoint: str = "Mixed_5c", scope: str = "InceptionV1"):
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with store.scope(scope):
        with scopes.arg_scope([ops.convolution2d, ops.fully_connected],
                              weights_initializer=V.truncated_normal_initializer(0.01)):
            net, hit = _run_table(inputs, _V1, lambda d: d, "SAME", final_endpoint, end_points)
    if not hit:
        raise ValueError("Unknown final endpoint %s" % final_endpoint)
    return net, end_points


def inception_v1(inputs, num_classes: int = 1000, is_training: bool = True, dropout_keep_prob: float = 0.8,
                 spatial_squeeze: bool = True, global_pool: bool = False, scope: str = "InceptionV1"):
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        net, end_points = inception_v1_base(inputs, scope=scope)
        with store.scope(scope + "/Logits"):
            if global_pool:
                net = end_points["global_pool"] = net.mean((1, 2), keepdim=True)
            else:
                net = end_points["AvgPool_0a_7x7"] = ops.avg_pool2d(net, _small(net, [7, 7]), stride=1,
                                                                    scope="AvgPool_0a_7x7")
            if not num_classes:
                return net, end_points
            net = ops.dropout(net, dropout_keep_prob, scope="Dropout_0b")
            logits = ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                       scope="Conv2d_0c_1x1")
            logits = _squeeze(logits, spatial_squeeze)
        end_points["Logits"] = logits
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


inception_v1.default_image_size = 224


# ---- Inception v2 (BN-Inception) -------------------------------------------------------------------------
def _v2_mixed(b0, b1r, b1, b2r, b2, b3, pool="avg"):
    P = AP if pool == "avg" else MP
    return ([C(b0, 1, "Conv2d_0a_1x1")],
            [C(b1r, 1, "Conv2d_0a_1x1", init=0.09), C(b1, 3, "Conv2d_0b_3x3")],
            [C(b2r, 1, "Conv2d_0a_1x1", init=0.09), C(b2, 3, "Conv2d_0b_3x3"), C(b2, 3, "Conv2d_0c_3x3")],
            [P(3, "AvgPool_0a_3x3" if pool == "avg" else "MaxPool_0a_3x3"), C(b3, 1, "Conv2d_0b_1x1", init=0.1)])


def _v2_reduce(b0r, b0, b1r, b1):
    return ([C(b0r, 1, "Conv2d_0a_1x1", init=0.09), C(b0, 3, "Conv2d_1a_3x3", 2)],
            [C(b1r, 1, "Conv2d_0a_1x1", init=0.09), C(b1, 3, "Conv2d_0b_3x3"), C(b1, 3, "Conv2d_1a_3x3", 2)],
            [MP(3, "MaxPool_1a_3x3", 2)])


_V2_TAIL = [
    ("MaxPool_2a_3x3", MP(3, "MaxPool_2a_3x3", 2)),
    ("Conv2d_2b_1x1", C(64, 1, "Conv2d_2b_1x1", init=0.1)), ("Conv2d_2c_3x3", C(192, 3, "Conv2d_2c_3x3")),
    ("MaxPool_3a_3x3", MP(3, "MaxPool_3a_3x3", 2)),
    ("Mixed_3b", _v2_mixed(64, 64, 64, 64, 96, 32)), ("Mixed_3c", _v2_mixed(64, 64, 96, 64, 96, 64)),
    ("Mixed_4a", _v2_reduce(128, 160, 64, 96)),
    ("Mixed_4b", _v2_mixed(224, 64, 96, 96, 128, 128)), ("Mixed_4c", _v2_mixed(192, 96, 128, 96, 128, 128)),
    ("Mixed_4d", _v2_mixed(160, 128, 160, 128, 160, 96)), ("Mixed_4e", _v2_mixed(96, 128, 192, 160, 192, 96)),
    ("Mixed_5a", _v2_reduce(128, 192, 192, 256)),
    ("Mixed_5b", _v2_mixed(352, 192, 320, 160, 224, 128)),
    ("Mixed_5c", _v2_mixed(352, 192, 320, 192, 224, 128, pool="max")),
]


def inception_v2_base(inputs, final_endpoint: str = "Mixed_5c", min_depth: int = 16, depth_multiplier: float = 1.0,
                      use_separable_conv: bool = True, scope: str = "InceptionV2"):
    if depth_multiplier <= 0:
        raise ValueError("depth_multiplier is not greater than zero.")
    depth = lambda d: max(int(d * depth_multiplier), min_depth)
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with store.scope(scope):
        # the 7x7 stem is a separable convolution: 3 input planes x min(depth(64)/3, 8) filters each, then 1x1
        name = "Conv2d_1a_7x7"
        if use_separable_conv:
            net = ops.separable_convolution2d(inputs, depth(64), [7, 7], depth_multiplier=min(int(depth(64) / 3), 8),
                                              stride=2, padding="SAME",
                                              weights_initializer=V.truncated_normal_initializer(1.0), scope=name)
        else:
            net = ops.convolution2d(inputs, depth(64), [7, 7], stride=2, padding="SAME",
                                    weights_initializer=V.truncated_normal_initializer(1.0), scope=name)
        end_points[name] = net
        if final_endpoint == name:
            return net, end_points
        net, hit = _run_table(net, _V2_TAIL, depth, "SAME", final_endpoint, end_points)
    if not hit:
        raise ValueError("Unknown final endpoint %s" % final_endpoint)
    return net, end_points


def inception_v2(inputs, num_classes: int = 1000, is_training: bool = True, dropout_keep_prob: float = 0.8,
                 min_depth: int = 16, depth_multiplier: float = 1.0, spatial_squeeze: bool = True,
                 global_pool: bool = False, scope: str = "InceptionV2"):
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        # (the shared arg scope covers conv2d only, so the separable stem keeps slim's defaults: bias + ReLU, no BN)
        net, end_points = inception_v2_base(inputs, min_depth=min_depth, depth_multiplier=depth_multiplier,
                                            scope=scope)
        with store.scope(scope + "/Logits"):
            if global_pool:
                net = end_points["global_pool"] = net.mean((1, 2), keepdim=True)
            else:
                k = _small(net, [7, 7])
                net = end_points["AvgPool_1a"] = ops.avg_pool2d(net, k, stride=1, padding="VALID",
                                                                scope="AvgPool_1a_{}x{}".format(*k))
            if not num_classes:
                return net, end_points
            net = ops.dropout(net, dropout_keep_prob, scope="Dropout_1b")
            logits = _squeeze(ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                                scope="Conv2d_1c_1x1"), spatial_squeeze)
        end_points["Logits"] = logits
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


inception_v2.default_image_size = 224


# ---- Inception v3 (nets edition, with the reference's SAME-padding edits) --------------------------------
def _v3_35(pool, names=("Conv2d_0a_1x1", "Conv2d_0b_5x5")):
    return ([C(64, 1, "Conv2d_0a_1x1")],
            [C(48, 1, names[0]), C(64, 5, names[1])],
            [C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_0b_3x3"), C(96, 3, "Conv2d_0c_3x3")],
            [AP(3, "AvgPool_0a_3x3"), C(pool, 1, "Conv2d_0b_1x1")])


def _v3_17(mid):
    return ([C(192, 1, "Conv2d_0a_1x1")],
            [C(mid, 1, "Conv2d_0a_1x1"), C(mid, (1, 7), "Conv2d_0b_1x7"), C(192, (7, 1), "Conv2d_0c_7x1")],
            [C(mid, 1, "Conv2d_0a_1x1"), C(mid, (7, 1), "Conv2d_0b_7x1"), C(mid, (1, 7), "Conv2d_0c_1x7"),
             C(mid, (7, 1), "Conv2d_0d_7x1"), C(192, (1, 7), "Conv2d_0e_1x7")],
            [AP(3, "AvgPool_0a_3x3"), C(192, 1, "Conv2d_0b_1x1")])


def _v3_8(second_3x1):
    return ([C(320, 1, "Conv2d_0a_1x1")],
            [C(384, 1, "Conv2d_0a_1x1"), FORK(C(384, (1, 3), "Conv2d_0b_1x3"), C(384, (3, 1), second_3x1))],
            [C(448, 1, "Conv2d_0a_1x1"), C(384, 3, "Conv2d_0b_3x3"),
             FORK(C(384, (1, 3), "Conv2d_0c_1x3"), C(384, (3, 1), "Conv2d_0d_3x1"))],
            [AP(3, "AvgPool_0a_3x3"), C(192, 1, "Conv2d_0b_1x1")])


_V3 = [
    ("Conv2d_1a_3x3", C(32, 3, "Conv2d_1a_3x3", 2)), ("Conv2d_2a_3x3", C(32, 3, "Conv2d_2a_3x3")),
    ("Conv2d_2b_3x3", C(64, 3, "Conv2d_2b_3x3", padding="SAME")), ("MaxPool_3a_3x3", MP(3, "MaxPool_3a_3x3", 2)),
    ("Conv2d_3b_1x1", C(80, 1, "Conv2d_3b_1x1")), ("Conv2d_4a_3x3", C(192, 3, "Conv2d_4a_3x3")),
    ("MaxPool_5a_3x3", MP(3, "MaxPool_5a_3x3", 2)),
    ("Mixed_5b", _v3_35(32)), ("Mixed_5c", _v3_35(64, ("Conv2d_0b_1x1", "Conv_1_0c_5x5"))), ("Mixed_5d", _v3_35(64)),
    ("Mixed_6a", ([C(384, 3, "Conv2d_1a_1x1", 2, "SAME")],
                  [C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_0b_3x3"), C(96, 3, "Conv2d_1a_1x1", 2, "SAME")],
                  [MP(3, "MaxPool_1a_3x3", 2, "SAME")])),
    ("Mixed_6b", _v3_17(128)), ("Mixed_6c", _v3_17(160)), ("Mixed_6d", _v3_17(160)), ("Mixed_6e", _v3_17(192)),
    ("Mixed_7a", ([C(192, 1, "Conv2d_0a_1x1"), C(320, 3, "Conv2d_1a_3x3", 2, "SAME")],
                  [C(192, 1, "Conv2d_0a_1x1"), C(192, (1, 7), "Conv2d_0b_1x7"), C(192, (7, 1), "Conv2d_0c_7x1"),
                   C(192, 3, "Conv2d_1a_3x3", 2, "SAME")],
                  [MP(3, "MaxPool_1a_3x3", 2, "SAME")])),
    ("Mixed_7b", _v3_8("Conv2d_0b_3x1")), ("Mixed_7c", _v3_8("Conv2d_0c_3x1")),
]


def inception_v3_base(inputs, final_endpoint: str = "Mixed_7c", min_depth: int = 16, depth_multiplier: float = 1.0,
                      scope: str = "InceptionV3"):
    if depth_multiplier <= 0:
        raise ValueError("depth_multiplier is not greater than zero.")
    depth = lambda d: max(int(d * depth_multiplier), min_depth)
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with store.scope(scope):
        net, hit = _run_table(inputs, _V3, depth, "SAME", final_endpoint, end_points)
    if not hit:
        raise ValueError("Unknown final endpoint %s" % final_endpoint)
    return net, end_points


def inception_v3(inputs, num_classes: int = 1000, is_training: bool = True, dropout_keep_prob: float = 0.8,
                 min_depth: int = 16, depth_multiplier: float = 1.0, spatial_squeeze: bool = True,
                 create_aux_logits: bool = True, global_pool: bool = False, scope: str = "InceptionV3"):
    depth = lambda d: max(int(d * depth_multiplier), min_depth)
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        net, end_points = inception_v3_base(inputs, min_depth=min_depth, depth_multiplier=depth_multiplier,
                                            scope=scope)
        if create_aux_logits and num_classes:
            with store.scope(scope + "/AuxLogits"):
                aux = ops.avg_pool2d(end_points["Mixed_6e"], [5, 5], stride=3, padding="SAME", scope="AvgPool_1a_5x5")
                aux = ops.convolution2d(aux, depth(128), [1, 1], scope="Conv2d_1b_1x1")
                k = _small(aux, [5, 5])
                aux = ops.convolution2d(aux, depth(768), k, padding="SAME",
                                        weights_initializer=V.truncated_normal_initializer(0.01),
                                        scope="Conv2d_2a_{}x{}".format(*k))
                aux = ops.convolution2d(aux, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                        weights_initializer=V.truncated_normal_initializer(0.001),
                                        scope="Conv2d_2b_1x1")
                end_points["AuxLogits"] = _squeeze(aux, spatial_squeeze)
        with store.scope(scope + "/Logits"):
            if global_pool:
                net = end_points["global_pool"] = net.mean((1, 2), keepdim=True)
            else:
                k = _small(net, [8, 8])
                net = end_points["AvgPool_1a"] = ops.avg_pool2d(net, k, stride=1, padding="SAME",
                                                                scope="AvgPool_1a_{}x{}".format(*k))
            if not num_classes:
                return net, end_points
            net = end_points["PreLogits"] = ops.dropout(net, dropout_keep_prob, scope="Dropout_1b")
            logits = _squeeze(ops.convolution2d(net, num_classes, [1, 1], activation_fn=None, normalizer_fn=None,
                                                scope="Conv2d_1c_1x1"), spatial_squeeze)
        end_points["Logits"] = logits
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


inception_v3.default_image_size = 299


# ---- Inception v4 ----------------------------------------------------------------------------------------
_V4_A = ([C(96, 1, "Conv2d_0a_1x1")],
         [C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_0b_3x3")],
         [C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_0b_3x3"), C(96, 3, "Conv2d_0c_3x3")],
         [AP(3, "AvgPool_0a_3x3"), C(96, 1, "Conv2d_0b_1x1")])
_V4_RA = ([C(384, 3, "Conv2d_1a_3x3", 2, "VALID")],
          [C(192, 1, "Conv2d_0a_1x1"), C(224, 3, "Conv2d_0b_3x3"), C(256, 3, "Conv2d_1a_3x3", 2, "VALID")],
          [MP(3, "MaxPool_1a_3x3", 2, "VALID")])
_V4_B = ([C(384, 1, "Conv2d_0a_1x1")],
         [C(192, 1, "Conv2d_0a_1x1"), C(224, (1, 7), "Conv2d_0b_1x7"), C(256, (7, 1), "Conv2d_0c_7x1")],
         [C(192, 1, "Conv2d_0a_1x1"), C(192, (7, 1), "Conv2d_0b_7x1"), C(224, (1, 7), "Conv2d_0c_1x7"),
          C(224, (7, 1), "Conv2d_0d_7x1"), C(256, (1, 7), "Conv2d_0e_1x7")],
         [AP(3, "AvgPool_0a_3x3"), C(128, 1, "Conv2d_0b_1x1")])
_V4_RB = ([C(192, 1, "Conv2d_0a_1x1"), C(192, 3, "Conv2d_1a_3x3", 2, "VALID")],
          [C(256, 1, "Conv2d_0a_1x1"), C(256, (1, 7), "Conv2d_0b_1x7"), C(320, (7, 1), "Conv2d_0c_7x1"),
           C(320, 3, "Conv2d_1a_3x3", 2, "VALID")],
          [MP(3, "MaxPool_1a_3x3", 2, "VALID")])
_V4_C = ([C(256, 1, "Conv2d_0a_1x1")],
         [C(384, 1, "Conv2d_0a_1x1"), FORK(C(256, (1, 3), "Conv2d_0b_1x3"), C(256, (3, 1), "Conv2d_0c_3x1"))],
         [C(384, 1, "Conv2d_0a_1x1"), C(448, (3, 1), "Conv2d_0b_3x1"), C(512, (1, 3), "Conv2d_0c_1x3"),
          FORK(C(256, (1, 3), "Conv2d_0d_1x3"), C(256, (3, 1), "Conv2d_0e_3x1"))],
         [AP(3, "AvgPool_0a_3x3"), C(256, 1, "Conv2d_0b_1x1")])

_V4 = ([("Conv2d_1a_3x3", C(32, 3, "Conv2d_1a_3x3", 2, "VALID")), ("Conv2d_2a_3x3", C(32, 3, "Conv2d_2a_3x3", 1, "VALID")),
        ("Conv2d_2b_3x3", C(64, 3, "Conv2d_2b_3x3")),
        ("Mixed_3a", ([MP(3, "MaxPool_0a_3x3", 2, "VALID")], [C(96, 3, "Conv2d_0a_3x3", 2, "VALID")])),
        ("Mixed_4a", ([C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_1a_3x3", 1, "VALID")],
                      [C(64, 1, "Conv2d_0a_1x1"), C(64, (1, 7), "Conv2d_0b_1x7"), C(64, (7, 1), "Conv2d_0c_7x1"),
                       C(96, 3, "Conv2d_1a_3x3", 1, "VALID")])),
        ("Mixed_5a", ([C(192, 3, "Conv2d_1a_3x3", 2, "VALID")], [MP(3, "MaxPool_1a_3x3", 2, "VALID")]))]
       + [("Mixed_5" + c, _V4_A) for c in "bcde"] + [("Mixed_6a", _V4_RA)]
       + [("Mixed_6" + c, _V4_B) for c in "bcdefgh"] + [("Mixed_7a", _V4_RB)]
       + [("Mixed_7" + c, _V4_C) for c in "bcd"])


def inception_v4_base(inputs, final_endpoint: str = "Mixed_7d", scope: str = "InceptionV4"):
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with store.scope(scope):
        net, hit = _run_table(inputs, _V4, lambda d: d, "SAME", final_endpoint, end_points)
    if not hit:
        raise ValueError("Unknown final endpoint %s" % final_endpoint)
    return net, end_points


def inception_v4(inputs, num_classes: int = 1001, is_training: bool = True, dropout_keep_prob: float = 0.8,
                 create_aux_logits: bool = True, scope: str = "InceptionV4"):
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        net, end_points = inception_v4_base(inputs, scope=scope)
        if create_aux_logits and num_classes:
            with store.scope(scope + "/AuxLogits"):
                aux = ops.avg_pool2d(end_points["Mixed_6h"], [5, 5], stride=3, padding="VALID", scope="AvgPool_1a_5x5")
                aux = ops.convolution2d(aux, 128, [1, 1], padding="SAME", scope="Conv2d_1b_1x1")
                aux = ops.convolution2d(aux, 768, [aux.shape[1], aux.shape[2]], padding="VALID", scope="Conv2d_2a")
                aux = ops.fully_connected(ops.flatten(aux), num_classes, activation_fn=None, scope="Aux_logits")
                end_points["AuxLogits"] = aux
        with store.scope(scope + "/Logits"):
            net = end_points["global_pool"] = ops.avg_pool2d(net, [net.shape[1], net.shape[2]], stride=1,
                                                             padding="VALID", scope="AvgPool_1a")
            if not num_classes:
                return net, end_points
            net = ops.dropout(net, dropout_keep_prob, scope="Dropout_1b")
            net = end_points["PreLogitsFlatten"] = ops.flatten(net)
            logits = end_points["Logits"] = ops.fully_connected(net, num_classes, activation_fn=None, scope="Logits")
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


inception_v4.default_image_size = 299


# ---- Inception-ResNet v2 ---------------------------------------------------------------------------------
def _res_block(default_name: str, branches):
    """Residual Inception unit: concat(branches) -> linear 1x1 back to the input depth (bias, no BN) -> scaled add."""
    def block(net, scale: float = 1.0, activation_fn=ops.relu, scope: Optional[str] = None):
        store = V.current_store()
        with store.scope(scope, default_name):
            outs = []
            for i, branch in enumerate(branches):
                with store.scope(f"Branch_{i}"):
                    b = net
                    for spec in branch:
                        b = _run_layer(b, spec, lambda d: d, "SAME")
                    outs.append(b)
            up = ops.convolution2d(torch.cat(outs, dim=3), net.shape[-1], 1, normalizer_fn=None, activation_fn=None,
                                   scope="Conv2d_1x1")
            up = up * scale
            if activation_fn is ops.relu6:
                up = up.clamp(-6.0, 6.0)
            net = net + up
            return activation_fn(net) if activation_fn else net
    block.__name__ = default_name.lower()
    return block


block35 = _res_block("Block35", ([C(32, 1, "Conv2d_1x1")],
                                 [C(32, 1, "Conv2d_0a_1x1"), C(32, 3, "Conv2d_0b_3x3")],
                                 [C(32, 1, "Conv2d_0a_1x1"), C(48, 3, "Conv2d_0b_3x3"), C(64, 3, "Conv2d_0c_3x3")]))
block17 = _res_block("Block17", ([C(192, 1, "Conv2d_1x1")],
                                 [C(128, 1, "Conv2d_0a_1x1"), C(160, (1, 7), "Conv2d_0b_1x7"),
                                  C(192, (7, 1), "Conv2d_0c_7x1")]))
block8 = _res_block("Block8", ([C(192, 1, "Conv2d_1x1")],
                               [C(192, 1, "Conv2d_0a_1x1"), C(224, (1, 3), "Conv2d_0b_1x3"),
                                C(256, (3, 1), "Conv2d_0c_3x1")]))


def inception_resnet_v2_base(inputs, final_endpoint: str = "Conv2d_7b_1x1", output_stride: int = 16,
                             align_feature_maps: bool = False, activation_fn=ops.relu,
                             scope: str = "InceptionResnetV2"):
    if output_stride not in (8, 16):
        raise ValueError("output_stride must be 8 or 16.")
    pad = "SAME" if align_feature_maps else "VALID"
    atrous = output_stride == 8
    store = V.current_store()
    ep: Dict[str, torch.Tensor] = {}
    stem = [("Conv2d_1a_3x3", C(32, 3, "Conv2d_1a_3x3", 2, pad)), ("Conv2d_2a_3x3", C(32, 3, "Conv2d_2a_3x3", 1, pad)),
            ("Conv2d_2b_3x3", C(64, 3, "Conv2d_2b_3x3")), ("MaxPool_3a_3x3", MP(3, "MaxPool_3a_3x3", 2, pad)),
            ("Conv2d_3b_1x1", C(80, 1, "Conv2d_3b_1x1", 1, pad)), ("Conv2d_4a_3x3", C(192, 3, "Conv2d_4a_3x3", 1, pad)),
            ("MaxPool_5a_3x3", MP(3, "MaxPool_5a_3x3", 2, pad)),
            ("Mixed_5b", ([C(96, 1, "Conv2d_1x1")], [C(48, 1, "Conv2d_0a_1x1"), C(64, 5, "Conv2d_0b_5x5")],
                          [C(64, 1, "Conv2d_0a_1x1"), C(96, 3, "Conv2d_0b_3x3"), C(96, 3, "Conv2d_0c_3x3")],
                          [AP(3, "AvgPool_0a_3x3", 1, "SAME"), C(64, 1, "Conv2d_0b_1x1")]))]
    s6 = 1 if atrous else 2
    with store.scope(scope):
        net, hit = _run_table(inputs, stem, lambda d: d, "SAME", final_endpoint, ep)
        if hit:
            return net, ep
        net = ops.repeat(net, 10, block35, scale=0.17, activation_fn=activation_fn)
        net = _run_mixed(net, ([C(384, 3, "Conv2d_1a_3x3", s6, pad)],
                               [C(256, 1, "Conv2d_0a_1x1"), C(256, 3, "Conv2d_0b_3x3"),
                                C(384, 3, "Conv2d_1a_3x3", s6, pad)],
                               [MP(3, "MaxPool_1a_3x3", s6, pad)]), lambda d: d, "SAME", "Mixed_6a")
        ep["Mixed_6a"] = net
        if final_endpoint == "Mixed_6a":
            return net, ep
        with scopes.arg_scope([ops.convolution2d], rate=2 if atrous else 1):
            net = ops.repeat(net, 20, block17, scale=0.10, activation_fn=activation_fn)
        ep["PreAuxLogits"] = net
        if final_endpoint == "PreAuxLogits":
            return net, ep
        if atrous:
            raise ValueError("output_stride==8 is only supported up to the PreAuxlogits end_point for now.")
        net = _run_mixed(net, ([C(256, 1, "Conv2d_0a_1x1"), C(384, 3, "Conv2d_1a_3x3", 2, pad)],
                               [C(256, 1, "Conv2d_0a_1x1"), C(288, 3, "Conv2d_1a_3x3", 2, pad)],
                               [C(256, 1, "Conv2d_0a_1x1"), C(288, 3, "Conv2d_0b_3x3"),
                                C(320, 3, "Conv2d_1a_3x3", 2, pad)],
                               [MP(3, "MaxPool_1a_3x3", 2, pad)]), lambda d: d, "SAME", "Mixed_7a")
        ep["Mixed_7a"] = net
        if final_endpoint == "Mixed_7a":
            return net, ep
        net = ops.repeat(net, 9, block8, scale=0.20, activation_fn=activation_fn)
        net = block8(net, activation_fn=None)
        net = ep["Conv2d_7b_1x1"] = ops.convolution2d(net, 1536, 1, scope="Conv2d_7b_1x1")
        if final_endpoint == "Conv2d_7b_1x1":
            return net, ep
    raise ValueError("final_endpoint (%s) not recognized" % final_endpoint)


def inception_resnet_v2(inputs, num_classes: int = 1001, is_training: bool = True, dropout_keep_prob: float = 0.8,
                        create_aux_logits: bool = True, activation_fn=ops.relu, scope: str = "InceptionResnetV2"):
    store = V.current_store()
    with scopes.arg_scope([ops.batch_norm, ops.dropout], is_training=is_training):
        net, end_points = inception_resnet_v2_base(inputs, activation_fn=activation_fn, scope=scope)
        if create_aux_logits and num_classes:
            with store.scope(scope + "/AuxLogits"):
                aux = ops.avg_pool2d(end_points["PreAuxLogits"], 5, stride=3, padding="VALID", scope="Conv2d_1a_3x3")
                aux = ops.convolution2d(aux, 128, 1, scope="Conv2d_1b_1x1")
                aux = ops.convolution2d(aux, 768, [aux.shape[1], aux.shape[2]], padding="VALID", scope="Conv2d_2a_5x5")
                aux = ops.fully_connected(ops.flatten(aux), num_classes, activation_fn=None, scope="Logits")
                end_points["AuxLogits"] = aux
        with store.scope(scope + "/Logits"):
            net = end_points["global_pool"] = ops.avg_pool2d(net, [net.shape[1], net.shape[2]], stride=1,
                                                             padding="VALID", scope="AvgPool_1a_8x8")
            if not num_classes:
                return net, end_points
            net = ops.flatten(net)
            net = end_points["PreLogitsFlatten"] = ops.dropout(net, dropout_keep_prob, scope="Dropout")
            logits = end_points["Logits"] = ops.fully_connected(net, num_classes, activation_fn=None, scope="Logits")
        end_points["Predictions"] = ops.softmax(logits)
    store.end_points = end_points
    return logits, end_points


inception_resnet_v2.default_image_size = 299
