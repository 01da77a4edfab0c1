"""Inception-v3 (the 2016 in-tree slim variant the reference trains on ImageNet).

Architecture, scopes and variable names follow ``inception/slim/inception_model.py:54-332``; the training
wrapper (`inference` / `loss`) follows ``inception/inception_model.py:49-136``.  The network is described
here as data — a table of blocks, each a set of named branches made of layer specs — and a small
interpreter walks it.  That keeps the census the reference's tests pin down (388 variables: 98 weights, 2
biases, 96 beta, 96+96 moving statistics) in one auditable place and lets the native backend
recognise whole blocks: ``_run_block`` allocates the block's output once and every branch's last kernel (the
batch-norm apply behind its final convolution, or its pooling kernel) writes its channel slice of it with the
block's pixel pitch, so no concat kernel runs; the backward pass reads the slices of the incoming
gradient through the same pitch.

Layer spec grammar (tuples):
    ("conv", C_out, (kh, kw))                      stride 1, block-default padding
    ("conv", C_out, (kh, kw), stride, padding)
    ("avg", (kh, kw)) / ("max", (kh, kw), stride, padding)
    ("fork", [spec_a...], [spec_b...])             run both on the same input and concat (the 1x3 | 3x1 pair)
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..ops import functional as Fn
from ..ops import streams
from ..slim import losses, ops, scopes
from ..slim import variables as V

# ---- block table ------------------------------------------------------------------------------------


def _mixed_35(pool_ch: int):
    return [
        ("branch1x1", [("conv", 64, (1, 1))]),
        ("branch5x5", [("conv", 48, (1, 1)), ("conv", 64, (5, 5))]),
        ("branch3x3dbl", [("conv", 64, (1, 1)), ("conv", 96, (3, 3)), ("conv", 96, (3, 3))]),
        ("branch_pool", [("avg", (3, 3)), ("conv", pool_ch, (1, 1))]),
    ]


def _mixed_17(mid: int):
    return [
        ("branch1x1", [("conv", 192, (1, 1))]),
        ("branch7x7", [("conv", mid, (1, 1)), ("conv", mid, (1, 7)), ("conv", 192, (7, 1))]),
        ("branch7x7dbl", [("conv", mid, (1, 1)), ("conv", mid, (7, 1)), ("conv", mid, (1, 7)),
                          ("conv", mid, (7, 1)), ("conv", 192, (1, 7))]),
        ("branch_pool", [("avg", (3, 3)), ("conv", 192, (1, 1))]),
    ]


def _mixed_8():
    pair = ("fork", [("conv", 384, (1, 3))], [("conv", 384, (3, 1))])
    return [
        ("branch1x1", [("conv", 320, (1, 1))]),
        ("branch3x3", [("conv", 384, (1, 1)), pair]),
        ("branch3x3dbl", [("conv", 448, (1, 1)), ("conv", 384, (3, 3)), pair]),
        ("branch_pool", [("avg", (3, 3)), ("conv", 192, (1, 1))]),
    ]


BLOCKS: List[Tuple[str, list]] = [
    ("mixed_35x35x256a", _mixed_35(32)),
    ("mixed_35x35x288a", _mixed_35(64)),
    ("mixed_35x35x288b", _mixed_35(64)),
    ("mixed_17x17x768a", [
        ("branch3x3", [("conv", 384, (3, 3), 2, "VALID")]),
        ("branch3x3dbl", [("conv", 64, (1, 1)), ("conv", 96, (3, 3)), ("conv", 96, (3, 3), 2, "VALID")]),
        ("branch_pool", [("max", (3, 3), 2, "VALID")]),
    ]),
    ("mixed_17x17x768b", _mixed_17(128)),
    ("mixed_17x17x768c", _mixed_17(160)),
    ("mixed_17x17x768d", _mixed_17(160)),
    ("mixed_17x17x768e", _mixed_17(192)),
    ("mixed_17x17x1280a", [
        ("branch3x3", [("conv", 192, (1, 1)), ("conv", 320, (3, 3), 2, "VALID")]),
        ("branch7x7x3", [("conv", 192, (1, 1)), ("conv", 192, (1, 7)), ("conv", 192, (7, 1)),
                         ("conv", 192, (3, 3), 2, "VALID")]),
        ("branch_pool", [("max", (3, 3), 2, "VALID")]),
    ]),
    ("mixed_8x8x2048a", _mixed_8()),
    ("mixed_8x8x2048b", _mixed_8()),
]

AUX_AFTER = "mixed_17x17x768e"


def _spec_channels(specs: Sequence[tuple], cin: int) -> int:
    """Output channels of a spec chain."""
    c = cin
    for sp in specs:
        if sp[0] == "conv":
            c = sp[1]
        elif sp[0] == "fork":
            c = _spec_channels(sp[1], c) + _spec_channels(sp[2], c)
    return c


def _run_specs(x: torch.Tensor, specs: Sequence[tuple], out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out``: the channel slice of the enclosing block's output buffer that this chain's result belongs in (native
    backend).  The chain's last op writes straight into it — its batch-norm apply / pooling kernel takes the slice
    through ``native_ops.set_output_slot`` — so the block never runs a concat kernel."""
    from ..ops import native_ops
    for i, sp in enumerate(specs):
        kind = sp[0]
        last = out is not None and i == len(specs) - 1
        if kind == "fork":
            if last:
                ca = _spec_channels(sp[1], x.shape[-1])
                oa, ob = out[..., :ca], out[..., ca:]
                outs = streams.run_branches([lambda a=sp[1], v=x, o=oa: _run_specs(v, a, o),
                                             lambda b=sp[2], v=x, o=ob: _run_specs(v, b, o)], x)
                if all(o_.data_ptr() == t_.data_ptr() for o_, t_ in zip(outs, (oa, ob))):
                    x = native_ops.concat_views(out, outs)
                else:
                    x = Fn.concat_channels(outs)
            else:
                x = Fn.concat_channels(streams.run_branches([lambda a=sp[1], v=x: _run_specs(v, a),
                                                             lambda b=sp[2], v=x: _run_specs(v, b)], x))
            continue
        if last:
            native_ops.set_output_slot(out)
        try:
            if kind == "conv":
                if len(sp) == 3:
                    x = ops.conv2d(x, sp[1], list(sp[2]))
                else:
                    x = ops.conv2d(x, sp[1], list(sp[2]), stride=sp[3], padding=sp[4])
            elif kind == "avg":
                x = ops.avg_pool(x, list(sp[1]))
            elif kind == "max":
                x = ops.max_pool(x, list(sp[1]), stride=sp[2], padding=sp[3])
            else:
                raise ValueError(kind)
        finally:
            if last:
                native_ops.set_output_slot(None)     # (an op that does not write into slices leaves it untouched)
    return x


def _block_geometry(x: torch.Tensor, branches) -> Optional[Tuple[int, int, List[int]]]:
    """(OH, OW, channels per branch) of a block whose branches keep or (3x3 / 2 VALID) reduce the map."""
    n, h, w, c = x.shape
    reduce = any(len(sp) > 3 and sp[0] in ("conv", "max") and (sp[3] if sp[0] == "conv" else sp[2]) == 2
                 for _, specs in branches for sp in specs if sp[0] != "fork")
    oh, ow = ((h - 3) // 2 + 1, (w - 3) // 2 + 1) if reduce else (h, w)
    return oh, ow, [_spec_channels(specs, c) for _, specs in branches]


def _run_block(store, net: torch.Tensor, branches) -> torch.Tensor:
    """All branches of one Inception block on ``net``; the outputs end up side by side along the channels."""
    sliced = net.is_cuda and Fn.use_native(net)
    if sliced:
        oh, ow, widths = _block_geometry(net, branches)
        buf = torch.empty((net.shape[0], oh, ow, sum(widths)), dtype=net.dtype, device=net.device)
        offs = [sum(widths[:i]) for i in range(len(widths))]
        slots = [buf[..., o:o + w_] for o, w_ in zip(offs, widths)]
    else:
        slots = [None] * len(branches)

    def _branch(bname, specs, x, slot):
        def run():
            with store.scope(bname):
                return _run_specs(x, specs, slot)
        return run
    # independent branches: side streams on CUDA (ops/streams.py), plain sequence otherwise
    outs = streams.run_branches([_branch(b, sp, net, sl) for (b, sp), sl in zip(branches, slots)], net)
    if sliced and all(o.shape == sl.shape and o.data_ptr() == sl.data_ptr() for o, sl in zip(outs, slots)):
        from ..ops import native_ops
        return native_ops.concat_views(buf, outs)
    return Fn.concat_channels(outs)


def inception_v3(inputs: torch.Tensor, dropout_keep_prob: float = 0.8, num_classes: int = 1000,
                 is_training: bool = True, restore_logits: bool = True, scope: str = ""):
    """-> (logits, end_points); end_points['aux_logits'] holds the auxiliary head.
    Call inside ``arg_scope`` frames that provide BN params / weight decay (see :func:`inference`)."""
    store = V.current_store()
    end_points: Dict[str, torch.Tensor] = {}
    with scopes.arg_scope([ops.conv2d, ops.fc, ops.batch_norm, ops.dropout], is_training=is_training):
        # ---- stem: 299 -> 35 (inception_model.py:88-108) -------------------------------------------
        with scopes.arg_scope([ops.conv2d, ops.max_pool, ops.avg_pool], stride=1, padding="VALID"):
            net = end_points["conv0"] = ops.conv2d(inputs, 32, [3, 3], stride=2, scope="conv0")
            net = end_points["conv1"] = ops.conv2d(net, 32, [3, 3], scope="conv1")
            net = end_points["conv2"] = ops.conv2d(net, 64, [3, 3], padding="SAME", scope="conv2")
            net = end_points["pool1"] = ops.max_pool(net, [3, 3], stride=2, scope="pool1")
            net = end_points["conv3"] = ops.conv2d(net, 80, [1, 1], scope="conv3")
            net = end_points["conv4"] = ops.conv2d(net, 192, [3, 3], scope="conv4")
            net = end_points["pool2"] = ops.max_pool(net, [3, 3], stride=2, scope="pool2")
        # ---- inception blocks ------------------------------------------------------------------------
        with scopes.arg_scope([ops.conv2d, ops.max_pool, ops.avg_pool], stride=1, padding="SAME"):
            for name, branches in BLOCKS:
                with store.scope(name):
                    net = end_points[name] = _run_block(store, net, branches)
                if name == AUX_AFTER:
                    # auxiliary classifier (inception_model.py:251-263); its handful of small kernels run beside the
                    # remaining blocks of the trunk (joined before the function returns)
                    def _aux(x=net):
                        with store.scope("aux_logits"):
                            aux = ops.avg_pool(x, [5, 5], stride=3, padding="VALID")
                            aux = ops.conv2d(aux, 128, [1, 1], scope="proj")
                            aux = ops.conv2d(aux, 768, [aux.shape[1], aux.shape[2]], stddev=0.01, padding="VALID")
                            aux = ops.flatten(aux)
                            return ops.fc(aux, num_classes, activation=None, stddev=0.001, restore=restore_logits)
                    aux_head = streams.Detached(_aux, net)
            with store.scope("logits"):
                net = ops.avg_pool(net, [net.shape[1], net.shape[2]], padding="VALID", scope="pool")
                net = ops.dropout(net, dropout_keep_prob, scope="dropout")
                net = ops.flatten(net, scope="flatten")
                logits = ops.fc(net, num_classes, activation=None, scope="logits", restore=restore_logits)
                end_points["logits"] = logits
                end_points["predictions"] = torch.softmax(logits.detach(), dim=-1)
            end_points["aux_logits"] = aux_head.join()
    store.end_points = end_points
    return logits, end_points


def inception_v3_parameters(weight_decay: float = 0.00004, stddev: float = 0.1,
                            batch_norm_decay: float = 0.9997, batch_norm_epsilon: float = 0.001):
    """Context manager yielding the arg_scope with the stock Inception-v3 hyper-parameters
    (reference slim/inception_model.py:335-358)."""
    import contextlib

    @contextlib.contextmanager
    def cm():
        with scopes.arg_scope([ops.conv2d, ops.fc], weight_decay=weight_decay):
            with scopes.arg_scope([ops.conv2d], stddev=stddev, activation=ops.relu,
                                  batch_norm_params={"decay": batch_norm_decay,
                                                     "epsilon": batch_norm_epsilon}) as sc:
                yield sc
    return cm()


# ---- training wrapper (inception/inception_model.py) --------------------------------------------------
BATCHNORM_MOVING_AVERAGE_DECAY = 0.9997
MOVING_AVERAGE_DECAY = 0.9999


def inference(images: torch.Tensor, num_classes: int, for_training: bool = False, restore_logits: bool = True,
              scope: Optional[str] = None):
    """-> (logits, aux_logits).  conv = conv(no bias) -> BN(decay .9997, eps .001, centre only) -> ReLU;
    weight decay 4e-5 on conv and FC weights; conv init trunc-normal sigma 0.1
    (reference inception/inception_model.py:49-96)."""
    batch_norm_params = {"decay": BATCHNORM_MOVING_AVERAGE_DECAY, "epsilon": 0.001}
    with scopes.arg_scope([ops.conv2d, ops.fc], weight_decay=0.00004):
        with scopes.arg_scope([ops.conv2d], stddev=0.1, activation=ops.relu, batch_norm_params=batch_norm_params):
            logits, end_points = inception_v3(images, dropout_keep_prob=0.8, num_classes=num_classes,
                                              is_training=for_training, restore_logits=restore_logits,
                                              scope=scope or "")
    return logits, end_points["aux_logits"]


def loss(logits, labels: torch.Tensor, batch_size: Optional[int] = None) -> torch.Tensor:
    """Main + auxiliary label-smoothed (0.1) cross-entropies, weights 1.0 and 0.4
    (reference inception/inception_model.py:99-136).  ``logits`` = (logits, aux_logits).
    Returns their sum; the regularisation term is added by the trainer (…_bsp.py:112-115)."""
    main, aux = logits
    l_main = losses.cross_entropy_loss(main, labels, label_smoothing=0.1, weight=1.0)
    l_aux = losses.cross_entropy_loss(aux, labels, label_smoothing=0.1, weight=0.4, scope="aux_loss")
    return l_main + l_aux
