"""Evaluation daemons (role of ``cnn/cifar10_eval.py``, ``*/cifar10_*_eval.py``, ``inception/inception_eval.py``;
SURVEY §2.7, call stack §3.5).

``evaluate()`` loops every ``eval_interval_secs``: find the newest checkpoint through the ``checkpoint`` state
file, restore it (optionally the EMA shadow values — cnn and inception evals do, cnn/cifar10_eval.py:132-135,
inception_eval.py:157-160), run ``ceil(num_examples / batch)`` batches, print

    <datetime>: precision @ 1 = 0.xxx, global_step: N              (+ "recall @ 5" for ImageNet)

and append the numbers to ``<eval_dir>/eval.jsonl`` (the TF summary of the reference).  Training and eval
communicate only through the checkpoint directory, as in the reference.  Unlike the reference's wait loops
(which never re-read the state) the checkpoint state is re-polled.
"""
from __future__ import annotations

import json
import math
import os
import time
from typing import Optional

import torch

from .. import config as cfg
from ..ckpt import store as ckpt
from ..models.registry import MODEL_SPECS, spec_for
from ..ops import functional as Fn
from ..parallel.engine import TrainSetup, Trainer
from ..utils.logs import format_precision, log
from ..utils.tfevents import EventFileWriter

# per-model eval defaults (reference flag defaults)
EVAL_DEFAULTS = {
    "cnn": dict(num_examples=10000, interval=60.0, use_ema=True, batch=128),          # cnn/cifar10_eval.py:49-60
    "cifarnet": dict(num_examples=1000, interval=60.0, use_ema=False, batch=128),     # cifarnet eval :58
    "alexnet": dict(num_examples=1000, interval=60.0, use_ema=False, batch=128),
    "vgg": dict(num_examples=1000, interval=60.0, use_ema=False, batch=72),           # the reference has none
    "resnet": dict(num_examples=10000, interval=60.0, use_ema=False, batch=128),      # resnet eval :26
    "inception": dict(num_examples=200, interval=300.0, use_ema=True, batch=32),      # inception_eval.py:43,50
}


# any other zoo network: vgg/nets/mobilenet_v1_eval.py:31-39 (one pass over 50,000 validation images, batch 250,
# raw variables, accuracy + recall@5)
ZOO_EVAL_DEFAULTS = dict(num_examples=50000, interval=60.0, use_ema=False, batch=250)


def _eval_batches(args, spec, batch_size: int):
    if spec.dataset == "cifar10":
        from ..data import cifar10
        if not args.synthetic and cifar10.dataset_available(args.data_dir):
            return cifar10.inputs(args.eval_data == "test", args.data_dir, batch_size, spec.image_size, repeat=True)
        from ..data.synthetic import SyntheticBatches
        return SyntheticBatches(batch_size, spec.image_size, spec.num_classes, seed=999, pin=False)
    from ..data import imagenet
    ds = imagenet.ImagenetData("validation", args.data_dir)
    if not args.synthetic and ds.available():
        return imagenet.inputs(ds, batch_size, image_size=spec.image_size)
    from ..data.synthetic import SyntheticBatches
    return SyntheticBatches(batch_size, spec.image_size, spec.num_classes, seed=999, pin=False)


@torch.no_grad()
def eval_once(trainer: Trainer, args, ckpt_path: str, use_ema: bool, num_examples: int, batch_size: int,
              batches, training_bn: bool = False) -> dict:
    tensors = ckpt.load(ckpt_path)
    ckpt.restore(trainer.store, tensors, use_ema=use_ema, strict=False)
    global_step = ckpt.step_of(ckpt_path)
    num_iter = int(math.ceil(num_examples / batch_size))
    top1 = top5 = total = 0
    t0 = time.time()
    for it in range(num_iter):
        images, labels = next(batches)
        images, labels = images.to(trainer.device), labels.to(trainer.device)
        _, logits = trainer.evaluate_batch(images, labels, training_bn=training_bn)
        top1 += int(Fn.in_top_k(logits, labels, 1).sum())
        if logits.shape[1] > 5:
            top5 += int(Fn.in_top_k(logits, labels, 5).sum())
        total += labels.numel()
        if trainer.setup.model == "inception" and it % 20 == 19:
            dt = time.time() - t0
            log("[%d batches out of %d] (%.1f examples/sec; %.3f sec/batch)" % (it + 1, num_iter, 20 * batch_size / dt,
                                                                               dt / 20))
            t0 = time.time()
    res = {"precision_at_1": top1 / max(total, 1), "global_step": global_step, "examples": total, "time": time.time()}
    log(format_precision(res["precision_at_1"], global_step))
    if trainer.setup.model == "inception" or (trainer.spec.dataset == "imagenet" and logits.shape[1] > 5):
        res["recall_at_5"] = top5 / max(total, 1)
        log("precision @ 1 = %.4f recall @ 5 = %.4f [%d examples]" % (res["precision_at_1"], res["recall_at_5"], total))
    return res


def evaluate(args) -> Optional[dict]:
    d = EVAL_DEFAULTS.get(args.model, ZOO_EVAL_DEFAULTS)
    spec = spec_for(args.model, args.image_size, args.num_classes)
    num_examples = args.num_examples or d["num_examples"]
    interval = args.eval_interval_secs if args.eval_interval_secs is not None else d["interval"]
    batch_size = min(args.batch_size or d["batch"], d["batch"])
    checkpoint_dir = args.checkpoint_dir or args.train_dir
    device = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")   # train.sh runs eval on CPU
    trainer = Trainer(TrainSetup(model=args.model, batch_size=batch_size, resnet_size=args.resnet_size,
                                 ref_compat=args.ref_compat, cuda_graph=False, image_size=args.image_size,
                                 num_classes=args.num_classes, weight_decay=args.weight_decay,
                                 quantize=bool(args.quantize)), device=device)
    batches = _eval_batches(args, spec, batch_size)
    os.makedirs(args.eval_dir, exist_ok=True)
    summaries = EventFileWriter(args.eval_dir)       # TensorBoard event file, as the reference's FileWriter(eval_dir)
    # ResNet's reference eval builds the network in *training* mode (batch-statistics BN, Q3): only under --ref_compat
    training_bn = bool(args.ref_compat and args.model == "resnet")
    last, res = None, None
    while True:
        path = ckpt.wait_for_checkpoint(checkpoint_dir, poll_secs=1.0, timeout=None if not args.run_once else 120.0)
        if path is None:
            log("No checkpoint file found")
            return None
        if path != last:
            res = eval_once(trainer, args, path, d["use_ema"], num_examples, batch_size, batches, training_bn)
            with open(os.path.join(args.eval_dir, "eval.jsonl"), "a") as f:
                f.write(json.dumps(res) + "\n")
            scalars = {"Precision @ 1": res["precision_at_1"]}
            if "recall_at_5" in res:
                scalars["Recall @ 5"] = res["recall_at_5"]
            summaries.add_scalars(res["global_step"], scalars)
            summaries.flush()
            last = path
        if args.run_once:
            return res
        time.sleep(interval)


def main(argv=None, model: Optional[str] = None) -> int:
    args = cfg.parse_args(argv, model, "bsp")
    evaluate(args)
    return 0
