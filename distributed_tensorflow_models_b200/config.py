"""Flags and per-model hyper-parameter presets.

One argparse namespace replaces the reference's scattered ``tf.app.flags`` definitions: the
same flag names and defaults are accepted (``--job_name --task_id --ps_hosts --worker_hosts --batch_size
--data_dir --train_dir --max_steps --resnet_size --protocol --subset --save_interval_secs
--initial_learning_rate --num_epochs_per_decay --learning_rate_decay_factor --eval_dir --checkpoint_dir
--eval_interval_secs --num_examples --run_once …``) plus the switches that were module constants or did not
exist: ``--model --mode --staleness --synthetic --ref_compat --resume --trace --backend --placement``.

``PRESETS`` reproduces the constants buried in each trainer script: learning rate and
whether it is multiplied by the number of workers, decay factor / epochs, optimiser, weight-decay policy,
EMA, checkpoint policy, per-worker batch size.
"""
from __future__ import annotations

import argparse
from dataclasses import dataclass, field, replace
from typing import Callable, Dict, List, Optional

MODELS = ("cnn", "cifarnet", "alexnet", "vgg", "resnet", "inception")
MODES = ("bsp", "asp", "ssp")

PS_PORT = 22222       # train.sh:15
WORKER_PORT = 22223   # train.sh:23
SSP_PORT = 8000       # inception/ssp_manager.py:43


@dataclass(frozen=True)
class Preset:
    model: str
    batch_size: int
    max_steps: int
    initial_lr: float
    lr_times_workers_in_bsp: bool
    decay_factor: float
    epochs_per_decay: float
    examples_per_epoch: int
    optimizer: str = "sgd"                 # 'sgd' | 'rmsprop'
    weight_decay: Optional[float] = None   # None: keep the per-variable regularisers of the model
    decay_all_trainables: bool = False     # L2 over *all* trainable variables
    ema_decay: Optional[float] = 0.9999
    save_secs: Optional[float] = 60.0      # None: never checkpoint (VGG trainers pass saver=None)
    max_to_keep: int = 5
    wipe_train_dir: bool = True            # CIFAR chiefs delete train_dir at start
    nan_guard: bool = False
    accuracy_every: int = 0                # ResNet logs train-batch accuracy every 200 steps
    log_style: str = "cifar"               # 'cnn' | 'cifar' | 'resnet'
    decay_steps_divides_by_workers: bool = False   # inception: decay_steps / num_workers


PRESETS: Dict[str, Preset] = {
    # cnn/cifar10_cnn_bsp.py:9-13,46-67 — LR 0.32, *not* multiplied by N
    "cnn": Preset("cnn", 512, 20000, 0.32, False, 0.1, 350.0, 50000, log_style="cnn"),
    # cifarnet/cifar10_cifarnet_bsp.py:20-32
    "cifarnet": Preset("cifarnet", 512, 20000, 0.1, True, 0.1, 20.0, 50000, weight_decay=2e-4,
                       decay_all_trainables=True),
    # alexnet/cifar10_alexnet_bsp.py:18-30
    "alexnet": Preset("alexnet", 128, 20000, 0.01, True, 0.1, 350.0, 50000, weight_decay=2e-4,
                      decay_all_trainables=True),
    # vgg/cifar10_vgg_bsp.py:19-30,85-103 — no EMA in BSP, never saves
    "vgg": Preset("vgg", 72, 20000, 0.01, True, 0.1, 350.0, 50000, weight_decay=2e-4, decay_all_trainables=True,
                  ema_decay=None, save_secs=None),
    # resnet/cifar10_resnet_bsp.py:19-29,107,146-148
    "resnet": Preset("resnet", 128, 20000, 0.08, True, 0.1, 32.0, 50000, weight_decay=2e-3,
                     decay_all_trainables=True, max_to_keep=1, accuracy_every=200, log_style="resnet"),
    # inception/imagenet_inception_bsp.py:59-72,123-142
    "inception": Preset("inception", 32, 1000000, 0.045, True, 0.94, 2.0, 1281167, optimizer="rmsprop",
                        save_secs=600.0, wipe_train_dir=False, nan_guard=True, log_style="resnet",
                        decay_steps_divides_by_workers=True),
}


# Any other network of the model zoo trains with the constants of the zoo's own trainer script
# (vgg/nets/mobilenet_v1_train.py:36-54,133-146): plain SGD, LR 0.045 x0.94 every 2.5 epochs of its (sic) 1,271,167
# images, batch 64, the per-variable L2 regularisers of the network's arg scope, no EMA, checkpoints every 100 s.
ZOO_PRESET = Preset("zoo", 64, 1000000, 0.045, False, 0.94, 2.5, 1271167, ema_decay=None, save_secs=100.0,
                    wipe_train_dir=False, nan_guard=True, log_style="resnet")


def preset_for(model: str, mode: str) -> Preset:
    """Mode-specific deviations the reference hard-codes: ASP/SSP do not scale LR by N; ASP-VGG applies a
    per-worker EMA(0.9999) every step (vgg/cifar10.py:384-388); Inception ASP/SSP create but never apply the
    EMA (imagenet_inception_asp.py:120-123)."""
    if model not in PRESETS:
        return replace(ZOO_PRESET, model=model)
    p = PRESETS[model]
    if mode in ("asp", "ssp"):
        p = replace(p, lr_times_workers_in_bsp=False)
        if model == "vgg":
            p = replace(p, ema_decay=0.9999)
        if model == "inception":
            p = replace(p, ema_decay=None, decay_steps_divides_by_workers=False)
    return p


def _tf_bool(v) -> bool:
    if isinstance(v, bool):
        return v
    return str(v).lower() in ("1", "true", "t", "yes", "y")


def build_parser(model: Optional[str] = None, mode: Optional[str] = None) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description="B200-native parameter-server CNN trainer", allow_abbrev=False)
    # ---- new switches -----------------------------------------------------------------------------
    ap.add_argument("--model", default=model,
                    help="one of %s, or any network of the model-zoo registry (models/registry.py)" % (MODELS,))
    ap.add_argument("--mode", choices=MODES, default=mode or "bsp")
    ap.add_argument("--staleness", type=int, default=5, help="SSP bound (reference literal 5, _ssp.py:81)")
    ap.add_argument("--synthetic", type=_tf_bool, nargs="?", const=True, default=None,
                    help="synthetic data of the model's shape (default: on when data_dir has no dataset)")
    ap.add_argument("--ref_compat", "--ref-compat", type=_tf_bool, nargs="?", const=True, default=False,
                    help="reproduce reference quirks (CifarNet double softmax Q4, ResNet batch-stat eval Q3)")
    ap.add_argument("--resume", type=_tf_bool, nargs="?", const=True, default=False)
    ap.add_argument("--ckpt_format", default="npz", choices=("npz", "tf"),
                    help="checkpoint container: NumPy .npz, or TensorFlow's V2 bundle (.index + .data-00000-of-00001)")
    ap.add_argument("--init_checkpoint", default="",
                    help="fine-tuning: initialise the variables from this checkpoint (directory or model.ckpt-N prefix), "
                         "ignoring missing ones, and start at global step 0 (vgg/nets/mobilenet_v1_train.py:165-181)")
    ap.add_argument("--trace", default="", help="write a Chrome-trace JSON of step phases to this path (device-timed "
                                                 "with CUDA events on a GPU)")
    ap.add_argument("--summarize_activations", type=_tf_bool, nargs="?", const=True, default=False,
                    help="chief: with every summary also write the reference's activation histograms / sparsity "
                         "(one extra forward pass), variable and gradient histograms (cnn/cifar10.py:104-121,376-382)")
    ap.add_argument("--ps_timeout_secs", type=float, default=None,
                    help="--backend symm: give up a parameter-server flag wait (a peer died) after this many seconds "
                         "and abort the run (default 300, or $DTF_PS_TIMEOUT_S)")
    ap.add_argument("--ssp_timeout_secs", type=float, default=None,
                    help="--backend symm: give up an SSP staleness wait after this many seconds (default 30, or "
                         "$DTF_SSP_TIMEOUT_S)")
    ap.add_argument("--precision", default="tf32", choices=("tf32", "tf32x3"),
                    help="tensor-core multiplies: tf32 (default) or tf32x3 = split operands, three accumulating passes, "
                         "fp32-like products for parity runs (the reference is fp32 throughout)")
    ap.add_argument("--backend", default="auto", choices=("auto", "local", "thread", "gloo", "nccl", "symm"),
                    help="transport: symm = in-kernel NVLink push/pull (product path); nccl = baseline")
    ap.add_argument("--placement", default="flat", choices=("flat", "round_robin", "axis0"))
    ap.add_argument("--kernels", default="auto", choices=("auto", "native", "torch"))
    ap.add_argument("--cuda_graph", type=_tf_bool, nargs="?", const=True, default=True)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--log_every", type=int, default=1)
    ap.add_argument("--metrics_jsonl", default="")
    # ---- reference flags: cluster -------------------------------------------------------------------
    ap.add_argument("--job_name", default="worker", help="'ps' or 'worker'")
    ap.add_argument("--task_id", type=int, default=0)
    ap.add_argument("--ps_hosts", default="")
    ap.add_argument("--worker_hosts", default="")
    ap.add_argument("--protocol", default="grpc", help="accepted for compatibility; transport is --backend")
    # ---- reference flags: training ------------------------------------------------------------------
    ap.add_argument("--batch_size", type=int, default=None)
    ap.add_argument("--data_dir", default="/tmp/cifar10_data")
    ap.add_argument("--train_dir", default="/tmp/dtf_train")
    ap.add_argument("--max_steps", type=int, default=None)
    ap.add_argument("--log_device_placement", type=_tf_bool, nargs="?", const=True, default=False)
    ap.add_argument("--resnet_size", type=int, default=32)
    ap.add_argument("--subset", default="train")
    ap.add_argument("--save_interval_secs", type=float, default=None)
    ap.add_argument("--save_summaries_secs", type=float, default=180)
    ap.add_argument("--initial_learning_rate", type=float, default=None)
    ap.add_argument("--num_epochs_per_decay", type=float, default=None)
    ap.add_argument("--learning_rate_decay_factor", type=float, default=None)
    # ---- reference flags: input pipeline (image_processing.py:47-71) -----------------------------------
    ap.add_argument("--image_size", type=int, default=None,
                    help="input resolution (default: the model's own: 299 for inception, default_image_size for zoo nets)")
    ap.add_argument("--num_classes", type=int, default=None, help="zoo networks only (default 1001)")
    ap.add_argument("--weight_decay", type=float, default=None,
                    help="zoo networks only: L2 weight of the network's arg scope (default: the scope's own)")
    ap.add_argument("--quantize", type=_tf_bool, nargs="?", const=True, default=False,
                    help="quantisation-aware training / evaluation: 8-bit fake-quantised weights and activations with "
                         "folded batch norm (vgg/nets/mobilenet_v1_train.py:40,135-139, mobilenet_v1_eval.py:38,126)")
    ap.add_argument("--quant_delay", type=int, default=0,
                    help="global step at which fake quantisation switches on (the ranges are tracked from step 0)")
    ap.add_argument("--input_pipeline", default="host", choices=("host", "device"),
                    help="CIFAR-10: 'host' distorts on CPU threads (native pass per image) and copies float batches; "
                         "'device' keeps the uint8 records in HBM and distorts each batch with one kernel")
    ap.add_argument("--num_preprocess_threads", type=int, default=4)
    ap.add_argument("--num_readers", type=int, default=4)
    ap.add_argument("--input_queue_memory_factor", type=int, default=16)
    # ---- reference flags: eval (cnn/cifar10_eval.py:49-60, inception_eval.py:40-55) ----------------------
    ap.add_argument("--eval_dir", default="/tmp/dtf_eval")
    ap.add_argument("--eval_data", default="test")
    ap.add_argument("--checkpoint_dir", default="")
    ap.add_argument("--eval_interval_secs", type=float, default=None)
    ap.add_argument("--num_examples", type=int, default=None)
    ap.add_argument("--run_once", type=_tf_bool, nargs="?", const=True, default=False)
    return ap


def parse_hosts(s: str) -> List[str]:
    return [h for h in s.split(",") if h]


def parse_args(argv=None, model: Optional[str] = None, mode: Optional[str] = None) -> argparse.Namespace:
    args = build_parser(model, mode).parse_args(argv)
    if args.model is None:
        raise SystemExit("--model is required")
    if args.model not in MODELS:
        from .models import registry
        if not registry.is_zoo_model(args.model):
            raise SystemExit("--model: unknown network %r" % args.model)
    p = preset_for(args.model, args.mode)
    if args.batch_size is None:
        args.batch_size = p.batch_size
    if args.max_steps is None:
        args.max_steps = p.max_steps
    if args.initial_learning_rate is None:
        args.initial_learning_rate = p.initial_lr
    if args.num_epochs_per_decay is None:
        args.num_epochs_per_decay = p.epochs_per_decay
    if args.learning_rate_decay_factor is None:
        args.learning_rate_decay_factor = p.decay_factor
    if args.save_interval_secs is None:
        args.save_interval_secs = p.save_secs
    args.preset = p
    args.ps_host_list = parse_hosts(args.ps_hosts)
    args.worker_host_list = parse_hosts(args.worker_hosts)
    return args
