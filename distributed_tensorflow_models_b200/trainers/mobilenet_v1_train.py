"""MobileNet-v1 trainer of the model zoo.  Counterpart of the reference's ``vgg/nets/mobilenet_v1_train.py``.

    python -m distributed_tensorflow_models_b200.trainers.mobilenet_v1_train --dataset_dir=/data/imagenet \\
        --checkpoint_dir=/tmp/mobilenet_v1 --depth_multiplier=0.5 --image_size=160 --number_of_steps=1000

The reference script's flags (``:31-54``) are accepted under their own names and mapped onto the common trainer:
``--number_of_steps`` -> ``--max_steps``, ``--checkpoint_dir`` -> ``--train_dir``, ``--dataset_dir`` -> ``--data_dir``,
``--task`` -> ``--task_id``, ``--log_every_n_steps`` -> ``--log_every``; ``--depth_multiplier`` picks the registry
network (1.0 / 0.75 / 0.5 / 0.25); ``--fine_tune_checkpoint`` restores that checkpoint's variables, resets the global
step and drops the learning rate to 1e-4 (``:59-63,165-181``).  Optimiser, learning-rate schedule and loss are the
script's own constants (``config.ZOO_PRESET``).  ``--quantize`` switches on quantisation-aware training
(``slim/quantize.py``: 8-bit fake-quantised weights and activations, batch norm folded into the weights) with the
script's delay: none when fine-tuning, 250000 steps when training from scratch (``:66-74,135-139``);
``--quant_delay`` overrides it.
"""
import argparse
import sys

from .common import main as _main

_NETS = {1.0: "mobilenet_v1", 0.75: "mobilenet_v1_075", 0.5: "mobilenet_v1_050", 0.25: "mobilenet_v1_025"}


def translate(argv, for_eval: bool = False):
    """-> (argv for the common trainer / evaluator, registry network name, fine_tune_checkpoint)"""
    ap = argparse.ArgumentParser(add_help=False, allow_abbrev=False)
    ap.add_argument("--master", default="")
    ap.add_argument("--task", type=int, default=0)
    ap.add_argument("--ps_tasks", type=int, default=0)
    ap.add_argument("--number_of_steps", type=int, default=None)
    ap.add_argument("--depth_multiplier", type=float, default=1.0)
    ap.add_argument("--quantize", nargs="?", const="True", default="False")
    ap.add_argument("--fine_tune_checkpoint", default="")
    ap.add_argument("--checkpoint_dir", default="")
    ap.add_argument("--dataset_dir", default="")
    ap.add_argument("--log_every_n_steps", type=int, default=100)
    own, rest = ap.parse_known_args(argv)
    quantize = str(own.quantize).lower() in ("1", "true", "t", "yes", "")
    if own.depth_multiplier not in _NETS:
        raise SystemExit("--depth_multiplier must be one of %s" % sorted(_NETS))
    out = list(rest)
    if not for_eval:
        out += ["--task_id", str(own.task), "--log_every", str(own.log_every_n_steps)]
        if own.number_of_steps is not None:
            out += ["--max_steps", str(own.number_of_steps)]
    if own.checkpoint_dir:
        out += ["--checkpoint_dir" if for_eval else "--train_dir", own.checkpoint_dir]
    if own.dataset_dir:
        out += ["--data_dir", own.dataset_dir]
    if "--image_size" not in " ".join(rest):
        out += ["--image_size", "224"]
    if quantize:
        out += ["--quantize", "True"]
        if not for_eval and "--quant_delay" not in " ".join(rest):
            out += ["--quant_delay", "0" if own.fine_tune_checkpoint else "250000"]
    if own.fine_tune_checkpoint and "--initial_learning_rate" not in " ".join(rest):
        out += ["--initial_learning_rate", "1e-4"]
    return out, _NETS[own.depth_multiplier], own.fine_tune_checkpoint


def main(argv=None):
    argv = sys.argv[1:] if argv is None else list(argv)
    rest, net, fine_tune = translate(argv)
    if fine_tune:
        rest += ["--init_checkpoint", fine_tune]
    return _main(rest, model=net, mode="bsp")


if __name__ == "__main__":
    sys.exit(main())
