"""vgg evaluation entry point.  The reference has no VGG evaluator; this one follows the CifarNet pattern.

    python -m distributed_tensorflow_models_b200.eval.cifar10_vgg_eval --checkpoint_dir=/tmp/dtf_train --run_once
"""
import sys

from .evaluator import main as _main


def main(argv=None):
    return _main(argv, model="vgg")


if __name__ == "__main__":
    sys.exit(main())
