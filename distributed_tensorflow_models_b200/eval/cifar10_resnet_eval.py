"""resnet evaluation entry point.  Counterpart of the reference's ``resnet/cifar10_resnet_eval.py``.

    python -m distributed_tensorflow_models_b200.eval.cifar10_resnet_eval --checkpoint_dir=/tmp/dtf_train --run_once
"""
import sys

from .evaluator import main as _main


def main(argv=None):
    return _main(argv, model="resnet")


if __name__ == "__main__":
    sys.exit(main())
