"""cifarnet evaluation entry point.  Counterpart of the reference's ``cifarnet/cifar10_cifarnet_eval.py``.

    python -m distributed_tensorflow_models_b200.eval.cifar10_cifarnet_eval --checkpoint_dir=/tmp/dtf_train --run_once
"""
import sys

from .evaluator import main as _main


def main(argv=None):
    return _main(argv, model="cifarnet")


if __name__ == "__main__":
    sys.exit(main())
