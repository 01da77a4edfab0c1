"""vgg / ASP trainer entry point.  Counterpart of the reference's ``vgg/cifar10_vgg_asp.py``.

    python -m distributed_tensorflow_models_b200.trainers.cifar10_vgg_asp --job_name=worker --task_id=0 \\
        --ps_hosts=h1:22222 --worker_hosts=h2:22223,h3:22223

All behaviour lives in ``trainers/common.py``; hyper-parameters come from ``config.PRESETS['vgg']``.
"""
import sys

from .common import main as _main


def main(argv=None):
    return _main(argv, model="vgg", mode="asp")


if __name__ == "__main__":
    sys.exit(main())
