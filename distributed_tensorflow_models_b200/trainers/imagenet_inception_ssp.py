"""inception / SSP trainer entry point.  Counterpart of the reference's ``inception/imagenet_inception_ssp.py``.

    python -m distributed_tensorflow_models_b200.trainers.imagenet_inception_ssp --job_name=worker --task_id=0 \\
        --ps_hosts=h1:22222 --worker_hosts=h2:22223,h3:22223

All behaviour lives in ``trainers/common.py``; hyper-parameters come from ``config.PRESETS['inception']``.
"""
import sys

from .common import main as _main


def main(argv=None):
    return _main(argv, model="inception", mode="ssp")


if __name__ == "__main__":
    sys.exit(main())
