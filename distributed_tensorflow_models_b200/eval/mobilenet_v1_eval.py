"""MobileNet-v1 evaluation of the model zoo.  Counterpart of the reference's ``vgg/nets/mobilenet_v1_eval.py``:
one pass of ``ceil(num_examples / batch_size)`` batches over the validation split with the newest checkpoint of
``--checkpoint_dir``, reporting accuracy and recall@5 (``:76-88,115-126``).

    python -m distributed_tensorflow_models_b200.eval.mobilenet_v1_eval --checkpoint_dir=/tmp/mobilenet_v1 \\
        --dataset_dir=/data/imagenet --depth_multiplier=0.5 --image_size=160
"""
import sys

from ..trainers.mobilenet_v1_train import translate
from .evaluator import main as _main


def main(argv=None):
    argv = sys.argv[1:] if argv is None else list(argv)
    rest, net, _ = translate(argv, for_eval=True)
    if "--run_once" not in rest:
        rest.append("--run_once")
    return _main(rest, model=net)


if __name__ == "__main__":
    sys.exit(main())
