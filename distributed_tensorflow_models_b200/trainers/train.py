"""Generic entry point: ``python -m distributed_tensorflow_models_b200.trainers.train --model resnet --mode bsp ...``"""
import sys

from .common import main

if __name__ == "__main__":
    sys.exit(main())
