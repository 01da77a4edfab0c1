"""Move the flat ImageNet validation images into per-synset folders (role of
``inception/data/preprocess_imagenet_validation_data.py``): line i of the labels file is the synset of
``ILSVRC2012_val_%08d.JPEG`` (i starting at 1).

    python -m distributed_tensorflow_models_b200.data.preprocess_imagenet_validation_data <val_dir> <labels.txt>
"""
from __future__ import annotations

import os
import sys


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 2:
        print(__doc__)
        return 1
    data_dir, labels_file = argv[0], argv[1]
    labels = [l.strip() for l in open(labels_file) if l.strip()]
    for s in sorted(set(labels)):
        os.makedirs(os.path.join(data_dir, s), exist_ok=True)
    moved = 0
    for i, s in enumerate(labels, start=1):
        name = "ILSVRC2012_val_%08d.JPEG" % i
        src = os.path.join(data_dir, name)
        if os.path.exists(src):
            os.rename(src, os.path.join(data_dir, s, name))
            moved += 1
    print("moved %d validation images" % moved)
    return 0


if __name__ == "__main__":
    sys.exit(main())
