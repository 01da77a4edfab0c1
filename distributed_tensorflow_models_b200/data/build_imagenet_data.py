"""ImageNet (ILSVRC-2012) -> TFRecords.  Role of ``inception/data/build_imagenet_data.py``:

* input tree ``<dir>/n01440764/n01440764_10026.JPEG`` (validation images first moved into synset folders by
  ``preprocess_imagenet_validation_data``);
* labels 1..1000 in the order of the synsets file, 0 reserved for background (:506-521);
* human-readable text from the metadata file (``n01440764<TAB>tench, Tinca tinca``);
* bounding boxes joined from the CSV produced by ``process_bounding_boxes`` (:635-680);
* output ``train-%05d-of-01024`` / ``validation-%05d-of-00128`` (:105-108,375).

    python -m distributed_tensorflow_models_b200.data.build_imagenet_data --train_directory=... \\
        --validation_directory=... --output_directory=... --labels_file=synsets.txt \\
        --imagenet_metadata_file=metadata.txt --bounding_box_file=boxes.csv
"""
from __future__ import annotations

import argparse
import os
import random
import sys
from typing import Dict, List

from .build_image_data import process_image_files


def build_synset_lookup(metadata_file: str) -> Dict[str, str]:
    out = {}
    if metadata_file and os.path.exists(metadata_file):
        for line in open(metadata_file):
            parts = line.rstrip("\n").split("\t")
            if len(parts) >= 2:
                out[parts[0]] = parts[1]
    return out


def build_bounding_box_lookup(csv_file: str) -> Dict[str, List[List[float]]]:
    """``<file>.JPEG,xmin,ymin,xmax,ymax`` (normalised) -> file -> list of boxes."""
    out: Dict[str, List[List[float]]] = {}
    if csv_file and os.path.exists(csv_file):
        for line in open(csv_file):
            parts = line.strip().split(",")
            if len(parts) == 5:
                out.setdefault(parts[0], []).append([float(v) for v in parts[1:]])
    return out


def find_image_files(data_dir: str, synsets: List[str]):
    filenames, syn, labels = [], [], []
    for idx, s in enumerate(synsets, start=1):
        d = os.path.join(data_dir, s)
        if not os.path.isdir(d):
            continue
        for fn in sorted(os.listdir(d)):
            if fn.upper().endswith((".JPEG", ".JPG", ".PNG")):
                filenames.append(os.path.join(d, fn))
                syn.append(s)
                labels.append(idx)
    order = list(range(len(filenames)))
    random.Random(12345).shuffle(order)
    return [filenames[i] for i in order], [syn[i] for i in order], [labels[i] for i in order]


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--train_directory", default="/tmp/")
    ap.add_argument("--validation_directory", default="/tmp/")
    ap.add_argument("--output_directory", default="/tmp/")
    ap.add_argument("--train_shards", type=int, default=1024)
    ap.add_argument("--validation_shards", type=int, default=128)
    ap.add_argument("--num_threads", type=int, default=8)
    ap.add_argument("--labels_file", default="imagenet_lsvrc_2015_synsets.txt")
    ap.add_argument("--imagenet_metadata_file", default="imagenet_metadata.txt")
    ap.add_argument("--bounding_box_file", default="./imagenet_2012_bounding_boxes.csv")
    a = ap.parse_args(argv)
    if os.path.exists(a.labels_file):
        synsets = [l.strip() for l in open(a.labels_file) if l.strip()]
    else:
        # the ILSVRC synsets file is the sorted list of the 1000 class directories: without it, read them off the
        # training directory (labels then come out the same, 1..1000 in that order)
        synsets = sorted(d for d in os.listdir(a.train_directory)
                         if os.path.isdir(os.path.join(a.train_directory, d)))
        if not synsets:
            raise SystemExit("--labels_file %r not found and %r has no class directories" %
                             (a.labels_file, a.train_directory))
        print("labels file %r not found: %d classes taken from the directories of %s" %
              (a.labels_file, len(synsets), a.train_directory))
    humans = build_synset_lookup(a.imagenet_metadata_file)
    boxes = build_bounding_box_lookup(a.bounding_box_file)
    for name, d, shards in (("validation", a.validation_directory, a.validation_shards),
                            ("train", a.train_directory, a.train_shards)):
        files, syn, labels = find_image_files(d, synsets)
        texts = [humans.get(s, s) for s in syn]
        process_image_files(name, files, texts, labels, shards, a.output_directory, a.num_threads, boxes, syn)
        print("%s: wrote %d images in up to %d shards" % (name, len(files), shards))
    return 0


if __name__ == "__main__":
    sys.exit(main())
