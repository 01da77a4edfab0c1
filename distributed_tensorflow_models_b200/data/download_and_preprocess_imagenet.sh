#!/bin/bash
# ImageNet from the official archives to sharded TFRecords (role of the reference's
# inception/data/download_and_preprocess_imagenet.sh): unpack (download_imagenet.sh), move the validation images
# into per-synset directories, turn the bounding-box XML into one CSV, write train-?????-of-01024 /
# validation-?????-of-00128.
#   bash download_and_preprocess_imagenet.sh <data_dir> <synsets.txt> <metadata.txt> <validation_labels.txt>
# The three text files are the ones the ILSVRC-2012 development kit / the TensorFlow models repository provide
# (1000 synset ids in label order; synset id <TAB> human-readable names; one synset id per validation image).
set -e
DATA_DIR="${1%/}"; SYNSETS="$2"; META="$3"; VAL_LABELS="$4"
if [ -z "$VAL_LABELS" ]; then echo "usage: $0 <data_dir> <synsets.txt> <metadata.txt> <validation_labels.txt>"; exit 1; fi
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
bash "$HERE/download_imagenet.sh" "$DATA_DIR/raw-data" "$SYNSETS"
bash "$HERE/preprocess_imagenet.sh" "$DATA_DIR/raw-data" "$DATA_DIR" "$SYNSETS" "$META" "$VAL_LABELS"
