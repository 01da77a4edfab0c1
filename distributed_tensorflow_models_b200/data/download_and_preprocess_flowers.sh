#!/bin/bash
# Flowers example data set -> TFRecords: fetch flower_photos.tgz when it is not in <data_dir> yet, unpack it, hold
# out validation images per label and shard (role of the reference's
# inception/data/download_and_preprocess_flowers.sh and its _mac twin: only POSIX tools + curl are used, so one
# script serves both).
#   bash download_and_preprocess_flowers.sh <data_dir> [validation images per label = 100]
# Without network access put flower_photos.tgz (or an unpacked flower_photos/ directory) into <data_dir> first.
set -e
DATA_DIR="${1%/}"; HOLDOUT="${2:-100}"
if [ -z "$DATA_DIR" ]; then echo "usage: $0 <data_dir> [validation images per label]"; exit 1; fi
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
DATA_URL="${FLOWERS_URL:-http://download.tensorflow.org/example_images/flower_photos.tgz}"
mkdir -p "$DATA_DIR"
if [ ! -d "$DATA_DIR/flower_photos" ]; then
  if [ ! -f "$DATA_DIR/flower_photos.tgz" ]; then
    echo "Downloading the flowers data set from $DATA_URL"
    curl -fL --retry 3 -o "$DATA_DIR/flower_photos.tgz.part" "$DATA_URL" && mv "$DATA_DIR/flower_photos.tgz.part" "$DATA_DIR/flower_photos.tgz" \
      || { rm -f "$DATA_DIR/flower_photos.tgz.part"; echo "download failed: place flower_photos.tgz into $DATA_DIR and run again"; exit 2; }
  else
    echo "Skipping download of the flowers data set."
  fi
  tar xzf "$DATA_DIR/flower_photos.tgz" -C "$DATA_DIR"
fi
bash "$HERE/preprocess_flowers.sh" "$DATA_DIR/flower_photos" "$DATA_DIR" "$HOLDOUT"
