#!/bin/bash
# Folder-per-label photo collection -> sharded TFRecords (role of the reference's
# inception/data/download_and_preprocess_flowers.sh, minus the download: this box has no network, so the
# unpacked collection — e.g. flower_photos/<label>/*.jpg — is the input).
#   bash preprocess_flowers.sh <photos_dir with one sub-directory per label> <data_dir> [images per label held out]
# Produces <data_dir>/raw-data/{train,validation,labels.txt} and <data_dir>/{train,validation}-?????-of-0000?.
set -e
SRC="${1%/}"; DATA_DIR="${2%/}"; HOLDOUT="${3:-100}"
if [ -z "$DATA_DIR" ]; then echo "usage: $0 <photos_dir> <data_dir> [validation images per label = 100]"; exit 1; fi
SCRATCH="$DATA_DIR/raw-data"
TRAIN="$SCRATCH/train"; VAL="$SCRATCH/validation"; LABELS="$SCRATCH/labels.txt"
rm -rf "$TRAIN" "$VAL"; mkdir -p "$TRAIN" "$VAL"
# labels = the sub-directory names, sorted (files such as LICENSE.txt at the top level are not labels)
find "$SRC" -mindepth 1 -maxdepth 1 -type d -printf '%f\n' | sort > "$LABELS"
while read -r LABEL; do
  mkdir -p "$TRAIN/$LABEL" "$VAL/$LABEL"
  cp "$SRC/$LABEL"/* "$TRAIN/$LABEL/"
  # a random but reproducible hold-out: the first $HOLDOUT files in shuffled order go to validation
  ls -1 "$TRAIN/$LABEL" | shuf --random-source=<(yes) | head -n "$HOLDOUT" | while read -r F; do
    mv "$TRAIN/$LABEL/$F" "$VAL/$LABEL/"
  done
done < "$LABELS"
python -m distributed_tensorflow_models_b200.data.build_image_data --train_directory="$TRAIN" \
  --validation_directory="$VAL" --output_directory="$DATA_DIR" --labels_file="$LABELS"
