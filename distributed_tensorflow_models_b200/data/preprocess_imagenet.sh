#!/bin/bash
# Raw ImageNet tree -> TFRecords (role of the reference's download_and_preprocess_imagenet.sh, minus the
# download: this box has no network, and credentials never belong in a script).
#   bash preprocess_imagenet.sh <raw_dir with train/ validation/ bounding_boxes/> <output_dir> <synsets.txt> <metadata.txt> <val_labels.txt>
set -e
RAW="$1"; OUT="$2"; SYNSETS="$3"; META="$4"; VAL_LABELS="$5"
if [ -z "$OUT" ]; then echo "usage: $0 <raw_dir> <output_dir> <synsets.txt> <metadata.txt> <validation_labels.txt>"; exit 1; fi
PKG=distributed_tensorflow_models_b200.data
python -m $PKG.preprocess_imagenet_validation_data "$RAW/validation" "$VAL_LABELS"
python -m $PKG.process_bounding_boxes "$RAW/bounding_boxes" "$SYNSETS" > "$OUT/imagenet_2012_bounding_boxes.csv"
python -m $PKG.build_imagenet_data --train_directory="$RAW/train" --validation_directory="$RAW/validation" \
  --output_directory="$OUT" --labels_file="$SYNSETS" --imagenet_metadata_file="$META" \
  --bounding_box_file="$OUT/imagenet_2012_bounding_boxes.csv"
