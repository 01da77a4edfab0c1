#!/bin/bash
# ILSVRC-2012 tarballs -> raw tree  <out_dir>/{train/<synset>/*.JPEG, validation/*.JPEG, bounding_boxes/<synset>/*.xml}
# (role of the reference's inception/data/download_imagenet.sh).
#   bash download_imagenet.sh <out_dir> [synsets.txt]
# The three official archives — ILSVRC2012_img_train.tar, ILSVRC2012_img_val.tar, ILSVRC2012_bbox_train_v2.tar.gz —
# are taken from <out_dir> when they are there.  Missing ones are downloaded only if IMAGENET_USERNAME and
# IMAGENET_ACCESS_KEY are set in the environment (an image-net.org account is needed; credentials never belong in a
# script or a repository).  Without a synsets file every class archive inside the training tarball is unpacked.
set -e
OUT="${1%/}"; SYNSETS="$2"
if [ -z "$OUT" ]; then echo "usage: $0 <out_dir> [synsets.txt]"; exit 1; fi
mkdir -p "$OUT/train" "$OUT/validation" "$OUT/bounding_boxes"
BASE_URL="${IMAGENET_BASE_URL:-http://www.image-net.org/challenges/LSVRC/2012/nnoupb}"
fetch() {   # archive name
  if [ -f "$OUT/$1" ]; then return 0; fi
  if [ -z "$IMAGENET_USERNAME" ] || [ -z "$IMAGENET_ACCESS_KEY" ]; then
    echo "$1 is not in $OUT and IMAGENET_USERNAME / IMAGENET_ACCESS_KEY are not set: download it from image-net.org"
    echo "with your own account and place it there."; exit 2
  fi
  curl -fL --retry 3 -u "$IMAGENET_USERNAME:$IMAGENET_ACCESS_KEY" -o "$OUT/$1.part" "$BASE_URL/$1" && mv "$OUT/$1.part" "$OUT/$1"
}
fetch ILSVRC2012_bbox_train_v2.tar.gz
tar xzf "$OUT/ILSVRC2012_bbox_train_v2.tar.gz" -C "$OUT/bounding_boxes"
echo "bounding-box annotations: $(find "$OUT/bounding_boxes" -name '*.xml' | wc -l) files"
fetch ILSVRC2012_img_val.tar
tar xf "$OUT/ILSVRC2012_img_val.tar" -C "$OUT/validation"
fetch ILSVRC2012_img_train.tar
if [ -n "$SYNSETS" ] && [ -f "$SYNSETS" ]; then LIST=$(cat "$SYNSETS"); else LIST=$(tar tf "$OUT/ILSVRC2012_img_train.tar" | sed 's/\.tar$//'); fi
for SYNSET in $LIST; do
  mkdir -p "$OUT/train/$SYNSET"
  tar xf "$OUT/ILSVRC2012_img_train.tar" -O "$SYNSET.tar" | tar xf - -C "$OUT/train/$SYNSET"
done
echo "training images: $(find "$OUT/train" -name '*.JPEG' | wc -l); validation images: $(find "$OUT/validation" -name '*.JPEG' | wc -l)"
