"""Convert a folder-per-label image tree into sharded TFRecords of ``Example`` protos.

Role of ``inception/data/build_image_data.py`` (and the shared machinery of ``build_imagenet_data.py``):

    <train_directory>/<label>/<image>.jpg ...      labels_file: one label (sub-directory name) per line

-> ``<output_directory>/train-00000-of-01024`` ... with the feature keys the reader expects
(``image/encoded``, ``image/class/label`` (1-based; 0 is the background class), ``image/class/text``,
``image/filename``, ``image/height|width|channels|colorspace|format``, optional ``image/object/bbox/*``).
Files are shuffled with the fixed seed 12345 (build_imagenet_data.py:526-528), PNG / CMYK inputs are
re-encoded as RGB JPEG (:264-322), shards are written by a pool of threads (:412-461).

    python -m distributed_tensorflow_models_b200.data.build_image_data --train_directory=... \\
        --validation_directory=... --output_directory=... --labels_file=labels.txt
"""
from __future__ import annotations

import argparse
import io
import os
import random
import sys
import threading
from typing import Dict, List, Optional, Sequence, Tuple

from .tfrecord import TFRecordWriter, encode_example


def _is_jpeg_rgb(data: bytes) -> bool:
    return data[:3] == b"\xff\xd8\xff"


def process_image(path: str) -> Tuple[bytes, int, int]:
    """-> (RGB JPEG bytes, height, width).  Non-JPEG and CMYK files are converted (reference :264-322)."""
    from PIL import Image
    with open(path, "rb") as f:
        data = f.read()
    img = Image.open(io.BytesIO(data))
    if not _is_jpeg_rgb(data) or img.mode != "RGB":
        buf = io.BytesIO()
        img.convert("RGB").save(buf, format="JPEG", quality=100)
        data = buf.getvalue()
        img = Image.open(io.BytesIO(data))
    w, h = img.size
    return data, h, w


def convert_to_example(filename: str, image_buffer: bytes, label: int, text: str, height: int, width: int,
                       bboxes: Optional[Sequence[Sequence[float]]] = None, synset: Optional[str] = None) -> bytes:
    feats: Dict[str, list] = {
        "image/height": [height], "image/width": [width], "image/colorspace": [b"RGB"], "image/channels": [3],
        "image/class/label": [label], "image/class/text": [text.encode()], "image/format": [b"JPEG"],
        "image/filename": [os.path.basename(filename).encode()], "image/encoded": [image_buffer],
    }
    if synset is not None:
        feats["image/class/synset"] = [synset.encode()]
    if bboxes:
        feats["image/object/bbox/xmin"] = [float(b[0]) for b in bboxes]
        feats["image/object/bbox/ymin"] = [float(b[1]) for b in bboxes]
        feats["image/object/bbox/xmax"] = [float(b[2]) for b in bboxes]
        feats["image/object/bbox/ymax"] = [float(b[3]) for b in bboxes]
        feats["image/object/bbox/label"] = [label] * len(bboxes)
    return encode_example(feats)


def find_image_files(data_dir: str, labels: List[str]):
    """-> (filenames, texts, labels) shuffled with seed 12345; label index starts at 1 (0 = background)."""
    filenames, texts, ids = [], [], []
    for idx, text in enumerate(labels, start=1):
        d = os.path.join(data_dir, text)
        if not os.path.isdir(d):
            continue
        for fn in sorted(os.listdir(d)):
            if fn.lower().endswith((".jpg", ".jpeg", ".png")):
                filenames.append(os.path.join(d, fn))
                texts.append(text)
                ids.append(idx)
    order = list(range(len(filenames)))
    random.Random(12345).shuffle(order)
    return [filenames[i] for i in order], [texts[i] for i in order], [ids[i] for i in order]


def process_image_files(name: str, filenames, texts, labels, num_shards: int, output_directory: str,
                        num_threads: int = 4, bbox_lookup=None, synsets=None) -> List[str]:
    os.makedirs(output_directory, exist_ok=True)
    num_shards = max(1, min(num_shards, max(1, len(filenames))))
    bounds = [round(i * len(filenames) / num_shards) for i in range(num_shards + 1)]
    outputs = [os.path.join(output_directory, "%s-%.5d-of-%.5d" % (name, s, num_shards)) for s in range(num_shards)]

    def work(tid: int):
        for s in range(tid, num_shards, num_threads):
            with TFRecordWriter(outputs[s]) as w:
                for i in range(bounds[s], bounds[s + 1]):
                    try:
                        buf, h, wd = process_image(filenames[i])
                    except Exception as e:  # noqa: BLE001
                        sys.stderr.write("SKIPPED: %s (%s)\n" % (filenames[i], e))
                        continue
                    boxes = bbox_lookup.get(os.path.basename(filenames[i])) if bbox_lookup else None
                    syn = synsets[i] if synsets else None
                    w.write(convert_to_example(filenames[i], buf, labels[i], texts[i], h, wd, boxes, syn))

    ts = [threading.Thread(target=work, args=(t,)) for t in range(max(1, min(num_threads, num_shards)))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return outputs


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--train_directory", default="/tmp/")
    ap.add_argument("--validation_directory", default="/tmp/")
    ap.add_argument("--output_directory", default="/tmp/")
    ap.add_argument("--train_shards", type=int, default=2)
    ap.add_argument("--validation_shards", type=int, default=2)
    ap.add_argument("--num_threads", type=int, default=2)
    ap.add_argument("--labels_file", default="")
    a = ap.parse_args(argv)
    labels = [l.strip() for l in open(a.labels_file) if l.strip()]
    for name, d, shards in (("validation", a.validation_directory, a.validation_shards),
                            ("train", a.train_directory, a.train_shards)):
        f, t, l = find_image_files(d, labels)
        process_image_files(name, f, t, l, shards, a.output_directory, a.num_threads)
        print("%s: wrote %d images" % (name, len(f)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
