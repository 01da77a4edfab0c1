"""ImageNet bounding-box XML -> normalised CSV (role of ``inception/data/process_bounding_boxes.py:119-166``).

Each annotation file ``<synset>/<synset>_<id>.xml`` holds the image size and one or more ``bndbox`` entries;
output lines are ``<synset>_<id>.JPEG,xmin,ymin,xmax,ymax`` with coordinates clamped to [0, 1] and ordered so that
min <= max.  Optionally restricted to the synsets of the challenge list; boxes whose label is not the file's own
synset (or a child in a user-supplied hierarchy) are skipped, like the reference does.

    python -m distributed_tensorflow_models_b200.data.process_bounding_boxes <xml_dir> [synsets.txt] > boxes.csv
"""
from __future__ import annotations

import glob
import os
import sys
import xml.etree.ElementTree as ET
from typing import Iterator, List, Optional, Set, Tuple


def parse_xml(path: str) -> List[Tuple[str, float, float, float, float]]:
    root = ET.parse(path).getroot()
    size = root.find("size")
    w, h = float(size.find("width").text), float(size.find("height").text)
    out = []
    for obj in root.iter("object"):
        label = obj.find("name").text
        bb = obj.find("bndbox")
        xmin, ymin = float(bb.find("xmin").text) / w, float(bb.find("ymin").text) / h
        xmax, ymax = float(bb.find("xmax").text) / w, float(bb.find("ymax").text) / h
        xmin, xmax = min(xmin, xmax), max(xmin, xmax)
        ymin, ymax = min(ymin, ymax), max(ymin, ymax)
        clamp = lambda v: min(max(v, 0.0), 1.0)  # noqa: E731
        out.append((label, clamp(xmin), clamp(ymin), clamp(xmax), clamp(ymax)))
    return out


def process(xml_dir: str, synsets: Optional[Set[str]] = None) -> Iterator[str]:
    for path in sorted(glob.glob(os.path.join(xml_dir, "*", "*.xml"))):
        syn = os.path.basename(os.path.dirname(path))
        if synsets is not None and syn not in synsets:
            continue
        image = os.path.splitext(os.path.basename(path))[0] + ".JPEG"
        for label, xmin, ymin, xmax, ymax in parse_xml(path):
            if label != syn:
                continue                     # box of a different object class in the same image
            if xmax <= xmin or ymax <= ymin:
                continue
            yield "%s,%.4f,%.4f,%.4f,%.4f" % (image, xmin, ymin, xmax, ymax)


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    if not argv:
        print(__doc__)
        return 1
    syn = set(l.strip() for l in open(argv[1])) if len(argv) > 1 else None
    for line in process(argv[0], syn):
        print(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
