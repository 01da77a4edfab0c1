"""List or convert checkpoints (role of TensorFlow's ``inspect_checkpoint`` tool for the files this framework and
``tf.train.Saver`` write).

    python -m distributed_tensorflow_models_b200.ckpt.inspect /tmp/train                   # newest checkpoint of a dir
    python -m distributed_tensorflow_models_b200.ckpt.inspect /tmp/train/model.ckpt-300 --tensor conv1/biases
    python -m distributed_tensorflow_models_b200.ckpt.inspect /tmp/train/model.ckpt-300 --convert tf --out /tmp/x/model.ckpt-300
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from . import store, tf_bundle


def resolve(path: str) -> str:
    if os.path.isdir(path):
        latest = store.latest_checkpoint(path)
        if not latest:
            raise SystemExit("no checkpoint state file in %s" % path)
        return latest
    for suffix in (".npz", ".index"):
        if path.endswith(suffix):
            path = path[: -len(suffix)]
    if not store.exists(path):
        raise SystemExit("no checkpoint at %s" % path)
    return path


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("path", help="checkpoint prefix (model.ckpt-N) or a directory with a `checkpoint` state file")
    ap.add_argument("--tensor", default="", help="print this tensor's values")
    ap.add_argument("--convert", choices=("npz", "tf"), default="", help="rewrite the checkpoint in this container")
    ap.add_argument("--out", default="", help="output prefix for --convert (default: next to the input)")
    args = ap.parse_args(argv)
    prefix = resolve(args.path)
    tensors = store.load(prefix)
    kind = "tf bundle" if (tf_bundle.is_bundle(prefix) and not os.path.exists(prefix + ".npz")) else "npz"
    if args.tensor:
        if args.tensor not in tensors:
            raise SystemExit("%s is not in %s" % (args.tensor, prefix))
        t = tensors[args.tensor]
        print("%s %s %s" % (args.tensor, t.dtype, tuple(t.shape)))
        with np.printoptions(threshold=64, precision=6):
            print(t)
        return 0
    if args.convert:
        out = args.out or prefix
        os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
        if args.convert == "tf":
            tf_bundle.write_bundle(out, tensors)
        else:
            np.savez(out + ".npz", **{k.replace("/", "|"): v for k, v in tensors.items()})
        print("wrote %s (%s)" % (out, args.convert))
        return 0
    total = 0
    print("%s  [%s, global step %d]" % (prefix, kind, store.step_of(prefix)))
    for name in sorted(tensors):
        t = tensors[name]
        total += t.size
        print("  %-72s %-8s %s" % (name, t.dtype, tuple(t.shape)))
    print("%d tensors, %d values" % (len(tensors), total))
    return 0


if __name__ == "__main__":
    try:
        sys.exit(main())
    except BrokenPipeError:          # `... | head`
        sys.stderr.close()
        sys.exit(0)
