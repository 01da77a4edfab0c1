"""Host-side throughput of the input pipelines (no GPU needed): CIFAR-10 batches per second with the torch formulas and
with the native per-image pass, and the ImageNet per-image stages with and without the native passes.

    python scripts/bench_input_pipeline.py > profiles/input_pipeline_host_v1.jsonl

One JSON line per measurement.  Numbers depend on the box (this was written on an 8-core container)."""
import io
import json
import os
import random
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_tensorflow_models_b200.data import cifar10, imagenet, synthetic  # noqa: E402


def _time(fn, n):
    fn(0)
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    return (time.perf_counter() - t0) / n


def cifar(tmp):
    synthetic.write_cifar10_binary(tmp, train_per_file=10000, test=100)
    for size in (24, 32):
        for native in (False, True):
            for threads in (1, 4):
                it = cifar10.Cifar10Batches(tmp, 128, size, native=native, num_threads=threads, prefetch=8)
                for _ in range(20):
                    next(it)
                n = 300
                t0 = time.perf_counter()
                for _ in range(n):
                    next(it)
                dt = time.perf_counter() - t0
                it.close()
                print(json.dumps({"pipeline": "cifar10", "image_size": size, "batch": 128,
                                  "per_image_pass": "native" if native else "torch", "producer_threads": threads,
                                  "img_per_s": round(n * 128 / dt)}), flush=True)


def imagenet_stages():
    torch.set_num_threads(1)
    rng = np.random.default_rng(0)
    yy, xx = np.mgrid[0:375, 0:500]
    base = np.stack([(np.sin(xx / 40) + 1) * 100 + 30, (np.cos(yy / 30) + 1) * 90 + 40, (xx + yy) % 255], -1)
    arr = np.clip(base + rng.normal(0, 12, base.shape), 0, 255).astype(np.uint8)
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(arr).save(buf, format="JPEG", quality=90)
    jpg = buf.getvalue()
    bbox = np.zeros((0, 4), dtype=np.float32)
    img = imagenet.decode_jpeg_u8(jpg)
    real = cifar10._native_lib
    for native in (False, True):
        cifar10._native_lib = real if native else (lambda: None)
        tag = "native" if native else "numpy/torch"
        crop = imagenet._to_float(img[20:320, 50:400])
        res = {"pipeline": "imagenet", "image": "500x375 JPEG q90 -> 299x299", "per_image_passes": tag,
               "decode_ms": round(_time(lambda i: imagenet.decode_jpeg_u8(jpg), 40) * 1e3, 2)}
        for m, name in enumerate(("bilinear", "nearest", "bicubic", "area")):
            if native:
                fn = lambda i, m=m: imagenet.crop_resize_native(img, 20, 50, 300, 350, 299, 299, m, False)
            else:
                fn = lambda i, m=m: imagenet.resize(imagenet._to_float(img[20:320, 50:400]), 299, 299, m)
            res["resize_%s_ms" % name] = round(_time(fn, 30) * 1e3, 2)
        f = imagenet.resize(crop, 299, 299, 0)
        res["colour_ms"] = round(_time(lambda i: imagenet.distort_color(f, i, random.Random(i)), 40) * 1e3, 2)
        res["train_image_ms"] = round(_time(
            lambda i: imagenet.image_preprocessing(jpg, bbox, True, i, random.Random(i)), 60) * 1e3, 2)
        res["eval_image_ms"] = round(_time(
            lambda i: imagenet.image_preprocessing(jpg, bbox, False, i, random.Random(i)), 60) * 1e3, 2)
        print(json.dumps(res), flush=True)
    cifar10._native_lib = real


if __name__ == "__main__":
    print(json.dumps({"host": {"cpus": os.cpu_count(), "torch": torch.__version__}}), flush=True)
    with tempfile.TemporaryDirectory() as tmp:
        cifar(tmp)
    imagenet_stages()
