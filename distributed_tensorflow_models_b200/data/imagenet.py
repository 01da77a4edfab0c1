"""ImageNet input pipeline (role of ``inception/image_processing.py``).

* ``parse_example_proto`` (:339-407): fields ``image/encoded``, ``image/class/label``, ``image/class/text``,
  ``image/object/bbox/{xmin,ymin,xmax,ymax}`` -> (jpeg bytes, label, bbox [n,4] as ymin,xmin,ymax,xmax, text)
* training distortion (:198-276): ``sample_distorted_bounding_box(min_object_covered 0.1, aspect [0.75,1.33],
  area [0.05,1], 100 attempts)`` -> crop -> resize with method ``thread_id % 4`` -> random flip -> colour
  distortion in one of two orders by ``thread_id % 2`` (brightness 32/255, saturation [0.5,1.5], hue 0.2,
  contrast [0.5,1.5]) -> clip to [0,1]
* eval (:279-301): central crop 87.5 % + bilinear resize
* final scaling to [-1, 1] (:333-335)
* batching (:410-516): shards are dealt to ``num_preprocess_threads`` worker threads (the reference's readers +
  per-thread preprocessing graphs + batch_join); each fills a shuffle buffer and emits examples into a common
  bounded queue from which whole batches are assembled in pinned memory.  Ranks read disjoint shard subsets.

Images are decoded with Pillow (CMYK / PNG inputs are converted to RGB on decode, which covers the special
cases ``build_imagenet_data.py:264-322`` fixes offline).
"""
from __future__ import annotations

import io
import math
import queue
import random
import threading
from typing import List, Optional, Tuple

import numpy as np
import torch

from .dataset import Dataset, ImagenetData  # noqa: F401  (re-exported)
from .tfrecord import decode_example, read_records

BATCH_SIZE = 32
IMAGE_SIZE = 299
NUM_PREPROCESS_THREADS = 4
NUM_READERS = 4
INPUT_QUEUE_MEMORY_FACTOR = 16


# ---- example parsing ----------------------------------------------------------------------------------------
def parse_example_proto(record: bytes):
    f = decode_example(record)
    image = f["image/encoded"][0]
    label = int(f.get("image/class/label", [0])[0])
    text = f.get("image/class/text", [b""])[0]
    xmin, ymin = f.get("image/object/bbox/xmin", []), f.get("image/object/bbox/ymin", [])
    xmax, ymax = f.get("image/object/bbox/xmax", []), f.get("image/object/bbox/ymax", [])
    bbox = np.asarray([ymin, xmin, ymax, xmax], dtype=np.float32).T.reshape(-1, 4)
    return image, label, bbox, text


def decode_jpeg(buf: bytes) -> torch.Tensor:
    """-> float32 HWC in [0, 1] (tf.image.decode_jpeg + convert_image_dtype)."""
    from PIL import Image
    img = Image.open(io.BytesIO(buf)).convert("RGB")
    return torch.from_numpy(np.asarray(img, dtype=np.uint8).copy()).to(torch.float32) / 255.0


# ---- geometry ---------------------------------------------------------------------------------------------------
def sample_distorted_bounding_box(h: int, w: int, bboxes: np.ndarray, rng: random.Random,
                                  min_object_covered: float = 0.1, aspect_ratio_range=(0.75, 1.33),
                                  area_range=(0.05, 1.0), max_attempts: int = 100) -> Tuple[int, int, int, int]:
    """-> (y0, x0, crop_h, crop_w).  Falls back to the whole image like TF when no sample qualifies."""
    if bboxes.size == 0:
        bboxes = np.asarray([[0.0, 0.0, 1.0, 1.0]], dtype=np.float32)
    boxes = bboxes * np.asarray([h, w, h, w], dtype=np.float32)
    for _ in range(max_attempts):
        area = rng.uniform(*area_range) * h * w
        ratio = rng.uniform(*aspect_ratio_range)
        cw = int(round(math.sqrt(area * ratio)))
        ch = int(round(math.sqrt(area / ratio)))
        if cw < 1 or ch < 1 or cw > w or ch > h:
            continue
        y0 = rng.randint(0, h - ch)
        x0 = rng.randint(0, w - cw)
        # fraction of some object covered by the crop
        iy0 = np.maximum(boxes[:, 0], y0)
        ix0 = np.maximum(boxes[:, 1], x0)
        iy1 = np.minimum(boxes[:, 2], y0 + ch)
        ix1 = np.minimum(boxes[:, 3], x0 + cw)
        inter = np.clip(iy1 - iy0, 0, None) * np.clip(ix1 - ix0, 0, None)
        barea = np.clip((boxes[:, 2] - boxes[:, 0]) * (boxes[:, 3] - boxes[:, 1]), 1e-6, None)
        if np.any(inter / barea >= min_object_covered):
            return y0, x0, ch, cw
    return 0, 0, h, w


_RESIZE = ("bilinear", "nearest", "bicubic", "area")


def resize(img: torch.Tensor, height: int, width: int, method: int = 0) -> torch.Tensor:
    """HWC float -> HWC float; method index as in tf.image.ResizeMethod (0 bilinear, 1 nearest, 2 bicubic, 3 area)."""
    x = img.permute(2, 0, 1).unsqueeze(0)
    mode = _RESIZE[method % 4]
    if mode in ("bilinear", "bicubic"):
        y = torch.nn.functional.interpolate(x, size=(height, width), mode=mode, align_corners=False)
    else:
        y = torch.nn.functional.interpolate(x, size=(height, width), mode=mode)
    return y.squeeze(0).permute(1, 2, 0)


# ---- colour -----------------------------------------------------------------------------------------------------
def rgb_to_hsv(x: torch.Tensor) -> torch.Tensor:
    r, g, b = x[..., 0], x[..., 1], x[..., 2]
    mx, _ = x.max(-1)
    mn, _ = x.min(-1)
    d = mx - mn
    s = torch.where(mx > 0, d / mx.clamp_min(1e-12), torch.zeros_like(mx))
    dz = d.clamp_min(1e-12)
    h = torch.where(mx == r, ((g - b) / dz) % 6, torch.where(mx == g, (b - r) / dz + 2, (r - g) / dz + 4)) / 6.0
    h = torch.where(d > 0, h, torch.zeros_like(h))
    return torch.stack([h % 1.0, s, mx], -1)


def hsv_to_rgb(x: torch.Tensor) -> torch.Tensor:
    h, s, v = x[..., 0], x[..., 1], x[..., 2]
    i = torch.floor(h * 6)
    f = h * 6 - i
    p, q, t = v * (1 - s), v * (1 - f * s), v * (1 - (1 - f) * s)
    i = i.to(torch.int64) % 6
    r = torch.stack([v, q, p, p, t, v], -1).gather(-1, i.unsqueeze(-1)).squeeze(-1)
    g = torch.stack([t, v, v, q, p, p], -1).gather(-1, i.unsqueeze(-1)).squeeze(-1)
    b = torch.stack([p, p, t, v, v, q], -1).gather(-1, i.unsqueeze(-1)).squeeze(-1)
    return torch.stack([r, g, b], -1)


def distort_color(img: torch.Tensor, thread_id: int, rng: random.Random) -> torch.Tensor:
    """Two orderings of brightness / saturation / hue / contrast by thread parity (image_processing.py:164-195)."""
    def brightness(x):
        return x + rng.uniform(-32.0 / 255.0, 32.0 / 255.0)

    def saturation(x):
        hsv = rgb_to_hsv(x.clamp(0, 1))
        hsv[..., 1] = (hsv[..., 1] * rng.uniform(0.5, 1.5)).clamp(0, 1)
        return hsv_to_rgb(hsv)

    def hue(x):
        hsv = rgb_to_hsv(x.clamp(0, 1))
        hsv[..., 0] = (hsv[..., 0] + rng.uniform(-0.2, 0.2)) % 1.0
        return hsv_to_rgb(hsv)

    def contrast(x):
        m = x.mean(dim=(0, 1), keepdim=True)
        return (x - m) * rng.uniform(0.5, 1.5) + m

    order = (brightness, saturation, hue, contrast) if thread_id % 2 == 0 else (brightness, contrast, saturation, hue)
    for fn in order:
        img = fn(img)
    return img.clamp(0.0, 1.0)


def distort_image(img: torch.Tensor, height: int, width: int, bbox: np.ndarray, thread_id: int,
                  rng: random.Random) -> torch.Tensor:
    h, w = img.shape[0], img.shape[1]
    y0, x0, ch, cw = sample_distorted_bounding_box(h, w, bbox, rng)
    img = img[y0:y0 + ch, x0:x0 + cw]
    img = resize(img, height, width, thread_id % 4)
    if rng.random() < 0.5:
        img = img.flip(1)
    return distort_color(img, thread_id, rng)


def eval_image(img: torch.Tensor, height: int, width: int) -> torch.Tensor:
    h, w = img.shape[0], img.shape[1]
    ch, cw = int(h * 0.875), int(w * 0.875)
    y0, x0 = (h - ch) // 2, (w - cw) // 2
    return resize(img[y0:y0 + ch, x0:x0 + cw], height, width, 0)


def image_preprocessing(image_buffer: bytes, bbox: np.ndarray, train: bool, thread_id: int, rng: random.Random,
                        image_size: int = IMAGE_SIZE) -> torch.Tensor:
    img = decode_jpeg(image_buffer)
    img = distort_image(img, image_size, image_size, bbox, thread_id, rng) if train else eval_image(img, image_size,
                                                                                                    image_size)
    return (img - 0.5) * 2.0


# ---- batching ---------------------------------------------------------------------------------------------------
class ImagenetBatches:
    def __init__(self, dataset: Dataset, batch_size: int, train: bool, image_size: int = IMAGE_SIZE,
                 num_preprocess_threads: int = NUM_PREPROCESS_THREADS, rank: int = 0, world: int = 1, seed: int = 0,
                 shuffle_buffer: int = 256, repeat: bool = True):
        files = dataset.data_files()
        if not files:
            raise ValueError(dataset.download_message())
        files = files[rank::world] if len(files) >= world else files
        self.bs, self.size, self.train, self.repeat = batch_size, image_size, train, repeat
        self.q: "queue.Queue" = queue.Queue(maxsize=max(4 * batch_size, 64))
        self.stop = False
        self.n_threads = max(1, min(num_preprocess_threads, len(files)))
        self.done = 0
        self.lock = threading.Lock()
        self.threads = []
        for t in range(self.n_threads):
            th = threading.Thread(target=self._worker, args=(t, files[t::self.n_threads], seed + 1000 * rank + t,
                                                             shuffle_buffer), daemon=True)
            th.start()
            self.threads.append(th)
        self.pin = torch.cuda.is_available()

    def _emit(self, item) -> bool:
        while not self.stop:
            try:
                self.q.put(item, timeout=0.1)
                return True
            except queue.Full:
                continue
        return False

    def _worker(self, tid: int, files: List[str], seed: int, shuffle_buffer: int):
        rng = random.Random(seed)
        buf: list = []
        try:
            while not self.stop:
                order = list(files)
                if self.train:
                    rng.shuffle(order)
                for path in order:
                    for rec in read_records(path):
                        if self.stop:
                            return
                        if self.train:
                            buf.append(rec)
                            if len(buf) < shuffle_buffer:
                                continue
                            rec = buf.pop(rng.randrange(len(buf)))
                        image, label, bbox, _ = parse_example_proto(rec)
                        x = image_preprocessing(image, bbox, self.train, tid, rng, self.size)
                        if not self._emit((x, label)):
                            return
                if not self.repeat:
                    break
            while buf and not self.stop:
                image, label, bbox, _ = parse_example_proto(buf.pop())
                if not self._emit((image_preprocessing(image, bbox, self.train, tid, rng, self.size), label)):
                    return
        finally:
            with self.lock:
                self.done += 1
                if self.done == self.n_threads:
                    self._emit(None)

    def __iter__(self):
        return self

    def __next__(self):
        if getattr(self, "_exhausted", False):
            raise StopIteration
        xs, ys = [], []
        while len(xs) < self.bs:
            item = self.q.get()
            if item is None:
                self._exhausted = True
                if not xs:
                    raise StopIteration
                break
            xs.append(item[0])
            ys.append(item[1])
        x = torch.stack(xs)
        y = torch.tensor(ys, dtype=torch.int64)
        if self.pin:
            x, y = x.pin_memory(), y.pin_memory()
        return x, y

    def close(self):
        self.stop = True


def distorted_inputs(dataset: Dataset, batch_size: int = BATCH_SIZE, image_size: int = IMAGE_SIZE,
                     num_preprocess_threads: int = NUM_PREPROCESS_THREADS, **kw) -> ImagenetBatches:
    """Training batches (image_processing.py:107-137)."""
    return ImagenetBatches(dataset, batch_size, True, image_size, num_preprocess_threads, **kw)


def inputs(dataset: Dataset, batch_size: int = BATCH_SIZE, image_size: int = IMAGE_SIZE,
           num_preprocess_threads: int = NUM_PREPROCESS_THREADS, **kw) -> ImagenetBatches:
    """Evaluation batches (image_processing.py:74-104)."""
    return ImagenetBatches(dataset, batch_size, False, image_size, num_preprocess_threads, **kw)
