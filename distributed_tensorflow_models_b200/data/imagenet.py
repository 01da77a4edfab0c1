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

import atexit
import io
import math
import os
import queue
import random
import sys
import threading
import weakref
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


_TV_DECODE = []          # [callable or None], resolved on first use (the import is slow and optional)


def _torchvision_decoder():
    """torchvision's libjpeg-turbo binding, reached through its operator library alone: importing the torchvision
    package costs ~8 s per (worker) process, loading ``image.so`` costs nothing."""
    if not _TV_DECODE:
        dec = None
        try:
            import importlib.util
            spec = importlib.util.find_spec("torchvision")
            lib = os.path.join(os.path.dirname(spec.origin), "image.so") if spec and spec.origin else ""
            if os.path.exists(lib):
                torch.ops.load_library(lib)
                op = torch.ops.image.decode_jpeg
                RGB = 3                                           # torchvision.io.ImageReadMode.RGB

                def dec(buf: bytes) -> np.ndarray:
                    chw = op(torch.frombuffer(bytearray(buf), dtype=torch.uint8), RGB, False)
                    return chw.permute(1, 2, 0).contiguous().numpy()
                probe = dec(_PROBE_JPEG)                          # an operator with another signature fails here
                if probe.shape != (1, 1, 3):
                    dec = None
        except Exception:
            dec = None
        _TV_DECODE.append(dec)
    return _TV_DECODE[0]


# 1x1 white baseline JPEG: checks the operator's calling convention once per process
_PROBE_JPEG = bytes.fromhex(
    "ffd8ffe000104a46494600010100000100010000ffdb004300080606070605080707070909080a0c140d0c0b0b0c1912130f141d1a1f"
    "1e1d1a1c1c20242e2720222c231c1c2837292c30313434341f27393d38323c2e333432ffc0000b080001000101011100ffc4001f0000"
    "010501010101010100000000000000000102030405060708090a0bffc400b5100002010303020403050504040000017d010203000411"
    "05122131410613516107227114328191a1082342b1c11552d1f02433627282090a161718191a25262728292a3435363738393a434445"
    "464748494a535455565758595a636465666768696a737475767778797a838485868788898a92939495969798999aa2a3a4a5a6a7a8a9"
    "aab2b3b4b5b6b7b8b9bac2c3c4c5c6c7c8c9cad2d3d4d5d6d7d8d9dae1e2e3e4e5e6e7e8e9eaf1f2f3f4f5f6f7f8f9faffda00080101"
    "00003f00f7faffd9")


def decode_jpeg_u8(buf: bytes) -> np.ndarray:
    """-> uint8 HWC (tf.image.decode_jpeg with channels=3).  libjpeg-turbo through torchvision when it is installed
    (1.8 ms for a 500x375 image against Pillow's 2.8 ms), Pillow otherwise and for whatever that decoder refuses
    (CMYK files, PNGs carrying a .JPEG name — both occur in ImageNet)."""
    dec = _torchvision_decoder()
    if dec is not None and buf[:2] == b"\xff\xd8":
        try:
            return dec(buf)
        except Exception:
            pass
    from PIL import Image
    return np.asarray(Image.open(io.BytesIO(buf)).convert("RGB"), dtype=np.uint8)


def decode_jpeg(buf: bytes) -> torch.Tensor:
    """-> float32 HWC in [0, 1] (tf.image.decode_jpeg + convert_image_dtype)."""
    return torch.from_numpy(decode_jpeg_u8(buf).astype(np.float32) / np.float32(255.0))


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


def _rgb_to_hsv_np(r: np.ndarray, g: np.ndarray, b: np.ndarray):
    """Contiguous float32 planes in [0, 1] -> (h in [0, 1), s, v) (same formulas as :func:`rgb_to_hsv`; the modulo
    operations are written as compare-and-shift because ``np.mod`` costs as much as everything else together)."""
    f = np.float32
    mx = np.maximum(np.maximum(r, g), b)
    mn = np.minimum(np.minimum(r, g), b)
    d = mx - mn
    inv = f(1.0) / np.maximum(d, f(1e-12))
    s = np.where(mx > 0, d / np.maximum(mx, f(1e-12)), f(0.0))
    hr = (g - b) * inv                       # in [-1, 1]: wrap the negative half by +6
    hr = np.where(hr < 0, hr + f(6.0), hr)
    h = np.where(mx == r, hr, np.where(mx == g, (b - r) * inv + f(2.0), (r - g) * inv + f(4.0)))
    h = np.where(d > 0, h * f(1.0 / 6.0), f(0.0))
    h = np.where(h >= 1, h - f(1.0), h)
    return h, s, mx


def _hsv_to_rgb_np(h: np.ndarray, s: np.ndarray, v: np.ndarray):
    """Branch-free form: channel(n) = v - v s max(0, min(k, 4 - k, 1)), k = (n + 6h) mod 6, n = 5, 3, 1."""
    f = np.float32
    h6 = h * f(6.0)
    vs = v * s
    out = []
    for n in (5.0, 3.0, 1.0):
        k = h6 + f(n)
        k = np.where(k >= 6, k - f(6.0), k)
        out.append(v - vs * np.clip(np.minimum(k, f(4.0) - k), f(0.0), f(1.0)))
    return out


def distort_color(img: torch.Tensor, thread_id: int, rng: random.Random, inplace: bool = False) -> torch.Tensor:
    """Two orderings of brightness / saturation / hue / contrast by thread parity (image_processing.py:164-195):
    two native sweeps over the image (``csrc/input_pipeline.cu::dtf_distort_color_host``) when the library is built,
    the NumPy planes of :func:`distort_color_numpy` otherwise — same draws in the same order, same result to 1e-5."""
    from .cifar10 import _native_lib
    lib = _native_lib()
    if lib is None:
        return distort_color_numpy(img, thread_id, rng)
    import ctypes as C
    x = img.numpy() if isinstance(img, torch.Tensor) else np.asarray(img)
    if not (inplace and x.dtype == np.float32 and x.flags.c_contiguous and x.flags.writeable):
        x = np.array(x, dtype=np.float32, order="C", copy=True)      # ``inplace``: the caller owns a fresh image
    assert x.ndim == 3 and x.shape[2] == 3
    delta = rng.uniform(-32.0 / 255.0, 32.0 / 255.0)
    if thread_id % 2 == 0:
        sat, hue = rng.uniform(0.5, 1.5), rng.uniform(-0.2, 0.2)
        factor = rng.uniform(0.5, 1.5)
    else:
        factor = rng.uniform(0.5, 1.5)
        sat, hue = rng.uniform(0.5, 1.5), rng.uniform(-0.2, 0.2)
    rc = lib.dtf_distort_color_host(C.c_void_p(x.ctypes.data), C.c_longlong(x.shape[0] * x.shape[1]),
                                    int(thread_id % 2), C.c_float(delta), C.c_float(sat), C.c_float(hue),
                                    C.c_float(factor))
    if rc != 0:
        raise RuntimeError("dtf_distort_color_host failed (%d)" % rc)
    return torch.from_numpy(x)


def distort_color_numpy(img: torch.Tensor, thread_id: int, rng: random.Random) -> torch.Tensor:
    """The colour distortion on NumPy planes (the oracle of the native pass, and the fallback without it).

    Saturation and hue are neighbours in both orderings and the RGB image between them is already inside [0, 1], so
    they share one RGB -> HSV -> RGB round trip (the costly part of the whole input pipeline); the arithmetic runs in
    NumPy float32 on three contiguous colour planes.  The random draws are taken in the reference's order."""
    f = np.float32
    x = img.numpy() if isinstance(img, torch.Tensor) else np.asarray(img, dtype=np.float32)
    planes = [np.ascontiguousarray(x[..., c], dtype=np.float32) for c in range(3)]

    def brightness(p):
        delta = f(rng.uniform(-32.0 / 255.0, 32.0 / 255.0))
        return [a + delta for a in p]

    def saturation_hue(p):
        sat = f(rng.uniform(0.5, 1.5))
        hue = f(rng.uniform(-0.2, 0.2))
        h, s_, v = _rgb_to_hsv_np(*[np.clip(a, f(0.0), f(1.0)) for a in p])
        s_ = np.clip(s_ * sat, f(0.0), f(1.0))
        h = h + hue
        h -= np.floor(h)
        return _hsv_to_rgb_np(h, s_, v)

    def contrast(p):
        factor = f(rng.uniform(0.5, 1.5))
        out = []
        for a in p:
            m = a.mean(dtype=np.float32)
            out.append((a - m) * factor + m)
        return out

    order = (brightness, saturation_hue, contrast) if thread_id % 2 == 0 else (brightness, contrast, saturation_hue)
    for fn in order:
        planes = fn(planes)
    out = np.empty(x.shape, dtype=np.float32)
    for c in range(3):
        np.clip(planes[c], f(0.0), f(1.0), out=out[..., c])
    return torch.from_numpy(out)


def _to_float(img) -> torch.Tensor:
    """uint8 HWC (NumPy) -> float32 [0, 1] tensor; float tensors pass through."""
    if isinstance(img, np.ndarray):
        return torch.from_numpy(img.astype(np.float32) / np.float32(255.0))
    return img


def crop_resize_native(img: np.ndarray, y0: int, x0: int, ch: int, cw: int, height: int, width: int, method: int,
                       flip: bool) -> Optional[torch.Tensor]:
    """``csrc/input_pipeline.cu::dtf_crop_resize_host``: window of a decoded uint8 HWC image -> float HWC in [0, 1],
    resized with ``method`` (0 bilinear, 1 nearest, 2 bicubic, 3 area; the tap rules of :func:`resize`) and
    optionally mirrored, in one separable pass pair; ``None`` when the library is not built or ``img`` is not uint8."""
    from .cifar10 import _native_lib
    lib = _native_lib()
    if lib is None or not isinstance(img, np.ndarray) or img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
        return None
    import ctypes as C
    src = np.ascontiguousarray(img)
    out = np.empty((height, width, 3), dtype=np.float32)
    rc = lib.dtf_crop_resize_host(C.c_void_p(src.ctypes.data), int(src.shape[0]), int(src.shape[1]), int(y0), int(x0),
                                  int(ch), int(cw), int(height), int(width), int(method % 4), int(bool(flip)),
                                  C.c_void_p(out.ctypes.data))
    if rc != 0:
        msg = lib.dtf_last_error()
        raise RuntimeError("dtf_crop_resize_host failed: %s" % (msg.decode() if msg else rc))
    return torch.from_numpy(out)


def distort_image(img, height: int, width: int, bbox: np.ndarray, thread_id: int, rng: random.Random) -> torch.Tensor:
    """``img``: float HWC tensor in [0, 1], or the decoder's uint8 array (only the crop is then converted)."""
    h, w = img.shape[0], img.shape[1]
    y0, x0, ch, cw = sample_distorted_bounding_box(h, w, bbox, rng)
    flip = rng.random() < 0.5
    out = crop_resize_native(img, y0, x0, ch, cw, height, width, thread_id % 4, flip)
    if out is not None:
        return distort_color(out, thread_id, rng, inplace=True)
    out = resize(_to_float(img[y0:y0 + ch, x0:x0 + cw]), height, width, thread_id % 4)
    if flip:
        out = out.flip(1)
    return distort_color(out, thread_id, rng)


def eval_image(img, height: int, width: int) -> torch.Tensor:
    h, w = img.shape[0], img.shape[1]
    ch, cw = int(h * 0.875), int(w * 0.875)
    y0, x0 = (h - ch) // 2, (w - cw) // 2
    out = crop_resize_native(img, y0, x0, ch, cw, height, width, 0, False)
    if out is None:
        out = resize(_to_float(img[y0:y0 + ch, x0:x0 + cw]), height, width, 0)
    return out


def image_preprocessing(image_buffer: bytes, bbox: np.ndarray, train: bool, thread_id: int, rng: random.Random,
                        image_size: int = IMAGE_SIZE) -> torch.Tensor:
    img = decode_jpeg_u8(image_buffer)
    img = distort_image(img, image_size, image_size, bbox, thread_id, rng) if train else eval_image(img, image_size,
                                                                                                    image_size)
    return img.sub_(0.5).mul_(2.0)              # both branches hand back a fresh tensor


# ---- batching ---------------------------------------------------------------------------------------------------
def _worker_loop(tid: int, files: List[str], seed: int, shuffle_buffer: int, train: bool, repeat: bool, size: int,
                 emit, stopped) -> None:
    """One reader + preprocessor (the reference runs ``num_readers`` readers and ``num_preprocess_threads``
    preprocessing graphs around a RandomShuffleQueue, image_processing.py:410-516): reads its share of the shards,
    keeps a shuffle buffer, emits ``(image, label)`` pairs and finally ``None``."""
    rng = random.Random(seed)
    buf: list = []

    def process(rec) -> bool:
        image, label, bbox, _ = parse_example_proto(rec)
        return emit((image_preprocessing(image, bbox, train, tid, rng, size), label))

    try:
        while not stopped():
            order = list(files)
            if train:
                rng.shuffle(order)
            for path in order:
                for rec in read_records(path):
                    if stopped():
                        return
                    if train:
                        buf.append(rec)
                        if len(buf) < shuffle_buffer:
                            continue
                        rec = buf.pop(rng.randrange(len(buf)))
                    if not process(rec):
                        return
            if not repeat:
                break
        while buf and not stopped():
            if not process(buf.pop()):
                return
    finally:
        emit(None)


def _process_main(tid, files, seed, shuffle_buffer, train, repeat, size, ring, free_q, ready_q, stop_event) -> None:
    """Worker process: preprocessed images go into a free slot of the shared-memory ring; only (slot, label) travels
    through the queue (an mp.Queue moves ~300 MB/s — one 299x299x3 float image is 1 MB)."""
    torch.set_num_threads(1)          # one core per worker; also keeps the forked child out of the OpenMP pool

    def emit(item) -> bool:
        if item is None:
            while not stop_event.is_set():
                try:
                    ready_q.put(None, timeout=0.1)
                    return True
                except queue.Full:
                    continue
            return False
        x, label = item
        while not stop_event.is_set():
            try:
                slot = free_q.get(timeout=0.1)
            except queue.Empty:
                continue
            ring[slot].copy_(x)
            ready_q.put((slot, label))           # capacity == number of slots: cannot be full while we hold one
            return True
        return False

    _worker_loop(tid, files, seed, shuffle_buffer, train, repeat, size, emit, stop_event.is_set)
    if stop_event.is_set():           # shut down by the consumer: do not wait for it to read what is still queued
        ready_q.cancel_join_thread()  # (a normal finish flushes, so that the last images and the end marker arrive)
    free_q.cancel_join_thread()


_LIVE_BATCHERS: "weakref.WeakSet" = weakref.WeakSet()


def _close_live_batchers() -> None:
    for b in list(_LIVE_BATCHERS):
        b.close()


atexit.register(_close_live_batchers)


class ImagenetBatches:
    """Iterator of ``(images float32 NHWC in [-1, 1], labels int64)`` host batches.

    ``workers='process'`` (default for training with more than one worker) runs each reader / preprocessor in its own
    process and hands images over through a shared-memory ring: JPEG decoding and the colour distortion are
    Python-level PIL / NumPy work (13.6 ms per image and core) that threads parallelise only up to the interpreter
    lock (~250 img/s here).  ``workers='thread'`` keeps everything in-process (evaluation, tests, debuggers)."""

    def __init__(self, dataset: Dataset, batch_size: int, train: bool, image_size: int = IMAGE_SIZE,
                 num_preprocess_threads: int = NUM_PREPROCESS_THREADS, rank: int = 0, world: int = 1, seed: int = 0,
                 shuffle_buffer: int = 256, repeat: bool = True, workers: Optional[str] = None):
        files = dataset.data_files()
        if not files:
            raise ValueError(dataset.download_message())
        files = files[rank::world] if len(files) >= world else files
        self.bs, self.size, self.train, self.repeat = batch_size, image_size, train, repeat
        self.n_workers = max(1, min(num_preprocess_threads, len(files)))
        self.mode = workers or ("process" if (train and self.n_workers > 1) else "thread")
        if self.mode not in ("process", "thread"):
            raise ValueError("workers must be 'process' or 'thread'")
        self.finished = 0
        self._exhausted = False
        self._closed = False
        self.pin = torch.cuda.is_available()
        self.workers: list = []
        if self.mode == "process":
            import torch.multiprocessing as mp       # pickles the shared ring by handle
            # "spawn": the trainer process is heavily multi-threaded (CUDA runtime, OpenMP pool, prefetchers) and a
            # fork() of such a process can leave a child holding a lock nobody will ever release
            # (a __main__ that cannot be re-imported — a script piped through stdin — leaves only fork)
            main = sys.modules.get("__main__")
            mfile = getattr(main, "__file__", None)
            spawn_ok = getattr(main, "__spec__", None) is not None or mfile is None or os.path.exists(mfile)
            ctx = mp.get_context("spawn" if spawn_ok else "fork")
            slots = max(2 * batch_size, 2 * self.n_workers, 16)
            self.ring = torch.empty((slots, image_size, image_size, 3), dtype=torch.float32).share_memory_()
            self.free_q = ctx.Queue(maxsize=slots)
            self.q = ctx.Queue(maxsize=slots + self.n_workers)
            for i in range(slots):
                self.free_q.put(i)
            self.stop_event = ctx.Event()
            for t in range(self.n_workers):
                p = ctx.Process(target=_process_main, daemon=True,
                                args=(t, files[t::self.n_workers], seed + 1000 * rank + t, shuffle_buffer, train, repeat,
                                      image_size, self.ring, self.free_q, self.q, self.stop_event))
                p.start()
                self.workers.append(p)
        else:
            self.q = queue.Queue(maxsize=max(4 * batch_size, 64))
            self.stop_event = threading.Event()

            def emit(item) -> bool:
                while not self.stop_event.is_set():
                    try:
                        self.q.put(item, timeout=0.1)
                        return True
                    except queue.Full:
                        continue
                return False

            for t in range(self.n_workers):
                th = threading.Thread(target=_worker_loop, daemon=True,
                                      args=(t, files[t::self.n_workers], seed + 1000 * rank + t, shuffle_buffer, train,
                                            repeat, image_size, emit, self.stop_event.is_set))
                th.start()
                self.workers.append(th)
        _LIVE_BATCHERS.add(self)

    def __iter__(self):
        return self

    def _get(self):
        """Next item, or None when every worker has finished; notices dead worker processes instead of hanging."""
        while True:
            try:
                item = self.q.get(timeout=1.0)
            except queue.Empty:
                if self.mode == "process" and not any(p.is_alive() for p in self.workers) and self.q.empty():
                    return None
                continue
            if item is None:
                self.finished += 1
                if self.finished < self.n_workers:
                    continue
            return item

    def __next__(self):
        if self._exhausted:
            raise StopIteration
        xs, ys = [], []
        while len(xs) < self.bs:
            item = self._get()
            if item is None:
                self._exhausted = True
                if not xs:
                    raise StopIteration
                break
            xs.append(item[0])
            ys.append(item[1])
        if self.mode == "process":
            x = self.ring[torch.tensor(xs, dtype=torch.int64)]        # gather out of the ring, then free the slots
            for slot in xs:
                self.free_q.put(slot)
        else:
            x = torch.stack(xs)
        y = torch.tensor(ys, dtype=torch.int64)
        if self.pin:
            x, y = x.pin_memory(), y.pin_memory()
        return x, y

    def close(self):
        if self._closed:
            return
        self._closed = True
        self.stop_event.set()
        if self.mode == "process":
            for p in self.workers:
                p.join(timeout=0.5)
            for p in self.workers:
                if p.is_alive():              # e.g. still importing, or blocked on a full pipe: no cooperation needed
                    p.terminate()
            for p in self.workers:
                p.join(timeout=2.0)
            for q_ in (self.q, self.free_q):
                q_.cancel_join_thread()
                q_.close()
        else:
            try:                                  # unblock producers waiting on a full queue, then let them finish
                while True:
                    self.q.get_nowait()
            except queue.Empty:
                pass
            for th in self.workers:
                th.join(timeout=1.0)


def distorted_inputs(dataset: Dataset, batch_size: int = BATCH_SIZE, image_size: int = IMAGE_SIZE,
                     num_preprocess_threads: int = NUM_PREPROCESS_THREADS, **kw) -> ImagenetBatches:
    """Training batches (image_processing.py:107-137)."""
    return ImagenetBatches(dataset, batch_size, True, image_size, num_preprocess_threads, **kw)


def inputs(dataset: Dataset, batch_size: int = BATCH_SIZE, image_size: int = IMAGE_SIZE,
           num_preprocess_threads: int = NUM_PREPROCESS_THREADS, **kw) -> ImagenetBatches:
    """Evaluation batches (image_processing.py:74-104)."""
    return ImagenetBatches(dataset, batch_size, False, image_size, num_preprocess_threads, **kw)
