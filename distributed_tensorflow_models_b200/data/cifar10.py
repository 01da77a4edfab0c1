"""CIFAR-10 input pipeline (role of the five copies of ``cifar10_input.py`` in the reference).

* record format: 1 label byte + 3072 image bytes stored CHW -> HWC uint8 (cnn/cifar10_input.py:68-95);
  files ``data_batch_{1..5}.bin`` / ``test_batch.bin`` under ``<data_dir>/cifar-10-batches-bin`` (:156-157,223)
* training distortion (:145-203): random crop to ``image_size`` (24 for the plain CNN, 32 = identity for the
  others), random left-right flip, ``random_brightness(max_delta=63)``, ``random_contrast(0.2, 1.8)``,
  ``per_image_standardization``
* eval (:206-257): centre crop / pad to ``image_size`` + standardisation
* batching: a background thread keeps a bounded queue of ready batches (the reference's 16 queue-runner
  threads and ``capacity = 0.4*50000 + 3*batch`` shuffle queue); ``shuffle`` is honoured (the reference forces
  ``shuffle=False`` in four of five copies, Q9) and every rank reads a disjoint stride of the records when
  ``world > 1`` (the reference has no sharding; documented deviation).

All augmentation is vectorised torch-on-CPU over the whole batch; output is NHWC float32 in pinned memory
when CUDA is present so the trainer's H2D copy is asynchronous.
"""
from __future__ import annotations

import atexit
import os
import queue
import threading
import weakref
from typing import Iterator, List, Optional, Tuple

import numpy as np
import torch

IMAGE_SIZE = 24                 # cnn/cifar10_input.py:30 (the other models use 32)
NUM_CLASSES = 10
NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 50000
NUM_EXAMPLES_PER_EPOCH_FOR_EVAL = 10000
RECORD_BYTES = 1 + 32 * 32 * 3


def data_files(data_dir: str, eval_data: bool) -> List[str]:
    d = os.path.join(data_dir, "cifar-10-batches-bin")
    names = ["test_batch.bin"] if eval_data else [f"data_batch_{i}.bin" for i in range(1, 6)]
    files = [os.path.join(d, n) for n in names]
    for f in files:
        if not os.path.exists(f):
            raise ValueError("Failed to find file: " + f)
    return files


def dataset_available(data_dir: str) -> bool:
    try:
        data_files(data_dir, False)
        return True
    except ValueError:
        return False


DATA_URL = "http://www.cs.toronto.edu/~kriz/cifar-10-binary.tar.gz"      # the reference's source (cnn/cifar10.py:101)


def maybe_download_and_extract(data_dir: str, url: str = DATA_URL, timeout: float = 20.0) -> bool:
    """Make ``<data_dir>/cifar-10-batches-bin`` exist the way the reference does (cnn/cifar10.py:394-411): if the
    binary files are missing, unpack ``<data_dir>/cifar-10-binary.tar.gz``, downloading it first when it is not there
    either.  Returns whether the dataset is present afterwards — on a machine without network access (this build
    box) the download fails quietly and the caller falls back to synthetic data.  ``DTF_NO_DOWNLOAD=1`` skips the
    attempt."""
    if dataset_available(data_dir):
        return True
    import tarfile
    import urllib.request
    tarball = os.path.join(data_dir, url.split("/")[-1])
    if not os.path.exists(tarball):
        if os.environ.get("DTF_NO_DOWNLOAD") == "1":
            return False
        tmp = tarball + ".part"
        try:
            os.makedirs(data_dir, exist_ok=True)
            with urllib.request.urlopen(url, timeout=timeout) as r, open(tmp, "wb") as f:
                while True:
                    block = r.read(1 << 20)
                    if not block:
                        break
                    f.write(block)
            os.replace(tmp, tarball)
        except Exception:                    # no network, refused, disk full ...: not fatal, synthetic data it is
            try:
                os.remove(tmp)
            except OSError:
                pass
            return False
    try:
        with tarfile.open(tarball, "r:gz") as tf:
            members = [m for m in tf.getmembers()
                       if m.isfile() and not os.path.isabs(m.name) and ".." not in m.name.split("/")]
            tf.extractall(data_dir, members=members, filter="data")
    except (tarfile.TarError, OSError):
        return False
    return dataset_available(data_dir)


def read_records(files: List[str]) -> Tuple[np.ndarray, np.ndarray]:
    """-> (labels int64 [N], images uint8 [N,32,32,3])"""
    raw = np.concatenate([np.fromfile(f, dtype=np.uint8) for f in files])
    if raw.size % RECORD_BYTES:
        raise ValueError("corrupt CIFAR-10 binary file (size is not a multiple of the record length)")
    rec = raw.reshape(-1, RECORD_BYTES)
    labels = rec[:, 0].astype(np.int64)
    images = rec[:, 1:].reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)
    return labels, np.ascontiguousarray(images)


def per_image_standardization(x: torch.Tensor) -> torch.Tensor:
    """(x - mean) / max(stddev, 1/sqrt(num_elements)) per image (tf.image.per_image_standardization)."""
    n = x[0].numel()
    flat = x.reshape(x.shape[0], -1)
    mean = flat.mean(1, keepdim=True)
    std = flat.std(1, unbiased=False, keepdim=True).clamp_min(1.0 / n ** 0.5)
    return ((flat - mean) / std).reshape(x.shape)


def distort_batch_reference(images_u8: torch.Tensor, image_size: int, gen: torch.Generator) -> torch.Tensor:
    """The training distortion written out step by step as the reference applies it (cnn/cifar10_input.py:176-190):
    random crop -> random flip -> random brightness (+-63) -> random contrast (0.2 .. 1.8) -> standardisation.
    Kept as the oracle of :func:`distort_batch`."""
    b = images_u8.shape[0]
    x = images_u8.to(torch.float32)
    if image_size < 32:
        oy = torch.randint(0, 32 - image_size + 1, (b,), generator=gen)
        ox = torch.randint(0, 32 - image_size + 1, (b,), generator=gen)
        idx_y = oy[:, None] + torch.arange(image_size)[None, :]
        idx_x = ox[:, None] + torch.arange(image_size)[None, :]
        x = x[torch.arange(b)[:, None, None], idx_y[:, :, None], idx_x[:, None, :]]
    flip = torch.rand(b, generator=gen) < 0.5
    x = torch.where(flip[:, None, None, None], x.flip(2), x)
    delta = (torch.rand(b, generator=gen) * 2 - 1) * 63.0
    x = x + delta[:, None, None, None]
    factor = 0.2 + torch.rand(b, generator=gen) * 1.6
    mean = x.mean(dim=(1, 2), keepdim=True)
    x = (x - mean) * factor[:, None, None, None] + mean
    return per_image_standardization(x)


def draw_distortion(b: int, image_size: int, gen: torch.Generator):
    """The random draws of one training batch, consumed in the order of :func:`distort_batch_reference`:
    -> (crop rows int64 [b], crop columns int64 [b], flip bool [b], contrast factor float32 [b]).  The brightness
    delta is drawn too (stream parity) and dropped: the standardisation cancels it."""
    if image_size < 32:
        oy = torch.randint(0, 32 - image_size + 1, (b,), generator=gen)
        ox = torch.randint(0, 32 - image_size + 1, (b,), generator=gen)
    else:
        oy = ox = torch.zeros(b, dtype=torch.int64)
    flip = torch.rand(b, generator=gen) < 0.5
    _delta = torch.rand(b, generator=gen)
    factor = 0.2 + torch.rand(b, generator=gen) * 1.6
    return oy, ox, flip, factor


def _native_lib():
    """The native library when it has been built (``None`` otherwise: the torch formulas then do the work)."""
    from .. import _native as nat
    try:
        return nat.lib() if nat.available() else None
    except OSError:
        return None


def distort_records_native(images_u8: torch.Tensor, idx: torch.Tensor, oy: torch.Tensor, ox: torch.Tensor,
                           flip: torch.Tensor, factor: torch.Tensor, size: int, threads: int = 1,
                           out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """``csrc/input_pipeline.cu::dtf_cifar_distort_host``: gather records ``idx`` of the resident uint8 set
    ``[N,32,32,3]``, crop / flip / contrast / standardise in one pass per image on ``threads`` host threads
    -> float32 ``[b,size,size,3]``; ``None`` when the library is not built."""
    import ctypes as C
    lib = _native_lib()
    if lib is None:
        return None
    b = idx.numel()
    if out is None:
        out = torch.empty(b, size, size, 3, dtype=torch.float32)
    idx = idx.to(torch.int64).contiguous()
    oy32, ox32 = oy.to(torch.int32).contiguous(), ox.to(torch.int32).contiguous()
    fl, fac = flip.to(torch.uint8).contiguous(), factor.to(torch.float32).contiguous()
    assert images_u8.dtype == torch.uint8 and images_u8.is_contiguous() and tuple(images_u8.shape[1:]) == (32, 32, 3)
    rc = lib.dtf_cifar_distort_host(C.c_void_p(images_u8.data_ptr()), C.c_longlong(images_u8.shape[0]),
                                    C.c_void_p(idx.data_ptr()), C.c_void_p(oy32.data_ptr()),
                                    C.c_void_p(ox32.data_ptr()), C.c_void_p(fl.data_ptr()), C.c_void_p(fac.data_ptr()),
                                    int(b), int(size), C.c_void_p(out.data_ptr()), int(threads))
    if rc != 0:
        msg = lib.dtf_last_error()
        raise RuntimeError("dtf_cifar_distort_host failed: %s" % (msg.decode() if msg else rc))
    return out


def distort_batch(images_u8: torch.Tensor, image_size: int, gen: torch.Generator) -> torch.Tensor:
    """uint8 [B,32,32,3] -> float32 [B,image_size,image_size,3] with the training distortions, in three passes.

    Crop and flip are one gather on the uint8 records.  Brightness (x + d) and contrast ((x - m_c) f + m_c, with the
    per-channel mean m_c) are followed by per-image standardisation, which removes any shift and rescales: with
    c = x - m_c per channel, the standardised result is (c f + m_c - M) / max(sd(c f + m_c), 1/sqrt(n)), M the
    mean of the channel means — so the distortions never need to be materialised, only their effect on the
    statistics.  The same random draws are consumed in the same order as in :func:`distort_batch_reference`, and the
    outputs agree to rounding (tests/test_ckpt_data_cli.py)."""
    b = images_u8.shape[0]
    size = image_size if image_size < 32 else 32
    oy, ox, flip, factor = draw_distortion(b, image_size, gen)
    ar = torch.arange(size)
    idx_y = oy[:, None] + ar[None, :]
    cols = torch.where(flip[:, None], (size - 1 - ar)[None, :], ar[None, :])
    idx_x = ox[:, None] + cols
    x = images_u8[torch.arange(b)[:, None, None], idx_y[:, :, None], idx_x[:, None, :]].to(torch.float32)
    n = size * size * 3
    m_c = x.mean(dim=(1, 2), keepdim=True)                         # per-channel mean (what adjust_contrast uses)
    c = x - m_c
    f = factor[:, None, None, None]
    shift = m_c - m_c.mean(dim=3, keepdim=True)                    # channel mean minus image mean
    y = c * f + shift                                              # == contrast-adjusted image minus its mean
    var = (y * y).reshape(b, -1).mean(1)
    inv = 1.0 / torch.sqrt(var).clamp_min(1.0 / n ** 0.5)
    return y * inv[:, None, None, None]


def center_batch(images_u8: torch.Tensor, image_size: int) -> torch.Tensor:
    """Eval preprocessing: resize_image_with_crop_or_pad + standardisation (cnn/cifar10_input.py:242-247)."""
    x = images_u8.to(torch.float32)
    if image_size < 32:
        o = (32 - image_size) // 2
        x = x[:, o:o + image_size, o:o + image_size]
    elif image_size > 32:
        p = image_size - 32
        x = torch.nn.functional.pad(x, (0, 0, p // 2, p - p // 2, p // 2, p - p // 2))
    return per_image_standardization(x)


_LIVE_PREFETCHERS: "weakref.WeakSet" = weakref.WeakSet()


def _shutdown_prefetchers():
    for p in list(_LIVE_PREFETCHERS):
        p.close(join=True)


atexit.register(_shutdown_prefetchers)


class _OrderedPrefetcher:
    """``threads`` producer threads; thread t builds batches t, t + T, t + 2T, ... (``make(k)`` is a pure function of
    the batch number) into its own queue and the consumer drains the queues round-robin, so the batch sequence is the
    same for any thread count — the role of the reference's 16 ``tf.train.batch`` threads (cnn/cifar10_input.py:118-135),
    without their run-to-run reordering."""

    def __init__(self, make, threads: int, depth: int):
        self.T = max(1, threads)
        self.qs = [queue.Queue(maxsize=max(1, depth)) for _ in range(self.T)]
        self.stop = False
        self.next = 0
        self.ts = [threading.Thread(target=self._run, args=(make, t), daemon=True) for t in range(self.T)]
        _LIVE_PREFETCHERS.add(self)
        for t in self.ts:
            t.start()

    def _run(self, make, t: int):
        k = t
        while not self.stop:
            item = make(k)
            while not self.stop:
                try:
                    self.qs[t].put(item, timeout=0.1)
                    break
                except queue.Full:
                    continue
            if item is None:
                return
            k += self.T

    def get(self):
        item = self.qs[self.next % self.T].get()
        self.next += 1
        return item

    def close(self, join: bool = False):
        self.stop = True
        if join:
            for q_, t in zip(self.qs, self.ts):
                if t.is_alive():
                    try:
                        while True:
                            q_.get_nowait()
                    except queue.Empty:
                        pass
                    t.join(timeout=2.0)


class Cifar10Batches:
    """Iterator of ``(images float32 NHWC, labels int64)`` host batches.

    Batch ``k`` is a pure function of ``(seed, rank, k)``: its epoch's permutation and its distortion draws come from
    generators seeded by those numbers, so the stream does not depend on ``num_threads`` and a run can be repeated."""

    def __init__(self, data_dir: str, batch_size: int, image_size: int = IMAGE_SIZE, train: bool = True,
                 shuffle: bool = True, rank: int = 0, world: int = 1, seed: int = 0, prefetch: int = 4,
                 repeat: bool = True, num_threads: Optional[int] = None, native: bool = True):
        labels, images = read_records(data_files(data_dir, eval_data=not train))
        if world > 1 and train:
            labels, images = np.ascontiguousarray(labels[rank::world]), np.ascontiguousarray(images[rank::world])
        self.labels = torch.from_numpy(labels)
        self.images = torch.from_numpy(images)
        self.bs, self.size, self.train, self.shuffle, self.repeat = batch_size, image_size, train, shuffle, repeat
        self.seed = seed + 7919 * rank
        self.n = self.labels.numel()
        self.bpe = max(1, -(-self.n // batch_size))            # batches per epoch
        self._orders: dict = {}
        self._lock = threading.Lock()
        self.pin = torch.cuda.is_available()
        self.native = bool(native) and _native_lib() is not None
        if num_threads is None:
            num_threads = min(4, max(1, (os.cpu_count() or 2) // 2)) if train else 1
        self.pf = _OrderedPrefetcher(self._make, num_threads, max(1, prefetch // max(1, num_threads)) + 1)

    def _order(self, epoch: int) -> torch.Tensor:
        with self._lock:
            o = self._orders.get(epoch)
            if o is None:
                if self.shuffle and self.train:
                    o = torch.randperm(self.n, generator=torch.Generator().manual_seed(self.seed * 1000003 + epoch))
                else:
                    o = torch.arange(self.n)
                self._orders[epoch] = o
                for old in [e for e in self._orders if e < epoch - 1]:
                    del self._orders[old]
            return o

    def _indices(self, epoch: int, j: int) -> torch.Tensor:
        order = self._order(epoch)
        idx = order[j * self.bs:(j + 1) * self.bs]
        if idx.numel() < self.bs and self.repeat:
            idx = torch.cat([idx, order[: self.bs - idx.numel()]])
        return idx

    def _make(self, k: int):
        epoch, j = divmod(k, self.bpe)
        if epoch > 0 and not self.repeat:
            return None
        idx = self._indices(epoch, j)
        y = self.labels[idx].clone()
        x = self._make_native(idx, k) if self.native else None
        if x is None:
            img = self.images[idx]
            if self.train:
                x = distort_batch(img, self.size, torch.Generator().manual_seed(self.seed * 7368787 + k))
            else:
                x = center_batch(img, self.size)
            if self.pin:
                x = x.pin_memory()
        if self.pin:
            y = y.pin_memory()
        return x, y

    def _draws(self, b: int, k: int):
        """Crop / flip / contrast draws of batch ``k`` (training) or the fixed central crop (evaluation)."""
        if self.train:
            return draw_distortion(b, self.size, torch.Generator().manual_seed(self.seed * 7368787 + k))
        o = torch.full((b,), (32 - self.size) // 2, dtype=torch.int64)
        return o, o, torch.zeros(b, dtype=torch.bool), torch.ones(b)

    def _make_native(self, idx: torch.Tensor, k: int):
        """One native pass from the resident records straight into the (pinned) output batch."""
        if self.size > 32:
            return None
        b = idx.numel()
        oy, ox, flip, factor = self._draws(b, k)
        out = torch.empty(b, self.size, self.size, 3, dtype=torch.float32, pin_memory=self.pin)
        return distort_records_native(self.images, idx, oy, ox, flip, factor, self.size, threads=1, out=out)

    def __iter__(self):
        return self

    def __next__(self):
        item = self.pf.get()
        if item is None:
            raise StopIteration
        return item

    def close(self):
        self.pf.close()


class RecordBatch:
    """A batch that does not exist yet: record numbers and distortion draws, packed into one pinned staging buffer
    (``idx int64[b] | crop row int32[b] | crop column int32[b] | contrast float32[b] | flip uint8[b]``).
    :meth:`copy_into` produces the float images in a destination buffer on the device that holds the records —
    one small upload and one kernel (``dtf_cifar_distort``) on the current stream.  ``DevicePrefetcher`` calls it in
    place of the host-to-device image copy."""
    dtype = torch.float32
    BYTES_PER_IMAGE = 8 + 4 + 4 + 4 + 1

    def __init__(self, owner: "DeviceCifar10Batches", idx, oy, ox, flip, factor):
        b = idx.numel()
        self.owner, self.b = owner, b
        self.shape = torch.Size((b, owner.size, owner.size, 3))
        st = torch.empty(b * self.BYTES_PER_IMAGE, dtype=torch.uint8, pin_memory=owner.pin)
        st[:8 * b].view(torch.int64).copy_(idx)
        st[8 * b:12 * b].view(torch.int32).copy_(oy)
        st[12 * b:16 * b].view(torch.int32).copy_(ox)
        st[16 * b:20 * b].view(torch.float32).copy_(factor)
        st[20 * b:].copy_(flip)
        self.stage = st

    def fields(self, st: torch.Tensor):
        b = self.b
        return (st[:8 * b].view(torch.int64), st[8 * b:12 * b].view(torch.int32), st[12 * b:16 * b].view(torch.int32),
                st[20 * b:], st[16 * b:20 * b].view(torch.float32))

    def copy_into(self, dst: torch.Tensor) -> None:
        own = self.owner
        assert dst.shape == self.shape and dst.dtype == torch.float32 and dst.is_contiguous()
        if dst.is_cuda:
            import ctypes as C
            from .. import _native as nat
            if own.images_dev is None or own.images_dev.device != dst.device:
                raise RuntimeError("the records live on %s but the batch is wanted on %s: build DeviceCifar10Batches "
                                   "with device=<the training device>" % (own.device, dst.device))
            st = self.stage.to(dst.device, non_blocking=True)
            base, b = st.data_ptr(), self.b
            vp = C.c_void_p
            nat.check(nat.lib().dtf_cifar_distort(vp(own.images_dev.data_ptr()), vp(base), vp(base + 8 * b),
                                                  vp(base + 12 * b), vp(base + 20 * b), vp(base + 16 * b), int(b),
                                                  int(own.size), vp(dst.data_ptr()), nat.stream_ptr()),
                      "dtf_cifar_distort")
        else:
            idx, oy, ox, flip, factor = self.fields(self.stage)
            if distort_records_native(own.images, idx, oy, ox, flip, factor, own.size, out=dst) is None:
                raise RuntimeError("the device-resident CIFAR pipeline needs the native library")

    def to(self, device, non_blocking: bool = False) -> torch.Tensor:
        """Materialise as a fresh tensor (what a consumer without a destination buffer asks for)."""
        out = torch.empty(self.shape, dtype=torch.float32, device=device)
        self.copy_into(out)
        return out


class DeviceCifar10Batches(Cifar10Batches):
    """CIFAR-10 with the records resident on the training device.

    The uint8 records of this rank (150 MB for the whole training set) are uploaded once; after that a step sends
    21 bytes per image (record number + draws) instead of 12 KB of float pixels, and crop / flip / contrast /
    standardisation run in one kernel per batch next to the consumer — the host cores only draw random numbers.
    Batch ``k`` is the same function of ``(seed, rank, k)`` as in :class:`Cifar10Batches`: same permutation, same
    draws, same arithmetic (the two pipelines agree to float rounding).  Yields ``(RecordBatch, labels)``;
    ``utils.pipeline.DevicePrefetcher`` turns them into device tensors."""

    def __init__(self, data_dir: str, batch_size: int, image_size: int = IMAGE_SIZE, device="cuda", **kw):
        if image_size > 32:
            raise ValueError("the device-resident pipeline crops; image_size must not exceed 32")
        self.device = torch.device(device)
        kw["num_threads"] = 1
        kw["native"] = False
        self.images_dev = None
        super().__init__(data_dir, batch_size, image_size, **kw)
        if self.device.type == "cuda":
            self.images_dev = self.images.to(self.device)

    def _make(self, k: int):
        epoch, j = divmod(k, self.bpe)
        if epoch > 0 and not self.repeat:
            return None
        idx = self._indices(epoch, j)
        y = self.labels[idx].clone()
        if self.pin:
            y = y.pin_memory()
        return RecordBatch(self, idx, *self._draws(idx.numel(), k)), y


def distorted_inputs(data_dir: str, batch_size: int, image_size: int = IMAGE_SIZE, **kw) -> Cifar10Batches:
    """Training batches (reference ``cifar10.distorted_inputs``, cnn/cifar10.py:165)."""
    return Cifar10Batches(data_dir, batch_size, image_size, train=True, **kw)


def inputs(eval_data: bool, data_dir: str, batch_size: int, image_size: int = IMAGE_SIZE, **kw) -> Cifar10Batches:
    """Evaluation batches (reference ``cifar10.inputs``, cnn/cifar10.py:183)."""
    kw.setdefault("repeat", False)
    return Cifar10Batches(data_dir, batch_size, image_size, train=not eval_data, shuffle=False, **kw) if not eval_data \
        else Cifar10Batches(data_dir, batch_size, image_size, train=False, shuffle=False, **kw)
