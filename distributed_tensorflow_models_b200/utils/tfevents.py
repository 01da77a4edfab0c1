"""TensorBoard event files (``events.out.tfevents.<time>.<host>``) written without TensorFlow.

The reference's evaluators write a ``Precision @ 1`` (Inception: also ``Recall @ 5``) summary per evaluated
checkpoint into ``eval_dir`` through ``tf.summary.FileWriter`` (``cnn/cifar10_eval.py:106-109``,
``inception/inception_eval.py:125-129``) and the Supervisor logs ``global_step/sec``.  An event file is
a TFRecord stream (``data/tfrecord.py`` framing) of ``Event`` protocol buffers: the first record carries
``file_version = "brain.Event:2"``, every further one ``wall_time``, ``step`` and a ``Summary`` with
``(tag, simple_value)`` pairs.  The three messages are small enough to encode by hand.

Histogram summaries (``tf.summary.histogram``: the reference defines them for every activation, variable and gradient
— ``cnn/cifar10.py:104-121,376-382``, ``inception/inception_model.py:139-158`` — although its trainers run with
``summary_op=None``) are ``HistogramProto`` messages with TensorFlow's default bucket limits (±1e-12 · 1.1^k).
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Dict, Iterator, List, Optional, Tuple

from ..data.tfrecord import TFRecordWriter, _read_varint, _varint, read_records


def _tag(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _ld(field: int, payload: bytes) -> bytes:
    return _tag(field, 2) + _varint(len(payload)) + payload


def default_bucket_limits() -> List[float]:
    """TensorFlow's histogram buckets: 1e-12 * 1.1^k up to 1e20, mirrored for negative values, closed by DBL_MAX."""
    pos, v = [], 1e-12
    while v < 1e20:
        pos.append(v)
        v *= 1.1
    return [-x for x in reversed(pos)] + [0.0] + pos + [1.7976931348623157e308]


_LIMITS = None


def encode_histogram(values) -> bytes:
    """``HistogramProto`` of a tensor / array: min, max, num, sum, sum_squares, bucket_limit[], bucket[] (only the
    occupied range of buckets is written, as TensorFlow does)."""
    import numpy as np
    global _LIMITS
    if _LIMITS is None:
        _LIMITS = np.asarray(default_bucket_limits())
    a = np.asarray(values.detach().cpu().numpy() if hasattr(values, "detach") else values, dtype=np.float64).ravel()
    if a.size == 0:
        a = np.zeros(1)
    idx = np.searchsorted(_LIMITS, a, side="left")
    idx = np.minimum(idx, len(_LIMITS) - 1)
    counts = np.bincount(idx, minlength=len(_LIMITS)).astype(np.float64)
    nz = np.nonzero(counts)[0]
    lo, hi = int(nz[0]), int(nz[-1])
    lo = max(lo - 1, 0)                    # TensorFlow keeps one empty bucket in front of the first occupied one
    out = b"".join(_tag(f, 1) + struct.pack("<d", float(v)) for f, v in
                   ((1, a.min()), (2, a.max()), (3, a.size), (4, a.sum()), (5, (a * a).sum())))
    out += _ld(6, b"".join(struct.pack("<d", float(v)) for v in _LIMITS[lo:hi + 1]))
    out += _ld(7, b"".join(struct.pack("<d", float(v)) for v in counts[lo:hi + 1]))
    return out


def encode_event(wall_time: float, step: int = 0, file_version: Optional[str] = None,
                 scalars: Optional[Dict[str, float]] = None, histograms: Optional[Dict[str, object]] = None) -> bytes:
    out = _tag(1, 1) + struct.pack("<d", float(wall_time))
    if step:
        out += _tag(2, 0) + _varint(int(step))
    if file_version is not None:
        out += _ld(3, file_version.encode("utf-8"))
    if scalars or histograms:
        summary = b""
        for tag, value in (scalars or {}).items():
            summary += _ld(1, _ld(1, tag.encode("utf-8")) + _tag(2, 5) + struct.pack("<f", float(value)))
        for tag, values in (histograms or {}).items():
            summary += _ld(1, _ld(1, tag.encode("utf-8")) + _ld(5, encode_histogram(values)))
        out += _ld(5, summary)
    return out


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = _read_varint(buf, pos)
        elif wire == 1:
            val = buf[pos:pos + 8]
            pos += 8
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            val = buf[pos:pos + n]
            pos += n
        elif wire == 5:
            val = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError("unsupported wire type %d" % wire)
        yield field, wire, val


def _decode_histogram(buf: bytes) -> dict:
    h = {"bucket_limit": [], "bucket": []}
    names = {1: "min", 2: "max", 3: "num", 4: "sum", 5: "sum_squares"}
    for f, wire, v in _fields(buf):
        if f in names:
            h[names[f]] = struct.unpack("<d", v)[0]
        elif f in (6, 7):
            vals = [struct.unpack("<d", v[i:i + 8])[0] for i in range(0, len(v), 8)] if wire == 2 \
                else [struct.unpack("<d", v)[0]]
            h["bucket_limit" if f == 6 else "bucket"] += vals
    return h


def decode_event(record: bytes) -> dict:
    ev = {"wall_time": 0.0, "step": 0, "file_version": None, "scalars": {}, "histograms": {}}
    for field, _, val in _fields(record):
        if field == 1:
            ev["wall_time"] = struct.unpack("<d", val)[0]
        elif field == 2:
            ev["step"] = val
        elif field == 3:
            ev["file_version"] = val.decode("utf-8")
        elif field == 5:
            for f2, _, v2 in _fields(val):
                if f2 != 1:
                    continue
                tag, simple, histo = None, None, None
                for f3, _, v3 in _fields(v2):
                    if f3 == 1:
                        tag = v3.decode("utf-8")
                    elif f3 == 2:
                        simple = struct.unpack("<f", v3)[0]
                    elif f3 == 5:
                        histo = _decode_histogram(v3)
                if tag is not None and histo is not None:
                    ev["histograms"][tag] = histo
                if tag is not None and simple is not None:
                    ev["scalars"][tag] = simple
    return ev


class EventFileWriter:
    """``tf.summary.FileWriter(logdir)`` for scalar and histogram summaries."""

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.path = os.path.join(logdir, "events.out.tfevents.%010d.%s%s" % (int(time.time()), socket.gethostname(),
                                                                             filename_suffix))
        self._w = TFRecordWriter(self.path)
        self._w.write(encode_event(time.time(), file_version="brain.Event:2"))
        self.flush()

    def add_scalars(self, step: int, scalars: Dict[str, float], wall_time: Optional[float] = None) -> None:
        self._w.write(encode_event(time.time() if wall_time is None else wall_time, step, scalars=scalars))

    def add_summaries(self, step: int, scalars: Optional[Dict[str, float]] = None,
                      histograms: Optional[Dict[str, object]] = None, wall_time: Optional[float] = None) -> None:
        """Scalars and histograms (tensors / arrays) of one step in one event."""
        self._w.write(encode_event(time.time() if wall_time is None else wall_time, step, scalars=scalars,
                                   histograms=histograms))

    def flush(self) -> None:
        self._w.f.flush()

    def close(self) -> None:
        self._w.close()


def read_events(path: str) -> List[dict]:
    return [decode_event(r) for r in read_records(path, check_crc=True)]
