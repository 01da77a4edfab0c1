"""TensorBoard event files (``events.out.tfevents.<time>.<host>``) written without TensorFlow.

The reference's evaluators write a ``Precision @ 1`` (Inception: also ``Recall @ 5``) summary per evaluated
checkpoint into ``eval_dir`` through ``tf.summary.FileWriter`` (``cnn/cifar10_eval.py:106-109``,
``inception/inception_eval.py:125-129``) and the Supervisor logs ``global_step/sec``.  An event file is
a TFRecord stream (``data/tfrecord.py`` framing) of ``Event`` protocol buffers: the first record carries
``file_version = "brain.Event:2"``, every further one ``wall_time``, ``step`` and a ``Summary`` with
``(tag, simple_value)`` pairs.  The three messages are small enough to encode by hand.
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


def encode_event(wall_time: float, step: int = 0, file_version: Optional[str] = None,
                 scalars: Optional[Dict[str, float]] = None) -> bytes:
    out = _tag(1, 1) + struct.pack("<d", float(wall_time))
    if step:
        out += _tag(2, 0) + _varint(int(step))
    if file_version is not None:
        out += _ld(3, file_version.encode("utf-8"))
    if scalars:
        summary = b""
        for tag, value in scalars.items():
            summary += _ld(1, _ld(1, tag.encode("utf-8")) + _tag(2, 5) + struct.pack("<f", float(value)))
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


def decode_event(record: bytes) -> dict:
    ev = {"wall_time": 0.0, "step": 0, "file_version": None, "scalars": {}}
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
                tag, simple = None, None
                for f3, _, v3 in _fields(v2):
                    if f3 == 1:
                        tag = v3.decode("utf-8")
                    elif f3 == 2:
                        simple = struct.unpack("<f", v3)[0]
                if tag is not None and simple is not None:
                    ev["scalars"][tag] = simple
    return ev


class EventFileWriter:
    """``tf.summary.FileWriter(logdir)`` for scalar summaries."""

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.path = os.path.join(logdir, "events.out.tfevents.%010d.%s%s" % (int(time.time()), socket.gethostname(),
                                                                             filename_suffix))
        self._w = TFRecordWriter(self.path)
        self._w.write(encode_event(time.time(), file_version="brain.Event:2"))
        self.flush()

    def add_scalars(self, step: int, scalars: Dict[str, float], wall_time: Optional[float] = None) -> None:
        self._w.write(encode_event(time.time() if wall_time is None else wall_time, step, scalars=scalars))

    def flush(self) -> None:
        self._w.f.flush()

    def close(self) -> None:
        self._w.close()


def read_events(path: str) -> List[dict]:
    return [decode_event(r) for r in read_records(path, check_crc=True)]
