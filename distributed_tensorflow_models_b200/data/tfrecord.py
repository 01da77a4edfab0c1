"""TFRecord files and ``tf.train.Example`` protos without TensorFlow.

The reference stores ImageNet as sharded TFRecords of ``Example`` protos (inception/data/build_imagenet_data.py,
read back by inception/image_processing.py:339-407).  Both formats are simple enough to speak natively:

* TFRecord framing: ``uint64 length | uint32 masked_crc32c(length) | bytes data | uint32 masked_crc32c(data)``
  (CRC-32C, Castagnoli; mask = rotate-right 15 + 0xa282ead8).
* ``Example`` = ``Features{ map<string, Feature> }`` with ``Feature`` one of BytesList / FloatList / Int64List —
  encoded/decoded here with a ~60-line protobuf wire-format reader/writer.
"""
from __future__ import annotations

import struct
from typing import Dict, Iterator, List, Union

FeatureValue = Union[List[bytes], List[float], List[int]]

# ---- CRC-32C ------------------------------------------------------------------------------------------------
_TABLE = []
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ 0x82F63B78 if _c & 1 else _c >> 1
    _TABLE.append(_c)


_NATIVE = [None]


def _native_crc():
    """``dtf_crc32c`` of the native library (host code, slice-by-8) when the library is there; False otherwise."""
    if _NATIVE[0] is None:
        try:
            import ctypes as C

            from .. import _native as nat
            if not nat.available():
                raise OSError("native library not built")
            fn = nat.lib().dtf_crc32c
            fn.restype = C.c_uint32
            fn.argtypes = [C.c_char_p, C.c_ulonglong, C.c_uint32]
            _NATIVE[0] = fn
        except Exception:
            _NATIVE[0] = False
    return _NATIVE[0]


def crc32c(data, crc: int = 0) -> int:
    """CRC-32C of ``data`` (bytes-like), continuing from ``crc``."""
    data = bytes(data) if not isinstance(data, bytes) else data
    fn = _native_crc() if len(data) >= 64 else False
    if fn:
        return int(fn(data, len(data), crc))
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = _TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def mask_crc(c: int) -> int:
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def unmask_crc(m: int) -> int:
    r = (m - 0xA282EAD8) & 0xFFFFFFFF
    return ((r >> 17) | (r << 15)) & 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ---- record framing ------------------------------------------------------------------------------------------
class TFRecordWriter:
    def __init__(self, path: str):
        self.f = open(path, "wb")

    def write(self, record: bytes) -> None:
        hdr = struct.pack("<Q", len(record))
        self.f.write(hdr)
        self.f.write(struct.pack("<I", masked_crc(hdr)))
        self.f.write(record)
        self.f.write(struct.pack("<I", masked_crc(record)))

    def close(self) -> None:
        self.f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def read_records(path: str, check_crc: bool = False) -> Iterator[bytes]:
    with open(path, "rb") as f:
        while True:
            hdr = f.read(12)
            if len(hdr) < 12:
                return
            (n,) = struct.unpack("<Q", hdr[:8])
            data = f.read(n)
            crc = f.read(4)
            if len(data) < n or len(crc) < 4:
                raise IOError(f"truncated TFRecord file {path}")
            if check_crc and struct.unpack("<I", crc)[0] != masked_crc(data):
                raise IOError(f"corrupt record in {path}")
            yield data


# ---- protobuf wire format (just enough for Example) ----------------------------------------------------------
def _varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int):
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _ld(field: int, payload: bytes) -> bytes:          # length-delimited field
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_example(features: Dict[str, FeatureValue]) -> bytes:
    entries = b""
    for key in sorted(features):
        vals = features[key]
        if len(vals) and isinstance(vals[0], (bytes, bytearray, str)):
            inner = b"".join(_ld(1, v.encode() if isinstance(v, str) else bytes(v)) for v in vals)
            feat = _ld(1, inner)                                              # BytesList
        elif len(vals) and isinstance(vals[0], float):
            feat = _ld(2, _ld(1, struct.pack("<%df" % len(vals), *vals)))       # FloatList (packed)
        else:
            feat = _ld(3, _ld(1, b"".join(_varint(int(v)) for v in vals)))      # Int64List (packed)
        entry = _ld(1, key.encode()) + _ld(2, feat)                           # map entry: key, value
        entries += _ld(1, entry)
    return _ld(1, entries)                                                    # Example.features


def _fields(buf: bytes):
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 2:
            n, pos = _read_varint(buf, pos)
            yield field, wt, buf[pos:pos + n]
            pos += n
        elif wt == 0:
            v, pos = _read_varint(buf, pos)
            yield field, wt, v
        elif wt == 5:
            yield field, wt, buf[pos:pos + 4]
            pos += 4
        elif wt == 1:
            yield field, wt, buf[pos:pos + 8]
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wt}")


def decode_example(record: bytes) -> Dict[str, FeatureValue]:
    out: Dict[str, FeatureValue] = {}
    for f1, _, features in _fields(record):
        if f1 != 1:
            continue
        for f2, _, entry in _fields(features):
            if f2 != 1:
                continue
            key, feat = None, b""
            for f3, _, v in _fields(entry):
                if f3 == 1:
                    key = v.decode()
                elif f3 == 2:
                    feat = v
            vals: FeatureValue = []
            for kind, _, lst in _fields(feat):
                if kind == 1:
                    vals = [v for f, _, v in _fields(lst) if f == 1]
                elif kind == 2:
                    vals = []
                    for f, wt, v in _fields(lst):
                        if f == 1 and wt == 2:
                            vals.extend(struct.unpack("<%df" % (len(v) // 4), v))
                        elif f == 1 and wt == 5:
                            vals.append(struct.unpack("<f", v)[0])
                elif kind == 3:
                    vals = []
                    for f, wt, v in _fields(lst):
                        if f == 1 and wt == 2:
                            p = 0
                            while p < len(v):
                                x, p = _read_varint(v, p)
                                vals.append(x - (1 << 64) if x >= 1 << 63 else x)
                        elif f == 1 and wt == 0:
                            vals.append(v - (1 << 64) if v >= 1 << 63 else v)
            if key is not None:
                out[key] = vals
    return out
