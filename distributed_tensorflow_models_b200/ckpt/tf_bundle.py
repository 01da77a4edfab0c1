"""TensorFlow "V2" checkpoint bundles: ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``.

This is the on-disk format ``tf.train.Saver`` writes for the reference (``model.ckpt-<step>.{index,data-*}``,
SURVEY §5.4), implemented from its published layout so that checkpoints can be exchanged with TensorFlow-side
tooling without TensorFlow being installed here:

* the **data** file is the raw little-endian bytes of every tensor, back to back;
* the **index** file is an immutable sorted string table (the LevelDB table format TensorFlow uses in
  ``tensorflow/core/lib/io/table*``): data blocks of prefix-compressed ``key -> value`` entries with restart points
  every 16 entries, each block followed by a 1-byte compression tag (0 = none) and a masked CRC-32C; a (here empty)
  meta-index block; an index block mapping a separator key to each data block's (offset, size); a 48-byte footer with
  the two block handles and the magic number ``0xdb4775248b80fb57``;
* key ``""`` holds a ``BundleHeaderProto`` (``num_shards``, endianness, version), every other key is a tensor name
  whose value is a ``BundleEntryProto`` (dtype, shape, shard, offset, size, masked CRC-32C of the tensor bytes).

The protocol-buffer messages involved are tiny and are encoded / decoded by hand (varints and length-delimited
fields), like the ``Example`` protos of ``data/tfrecord.py``.

Partitioned variables (the reference's ``fixed_size_partitioner`` trainers: CifarNet, AlexNet, VGG) are stored by
TensorFlow as one entry under the full name that lists its ``slices`` (start / length per dimension) and carries no
data, plus one data entry per slice under an *ordered-code* key (a 0 byte, the escaped name, the slice extents in an
order-preserving integer encoding).  ``read_bundle`` reassembles them into the whole tensor — which is how this
framework keeps every variable — and ``write_bundle(..., partitions=...)`` can emit them (used by the tests, and for
handing a checkpoint back to a partitioned TensorFlow graph).

Verified here by round trips, by structural checks of the emitted bytes and by corruption tests
(``tests/test_ckpt_data_cli.py``); there is no TensorFlow on this box to cross-read the files with.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

from ..data.tfrecord import _read_varint, _varint, crc32c, mask_crc, unmask_crc

TABLE_MAGIC = 0xDB4775248B80FB57
FOOTER_LEN = 48
BLOCK_TRAILER_LEN = 5
RESTART_INTERVAL = 16
BLOCK_SIZE = 256 * 1024
HEADER_KEY = b""

# tensorflow/core/framework/types.proto
DT = {np.dtype("float32"): 1, np.dtype("float64"): 2, np.dtype("int32"): 3, np.dtype("uint8"): 4,
      np.dtype("int16"): 5, np.dtype("int8"): 6, np.dtype("int64"): 9, np.dtype("bool"): 10,
      np.dtype("uint16"): 17, np.dtype("float16"): 19, np.dtype("uint32"): 22, np.dtype("uint64"): 23}
DT_INV = {v: k for k, v in DT.items()}


# ---- sorted string table: writer -------------------------------------------------------------------------------
class _BlockBuilder:
    def __init__(self, restart_interval: int = RESTART_INTERVAL):
        self.interval = restart_interval
        self.reset()

    def reset(self) -> None:
        self.buf = bytearray()
        self.restarts = [0]
        self.counter = 0
        self.last_key = b""

    def empty(self) -> bool:
        return not self.buf

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.counter < self.interval:
            n = min(len(key), len(self.last_key))
            while shared < n and key[shared] == self.last_key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.counter = 0
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.counter += 1

    def finish(self) -> bytes:
        out = bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts)
        return out + struct.pack("<I", len(self.restarts))


def _handle(offset: int, size: int) -> bytes:
    return _varint(offset) + _varint(size)


class TableWriter:
    """Keys must be added in strictly increasing bytewise order."""

    def __init__(self, f, block_size: int = BLOCK_SIZE):
        self.f = f
        self.block_size = block_size
        self.offset = 0
        self.data = _BlockBuilder()
        self.index = _BlockBuilder(restart_interval=1)
        self.last_key: Optional[bytes] = None
        self.pending: Optional[Tuple[bytes, bytes]] = None      # (separator key, handle) of the block just flushed

    def _write_block(self, contents: bytes) -> Tuple[int, int]:
        trailer = b"\x00" + struct.pack("<I", mask_crc(crc32c(contents + b"\x00")))
        self.f.write(contents + trailer)
        off = self.offset
        self.offset += len(contents) + BLOCK_TRAILER_LEN
        return off, len(contents)

    def _flush(self) -> None:
        if self.data.empty():
            return
        off, size = self._write_block(self.data.finish())
        # any key >= every key of the block and < the next block's first key separates them; the block's own last
        # key is the simplest such key
        self.pending = (self.data.last_key, _handle(off, size))
        self.data.reset()

    def add(self, key: bytes, value: bytes) -> None:
        if self.last_key is not None and key <= self.last_key:
            raise ValueError("table keys must be added in increasing order: %r after %r" % (key, self.last_key))
        if self.pending is not None:
            self.index.add(*self.pending)
            self.pending = None
        self.data.add(key, value)
        self.last_key = key
        if self.data.size_estimate() >= self.block_size:
            self._flush()

    def finish(self) -> None:
        self._flush()
        if self.pending is not None:
            self.index.add(*self.pending)
            self.pending = None
        meta = self._write_block(_BlockBuilder().finish())           # no filter / properties: empty meta-index
        idx = self._write_block(self.index.finish())
        footer = _handle(*meta) + _handle(*idx)
        footer += b"\x00" * (40 - len(footer))
        footer += struct.pack("<II", TABLE_MAGIC & 0xFFFFFFFF, TABLE_MAGIC >> 32)
        self.f.write(footer)
        self.offset += len(footer)


# ---- sorted string table: reader -------------------------------------------------------------------------------
def _read_block(buf: bytes, offset: int, size: int, verify: bool = True) -> bytes:
    contents = buf[offset:offset + size]
    trailer = buf[offset + size:offset + size + BLOCK_TRAILER_LEN]
    if len(contents) != size or len(trailer) != BLOCK_TRAILER_LEN:
        raise ValueError("truncated table block")
    if trailer[0] != 0:
        raise NotImplementedError("compressed table blocks (type %d) are not supported" % trailer[0])
    if verify and unmask_crc(struct.unpack("<I", trailer[1:])[0]) != crc32c(contents + trailer[:1]):
        raise ValueError("table block checksum mismatch at offset %d" % offset)
    return contents


def _block_entries(block: bytes) -> Iterator[Tuple[bytes, bytes]]:
    if len(block) < 4:
        raise ValueError("bad table block")
    num_restarts = struct.unpack("<I", block[-4:])[0]
    limit = len(block) - 4 - 4 * num_restarts
    pos, key = 0, b""
    while pos < limit:
        shared, pos = _read_varint(block, pos)
        unshared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        key = key[:shared] + block[pos:pos + unshared]
        pos += unshared
        yield key, block[pos:pos + vlen]
        pos += vlen


def read_table(buf: bytes, verify: bool = True) -> List[Tuple[bytes, bytes]]:
    if len(buf) < FOOTER_LEN:
        raise ValueError("file too short to be a table")
    footer = buf[-FOOTER_LEN:]
    lo, hi = struct.unpack("<II", footer[40:])
    if (hi << 32 | lo) != TABLE_MAGIC:
        raise ValueError("not a TensorFlow table file (bad magic number)")
    pos = 0
    _, pos = _read_varint(footer, pos)          # meta-index handle (unused)
    _, pos = _read_varint(footer, pos)
    ioff, pos = _read_varint(footer, pos)
    isize, pos = _read_varint(footer, pos)
    out: List[Tuple[bytes, bytes]] = []
    for _, handle in _block_entries(_read_block(buf, ioff, isize, verify)):
        boff, p = _read_varint(handle, 0)
        bsize, _ = _read_varint(handle, p)
        out.extend(_block_entries(_read_block(buf, boff, bsize, verify)))
    return out


# ---- the two protocol buffers ----------------------------------------------------------------------------------
def _tag(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _encode_header(num_shards: int = 1) -> bytes:
    version = _tag(1, 0) + _varint(1)                                        # VersionDef.producer = 1
    return _tag(1, 0) + _varint(num_shards) + _tag(3, 2) + _varint(len(version)) + version   # endianness 0 = LITTLE


def _encode_entry(dtype: int, shape: Tuple[int, ...], shard_id: int, offset: int, size: int, crc: int) -> bytes:
    dims = b""
    for d in shape:
        dim = _tag(1, 0) + _varint(int(d))
        dims += _tag(2, 2) + _varint(len(dim)) + dim
    out = _tag(1, 0) + _varint(dtype) + _tag(2, 2) + _varint(len(dims)) + dims
    if shard_id:
        out += _tag(3, 0) + _varint(shard_id)
    if offset:
        out += _tag(4, 0) + _varint(offset)
    if size:
        out += _tag(5, 0) + _varint(size)
    return out + _tag(6, 5) + struct.pack("<I", crc)


def _encode_slice_proto(extents: List[Tuple[int, Optional[int]]]) -> bytes:
    """TensorSliceProto: one Extent {start, length} per dimension; a full dimension has neither set."""
    out = b""
    for start, length in extents:
        ext = b""
        if length is not None:
            if start:
                ext += _tag(1, 0) + _varint(start)
            ext += _tag(2, 0) + _varint(length)
        out += _tag(1, 2) + _varint(len(ext)) + ext
    return out


# ---- ordered code (tensorflow/core/lib/strings/ordered_code): the keys of slice data entries -----------------------
def _oc_num_increasing(v: int) -> bytes:
    body = v.to_bytes((v.bit_length() + 7) // 8, "big") if v else b""
    return bytes([len(body)]) + body


def _oc_string(s: bytes) -> bytes:
    esc = b"".join(b"\x00\xff" if c == 0 else (b"\xff\x00" if c == 0xFF else bytes([c])) for c in s)
    return esc + b"\x00\x01"


def _oc_signed_increasing(v: int) -> bytes:
    """Order-preserving signed encoding: 1 byte holds 6 payload bits, every further byte 7 more; the leading bits
    count the length in unary, inverted for negative numbers."""
    x = ~v if v < 0 else v
    n = 1
    while x >= (1 << (7 * n - 1)):
        n += 1
    raw = (v & ((1 << 80) - 1)).to_bytes(10, "big")[10 - n:] if n <= 10 else None
    buf = bytearray(raw)
    header = [(0, 0), (0x80, 0), (0xC0, 0), (0xE0, 0), (0xF0, 0), (0xF8, 0), (0xFC, 0), (0xFE, 0), (0xFF, 0),
              (0xFF, 0x80), (0xFF, 0xC0)][n]
    buf[0] ^= header[0]
    if n > 1:
        buf[1] ^= header[1]
    return bytes(buf)


def encode_slice_key(name: str, extents: List[Tuple[int, Optional[int]]]) -> bytes:
    """``checkpoint::EncodeTensorNameSlice``: 0, the name, the rank, then (start, length) per dimension with (-1, -1)
    standing for "all of it"."""
    key = _oc_num_increasing(0) + _oc_string(name.encode("utf-8")) + _oc_num_increasing(len(extents))
    for start, length in extents:
        if length is None:
            key += _oc_signed_increasing(-1) + _oc_signed_increasing(-1)
        else:
            key += _oc_signed_increasing(start) + _oc_signed_increasing(length)
    return key


def decode_slice_key_name(key: bytes) -> Optional[str]:
    """The tensor name inside a slice key (None for an ordinary key)."""
    if not key.startswith(b"\x00") or len(key) < 3:
        return None
    out, i = bytearray(), 1
    while i + 1 < len(key):
        c, d = key[i], key[i + 1]
        if c == 0 and d == 1:
            return out.decode("utf-8")
        if (c == 0 and d == 0xFF) or (c == 0xFF and d == 0):
            out.append(c)
            i += 2
        else:
            out.append(c)
            i += 1
    return None


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            val, pos = _read_varint(buf, pos)
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            val = buf[pos:pos + n]
            pos += n
        elif wire == 5:
            val = struct.unpack("<I", buf[pos:pos + 4])[0]
            pos += 4
        elif wire == 1:
            val = struct.unpack("<Q", buf[pos:pos + 8])[0]
            pos += 8
        else:
            raise ValueError("unsupported protobuf wire type %d" % wire)
        yield field, wire, val


def _decode_entry(buf: bytes) -> dict:
    e = {"dtype": 0, "shape": (), "shard_id": 0, "offset": 0, "size": 0, "crc32c": None, "slices": []}
    for field, _, val in _fields(buf):
        if field == 1:
            e["dtype"] = val
        elif field == 2:
            dims = []
            for f2, _, v2 in _fields(val):
                if f2 == 2:
                    size = 0
                    for f3, _, v3 in _fields(v2):
                        if f3 == 1:
                            size = v3 - (1 << 64) if v3 >= (1 << 63) else v3
                    dims.append(size)
                elif f2 == 3 and v2:
                    raise ValueError("tensor of unknown rank in a checkpoint")
            e["shape"] = tuple(dims)
        elif field == 3:
            e["shard_id"] = val
        elif field == 4:
            e["offset"] = val
        elif field == 5:
            e["size"] = val
        elif field == 6:
            e["crc32c"] = val
        elif field == 7:
            extents = []
            for f2, _, v2 in _fields(val):
                if f2 == 1:
                    start, length = 0, None
                    for f3, _, v3 in _fields(v2):
                        if f3 == 1:
                            start = v3
                        elif f3 == 2:
                            length = v3
                    extents.append((start, length))
            e["slices"].append(extents)
    return e


def _decode_header(buf: bytes) -> dict:
    h = {"num_shards": 0, "endianness": 0}
    for field, _, val in _fields(buf):
        if field == 1:
            h["num_shards"] = val
        elif field == 2:
            h["endianness"] = val
    return h


# ---- bundles -----------------------------------------------------------------------------------------------------
def data_path(prefix: str, shard: int = 0, num_shards: int = 1) -> str:
    return "%s.data-%05d-of-%05d" % (prefix, shard, num_shards)


def index_path(prefix: str) -> str:
    return prefix + ".index"


def is_bundle(prefix: str) -> bool:
    return os.path.exists(index_path(prefix))


def bundle_files(prefix: str) -> List[str]:
    """Every file of the bundle that exists (for keep-N deletion)."""
    d, base = os.path.dirname(prefix) or ".", os.path.basename(prefix)
    out = [index_path(prefix)] if os.path.exists(index_path(prefix)) else []
    if os.path.isdir(d):
        out += [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.startswith(base + ".data-")]
    return out


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray], block_size: int = BLOCK_SIZE,
                 partitions: Optional[Dict[str, List[int]]] = None) -> None:
    """Write ``tensors`` (name -> array) as one-shard bundle ``prefix``; files appear atomically (index last).
    ``partitions``: name -> slice lengths along axis 0 for variables to be stored the way TensorFlow stores a
    partitioned variable (a slice-listing entry + one data entry per slice)."""
    entries = []
    tmp_data = data_path(prefix) + ".tmp"
    offset = 0
    partitions = partitions or {}
    with open(tmp_data, "wb") as f:
        def put(key: bytes, a: np.ndarray) -> None:
            nonlocal offset
            raw = a.tobytes()
            f.write(raw)
            entries.append((key, _encode_entry(DT[a.dtype], a.shape, 0, offset, len(raw), mask_crc(crc32c(raw)))))
            offset += len(raw)

        for name in sorted(tensors, key=lambda s: s.encode("utf-8")):
            if not name:
                raise ValueError("the empty name is reserved for the bundle header")
            a = np.asarray(tensors[name])
            if not a.flags.c_contiguous:           # (np.ascontiguousarray would turn a scalar into shape (1,))
                a = a.copy(order="C")
            if a.dtype not in DT:
                raise TypeError("dtype %s of %r has no TensorFlow checkpoint encoding here" % (a.dtype, name))
            if a.dtype.byteorder == ">":
                a = a.astype(a.dtype.newbyteorder("<"))
            parts = partitions.get(name)
            if not parts or len(parts) < 2:
                put(name.encode("utf-8"), a)
                continue
            if a.ndim == 0 or sum(parts) != a.shape[0]:
                raise ValueError("partition lengths %s do not add up to dim 0 of %r %s" % (parts, name, a.shape))
            listing = _tag(1, 0) + _varint(DT[a.dtype])
            dims = b""
            for d in a.shape:
                dim = _tag(1, 0) + _varint(int(d))
                dims += _tag(2, 2) + _varint(len(dim)) + dim
            listing += _tag(2, 2) + _varint(len(dims)) + dims
            start = 0
            for n in parts:
                extents = [(start, n)] + [(0, None)] * (a.ndim - 1)
                sp = _encode_slice_proto(extents)
                listing += _tag(7, 2) + _varint(len(sp)) + sp
                put(encode_slice_key(name, extents), np.ascontiguousarray(a[start:start + n]))
                start += n
            entries.append((name.encode("utf-8"), listing))
        entries.sort(key=lambda kv: kv[0])
    tmp_index = index_path(prefix) + ".tmp"
    with open(tmp_index, "wb") as f:
        w = TableWriter(f, block_size)
        w.add(HEADER_KEY, _encode_header(1))
        for key, value in entries:
            w.add(key, value)
        w.finish()
    os.replace(tmp_data, data_path(prefix))
    os.replace(tmp_index, index_path(prefix))


def list_bundle(prefix: str) -> Dict[str, Tuple[np.dtype, Tuple[int, ...]]]:
    """name -> (dtype, shape), like ``tf.train.list_variables``."""
    out = {}
    with open(index_path(prefix), "rb") as f:
        for key, value in read_table(f.read()):
            if key == HEADER_KEY or decode_slice_key_name(key) is not None:
                continue
            e = _decode_entry(value)
            out[key.decode("utf-8")] = (DT_INV.get(e["dtype"]), e["shape"])
    return out


def read_bundle(prefix: str, names: Optional[List[str]] = None, verify: bool = True) -> Dict[str, np.ndarray]:
    with open(index_path(prefix), "rb") as f:
        table = read_table(f.read(), verify)
    if not table or table[0][0] != HEADER_KEY:
        raise ValueError("bundle index has no header entry")
    header = _decode_header(table[0][1])
    if header["endianness"] != 0:
        raise NotImplementedError("big-endian checkpoint bundles are not supported")
    num_shards = max(1, header["num_shards"])
    shards: Dict[int, np.memmap] = {}
    out: Dict[str, np.ndarray] = {}
    want = set(names) if names is not None else None
    def fetch(name: str, e: dict) -> np.ndarray:
        if e["dtype"] not in DT_INV:
            raise TypeError("tensor %r has TensorFlow dtype %d, which is not supported" % (name, e["dtype"]))
        if e["shard_id"] not in shards:
            path = data_path(prefix, e["shard_id"], num_shards)
            shards[e["shard_id"]] = np.memmap(path, dtype=np.uint8, mode="r") if os.path.getsize(path) else \
                np.zeros(0, np.uint8)
        raw = bytes(shards[e["shard_id"]][e["offset"]:e["offset"] + e["size"]])
        dtype = DT_INV[e["dtype"]]
        count = int(np.prod(e["shape"])) if e["shape"] else 1
        if len(raw) != count * dtype.itemsize:
            raise ValueError("tensor %r: %d bytes on disk, shape %s needs %d" % (name, len(raw), e["shape"],
                                                                                  count * dtype.itemsize))
        if verify and e["crc32c"] is not None and unmask_crc(e["crc32c"]) != crc32c(raw):
            raise ValueError("tensor %r: checksum mismatch (corrupt data file)" % name)
        return np.frombuffer(raw, dtype=dtype).reshape(e["shape"]).copy()

    # slice data entries (ordered-code keys sort before every ordinary name), grouped by tensor name in key order
    slice_data: Dict[str, List[dict]] = {}
    for key, value in table[1:]:
        sname = decode_slice_key_name(key)
        if sname is not None:
            slice_data.setdefault(sname, []).append(_decode_entry(value))
    for key, value in table[1:]:
        if decode_slice_key_name(key) is not None:
            continue
        name = key.decode("utf-8")
        if want is not None and name not in want:
            continue
        e = _decode_entry(value)
        if not e["slices"]:
            out[name] = fetch(name, e)
            continue
        # partitioned variable: paste the slices into the full tensor.  Slice keys encode (start, length) per
        # dimension with an order-preserving code, so key order == ascending start order of the listed extents.
        parts = slice_data.get(name, [])
        specs = sorted(e["slices"], key=lambda ext: [(-1 if ln is None else st) for st, ln in ext])
        if len(parts) != len(specs):
            raise ValueError("tensor %r lists %d slices but %d slice entries are stored" % (name, len(specs), len(parts)))
        full = np.zeros(e["shape"], dtype=DT_INV[e["dtype"]])
        filled = np.zeros(e["shape"], dtype=bool) if full.size else None
        for ext, pe in zip(specs, parts):
            idx = tuple(slice(None) if ln is None else slice(st, st + ln) for st, ln in ext)
            piece = fetch(name, pe)
            if full[idx].shape != piece.shape:
                raise ValueError("tensor %r: slice %s has shape %s, stored piece %s" % (name, ext, full[idx].shape,
                                                                                         piece.shape))
            full[idx] = piece
            if filled is not None:
                filled[idx] = True
        if filled is not None and not filled.all():
            raise ValueError("tensor %r: the stored slices do not cover the whole variable" % name)
        out[name] = full
    if want is not None and want - set(out):
        raise KeyError("not in the checkpoint: %s" % sorted(want - set(out)))
    return out
