"""Writer for torch.load-compatible checkpoint files fed from a pinned host arena.

The reference writes its vanilla checkpoint with `torch.save(state, path)`
(pyrecover/checkpoint.py:74) and reads it back with `torch.load` (:182), and the reference's
own test (tests/check_weights_equality.py:114) opens the file with a bare
`torch.load(path)["model"]`.  So the on-disk format is part of the boundary: a ZIP_STORED
archive `<stem>/data.pkl`, `<stem>/byteorder`, `<stem>/data/<key>` (64-byte aligned),
`<stem>/version`.

Here the big records are NOT serialised tensor by tensor: every tensor of the state is a view
(storage offset + shape + stride) into one of a few large records, and each record is written
straight from the pinned arena the snapshot kernel filled.  The pickle is produced by a small
Pickler whose persistent ids are exactly what torch.save emits, so `torch.load(...,
weights_only=True)` accepts it.
"""
from __future__ import annotations

import hashlib
import io
import os
import pickle
import struct
import zipfile
from collections import OrderedDict
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

_STORAGE_CLASS = {
    torch.bfloat16: torch.BFloat16Storage, torch.float32: torch.FloatStorage,
    torch.float16: torch.HalfStorage, torch.float64: torch.DoubleStorage,
    torch.int64: torch.LongStorage, torch.int32: torch.IntStorage, torch.uint8: torch.ByteStorage,
    torch.int16: torch.ShortStorage, torch.int8: torch.CharStorage, torch.bool: torch.BoolStorage,
}
RECORD_ALIGN = 64
WRITE_CHUNK = 64 << 20


@dataclass(frozen=True)
class RecordRef:
    """A typed storage of the archive: record `key`, `numel` elements of `dtype`."""
    key: str
    dtype: torch.dtype
    numel: int


class TensorStub:
    """Stands for a tensor that is a view into a record; pickles exactly like the tensor would."""

    __slots__ = ("record", "offset", "shape", "stride")

    def __init__(self, record: RecordRef, offset: int, shape: Sequence[int], stride: Optional[Sequence[int]] = None):
        self.record = record
        self.offset = int(offset)                 # in elements
        self.shape = tuple(int(s) for s in shape)
        if stride is None:
            stride, acc = [], 1
            for d in reversed(self.shape):
                stride.append(acc)
                acc *= d
            stride = tuple(reversed(stride))
        self.stride = tuple(int(s) for s in stride)

    def __reduce_ex__(self, protocol):
        return (torch._utils._rebuild_tensor_v2,
                (self.record, self.offset, self.shape, self.stride, False, OrderedDict()))


class _Pickler(pickle.Pickler):
    def persistent_id(self, obj):
        if isinstance(obj, RecordRef):
            return ("storage", _STORAGE_CLASS[obj.dtype], obj.key, "cpu", obj.numel)
        return None


def dumps_state(state: Any) -> bytes:
    """Pickle `state` (nested dicts/lists with TensorStub leaves and small real tensors turned
    into stubs by the caller) the way torch.save would."""
    buf = io.BytesIO()
    _Pickler(buf, protocol=2).dump(state)
    return buf.getvalue()


def _padding_extra(header_offset: int, name: str, zip64: bool) -> bytes:
    fixed = 30 + len(name.encode()) + (20 if zip64 else 0)
    pad = (-(header_offset + fixed + 4)) % RECORD_ALIGN
    return struct.pack("<HH", 0x4246, pad) + b"Z" * pad


def write_archive(path: str, pickle_bytes: bytes, records: Sequence[Tuple[str, Any]],
                  md5: bool = False, fsync: bool = True) -> Dict[str, Any]:
    """Write `<path>.tmp` then rename to `path` (atomic: a walltime kill mid-save can no longer
    leave a corrupt "latest", reference defect pyrecover/checkpoint.py:74).

    records: (key, buffer) pairs; buffer is a bytes-like (memoryview of the pinned arena) or a
    (total_bytes, iterator of bytes-likes) pair -- the iterator form lets the writer thread
    stream the image chunk by chunk while later chunks are still crossing PCIe.
    Returns {"md5": hex or None, "offsets": {key: file offset of the record's data}, "bytes": n}.
    """
    path = str(path)
    stem = os.path.splitext(os.path.basename(path))[0]
    tmp = path + ".tmp"
    offsets: Dict[str, int] = {}
    with open(tmp, "wb") as raw:
        with zipfile.ZipFile(raw, "w", compression=zipfile.ZIP_STORED, allowZip64=True) as zf:
            def put(name: str, data, big_ok: bool = True):
                if isinstance(data, tuple):
                    total, chunks = data
                else:
                    mv = memoryview(data).cast("B")
                    total = mv.nbytes
                    chunks = (mv[o:o + WRITE_CHUNK] for o in range(0, total, WRITE_CHUNK))
                zip64 = total >= (1 << 31)
                zi = zipfile.ZipInfo(f"{stem}/{name}", date_time=(1980, 1, 1, 0, 0, 0))
                zi.compress_type = zipfile.ZIP_STORED
                zi.file_size = total
                zi.extra = _padding_extra(raw.tell(), zi.filename, zip64)
                with zf.open(zi, "w", force_zip64=zip64) as w:
                    data_off = raw.tell()
                    written = 0
                    for chunk in chunks:
                        w.write(chunk)
                        written += memoryview(chunk).nbytes
                    if written != total:
                        raise IOError(f"record {name}: wrote {written} of {total} bytes")
                return data_off
            put("data.pkl", pickle_bytes)
            put("byteorder", b"little")
            for key, buf in records:
                offsets[key] = put(f"data/{key}", buf)
                assert offsets[key] % RECORD_ALIGN == 0
            put("version", b"3\n")
        raw.flush()
        if fsync:
            os.fsync(raw.fileno())
        nbytes = raw.tell()
    digest = None
    if md5:
        h = hashlib.md5()
        with open(tmp, "rb") as f:              # page-cache hot; hashlib releases the GIL
            while True:
                chunk = f.read(WRITE_CHUNK)
                if not chunk:
                    break
                h.update(chunk)
        digest = h.hexdigest()
    os.replace(tmp, path)
    return {"md5": digest, "offsets": offsets, "bytes": nbytes}


def record_offsets(path: str) -> Dict[str, Tuple[int, int]]:
    """{record name without the stem: (data offset, size)} of a torch zip checkpoint."""
    out = {}
    with zipfile.ZipFile(path, "r") as zf:
        with open(path, "rb") as f:
            for zi in zf.infolist():
                f.seek(zi.header_offset)
                hdr = f.read(30)
                nlen, elen = struct.unpack("<HH", hdr[26:30])
                name = zi.filename.split("/", 1)[1] if "/" in zi.filename else zi.filename
                out[name] = (zi.header_offset + 30 + nlen + elen, zi.file_size)
    return out
