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


# ------------------------------------------------------------------------- zip framing ----
# The container is written by hand (not with the zipfile module) so that the multi-GB records can
# go to disk with positional O_DIRECT writes straight from the pinned arena -- no page-cache copy
# competing with the training loop for memory bandwidth -- while a helper thread computes the
# CRC-32 the container needs; the small parts (local headers, pickle, central directory) are
# written last, once the CRCs are known.
_DOS_TIME, _DOS_DATE = 0, (1 << 5) | 1               # 1980-01-01 00:00:00
_ZIP64_LIMIT = (1 << 31) - 1                         # same threshold as the zipfile module
_BIG_RECORD = 1 << 20                                # records this large start on a 4096-byte file offset
DIRECT_ALIGN = 4096


class _Entry:
    __slots__ = ("name", "size", "data", "chunks", "zip64", "header_offset", "data_offset", "extra", "crc")

    def __init__(self, name, payload):
        self.name = name.encode()
        if isinstance(payload, tuple):               # (total_bytes, iterator of bytes-likes)
            self.size, self.chunks, self.data = int(payload[0]), payload[1], None
        else:
            self.data = memoryview(payload).cast("B")
            self.size, self.chunks = self.data.nbytes, None
        self.zip64 = self.size > _ZIP64_LIMIT
        self.crc = None

    def place(self, header_offset: int) -> int:
        """Fix this entry's offsets; returns the offset just past its data."""
        self.header_offset = header_offset
        align = DIRECT_ALIGN if self.size >= _BIG_RECORD else RECORD_ALIGN
        z64 = struct.pack("<HHQQ", 1, 16, self.size, self.size) if self.zip64 else b""
        fixed = header_offset + 30 + len(self.name) + len(z64) + 4
        pad = (-fixed) % align
        self.extra = z64 + struct.pack("<HH", 0x4246, pad) + b"Z" * pad
        self.data_offset = fixed + pad
        return self.data_offset + self.size

    def local_header(self) -> bytes:
        size32 = 0xFFFFFFFF if self.zip64 else self.size
        return struct.pack("<4s2B4HL2L2H", b"PK\x03\x04", 45 if self.zip64 else 20, 0, 0, 0, _DOS_TIME, _DOS_DATE,
                           self.crc, size32, size32, len(self.name), len(self.extra)) + self.name + self.extra

    def central_header(self) -> bytes:
        fields = []
        size32, off32 = self.size, self.header_offset
        if self.size > _ZIP64_LIMIT:
            fields += [self.size, self.size]
            size32 = 0xFFFFFFFF
        if self.header_offset > _ZIP64_LIMIT:
            fields.append(self.header_offset)
            off32 = 0xFFFFFFFF
        extra = struct.pack("<HH" + "Q" * len(fields), 1, 8 * len(fields), *fields) if fields else b""
        ver = 45 if fields else 20
        return struct.pack("<4s4B4HL2L5H2L", b"PK\x01\x02", ver, 3, ver, 0, 0, 0, _DOS_TIME, _DOS_DATE, self.crc, size32,
                           size32, len(self.name), len(extra), 0, 0, 0, 0o600 << 16, off32) + self.name + extra


def _end_records(count: int, cd_offset: int, cd_size: int) -> bytes:
    out = b""
    if count > 0xFFFF or cd_offset > 0xFFFFFFFF or cd_size > 0xFFFFFFFF or cd_offset > _ZIP64_LIMIT:
        out += struct.pack("<4sQ2H2L4Q", b"PK\x06\x06", 44, 45, 45, 0, 0, count, count, cd_size, cd_offset)
        out += struct.pack("<4sLQL", b"PK\x06\x07", 0, cd_offset + cd_size, 1)
    return out + struct.pack("<4s4H2LH", b"PK\x05\x06", 0, 0, min(count, 0xFFFF), min(count, 0xFFFF),
                             min(cd_size, 0xFFFFFFFF), min(cd_offset, 0xFFFFFFFF), 0)


class _CrcWorker:
    """CRC-32 of a stream of buffers, chained in order on a helper thread (zlib releases the GIL)."""

    def __init__(self):
        import queue
        import threading
        self.q: "queue.Queue" = queue.Queue()
        self.crc = 0
        self.err: Optional[BaseException] = None
        self.t = threading.Thread(target=self._run, name="pyrecover-b200-crc", daemon=True)
        self.t.start()

    def _run(self):
        import zlib
        while True:
            item = self.q.get()
            if item is None:
                return
            try:
                self.crc = zlib.crc32(item, self.crc)
            except BaseException as e:      # noqa: BLE001 -- surfaced by finish()
                self.err = e

    def feed(self, buf) -> None:
        self.q.put(buf)

    def finish(self) -> int:
        self.q.put(None)
        self.t.join()
        if self.err is not None:
            raise self.err
        return self.crc & 0xFFFFFFFF


def _pwrite_all(fd: int, mv: memoryview, off: int) -> None:
    done, n = 0, mv.nbytes
    while done < n:
        done += os.pwrite(fd, mv[done:], off + done)


def _buffer_address(mv: memoryview) -> int:
    import ctypes
    if mv.nbytes == 0 or mv.readonly:
        return 1                                      # treated as unaligned: goes through the buffered descriptor
    return ctypes.addressof(ctypes.c_char.from_buffer(mv))


def write_archive(path: str, pickle_bytes: bytes, records: Sequence[Tuple[str, Any]],
                  md5: bool = False, fsync: bool = True, direct: Optional[bool] = None) -> Dict[str, Any]:
    """Write `<path>.tmp` then rename to `path` (atomic: a walltime kill mid-save can no longer
    leave a corrupt "latest", reference defect pyrecover/checkpoint.py:74).

    records: (key, buffer) pairs; buffer is a bytes-like (memoryview of the pinned arena) or a
    (total_bytes, iterator of bytes-likes) pair -- the iterator form lets the writer thread
    stream the image chunk by chunk while later chunks are still crossing PCIe.  Every buffer
    handed over must stay valid until this function returns (the pinned arena does).

    Big records are written with O_DIRECT when the filesystem takes it and the buffers are
    4096-byte aligned (`direct=None`: try, fall back to buffered writes; False: never).
    md5=True also returns the md5 of the finished file -- the reference's `.md5` side-car,
    pyrecover/checkpoint.py:79-83 -- computed from the buffers in memory, not by re-reading the file.
    Returns {"md5": hex or None, "offsets": {key: file offset of the record's data}, "bytes": n}.
    """
    path = str(path)
    stem = os.path.splitext(os.path.basename(path))[0]
    tmp = path + ".tmp"
    entries = [_Entry(f"{stem}/data.pkl", pickle_bytes), _Entry(f"{stem}/byteorder", b"little")]
    entries += [_Entry(f"{stem}/data/{key}", buf) for key, buf in records]
    entries.append(_Entry(f"{stem}/version", b"3\n"))
    pos = 0
    for e in entries:
        pos = e.place(pos)
    cd_offset = pos

    fd = os.open(tmp, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
    dfd = -1
    if direct is not False and hasattr(os, "O_DIRECT"):
        try:
            dfd = os.open(tmp, os.O_WRONLY | os.O_DIRECT)
        except OSError:
            dfd = -1                                  # tmpfs and friends: buffered writes only
    kept: List[Tuple[_Entry, List[memoryview]]] = []   # streamed buffers, for the md5 pass
    try:
        # 1. the streamed records: data first (their CRC is only known afterwards)
        for e in entries:
            if e.chunks is None:
                import zlib
                e.crc = zlib.crc32(e.data) & 0xFFFFFFFF
                continue
            crc = _CrcWorker()
            seen: List[memoryview] = []
            off, written = e.data_offset, 0
            for chunk in e.chunks:
                mv = memoryview(chunk).cast("B")
                crc.feed(mv)
                if md5:
                    seen.append(mv)
                n = mv.nbytes
                bulk = n - n % DIRECT_ALIGN
                if dfd >= 0 and bulk and off % DIRECT_ALIGN == 0 and _buffer_address(mv) % DIRECT_ALIGN == 0:
                    try:
                        _pwrite_all(dfd, mv[:bulk], off)
                    except OSError:
                        os.close(dfd)
                        dfd = -1                      # this filesystem refuses O_DIRECT writes after all
                        bulk = 0
                    if bulk < n:
                        _pwrite_all(fd, mv[bulk:], off + bulk)
                else:
                    _pwrite_all(fd, mv, off)
                off += n
                written += n
            e.crc = crc.finish()
            if written != e.size:
                raise IOError(f"record {e.name.decode()}: wrote {written} of {e.size} bytes")
            kept.append((e, seen))
        # 2. everything small: local headers, in-memory payloads, central directory, end records
        for e in entries:
            _pwrite_all(fd, memoryview(e.local_header()), e.header_offset)
            if e.data is not None and e.size:
                _pwrite_all(fd, e.data, e.data_offset)
        cd = b"".join(e.central_header() for e in entries)
        tail = cd + _end_records(len(entries), cd_offset, len(cd))
        _pwrite_all(fd, memoryview(tail), cd_offset)
        nbytes = cd_offset + len(tail)
        if fsync:
            os.fsync(fd)
    except BaseException:
        for d in (fd, dfd):
            if d >= 0:
                os.close(d)
        try:
            os.remove(tmp)
        except OSError:
            pass
        raise
    os.close(fd)
    if dfd >= 0:
        os.close(dfd)
    digest = None
    if md5:
        h = hashlib.md5()                             # file order; hashlib releases the GIL on big buffers
        streamed = {id(e): seen for e, seen in kept}
        for e in entries:
            h.update(e.local_header())
            if e.data is not None:
                h.update(e.data)
            else:
                for mv in streamed[id(e)]:
                    h.update(mv)
        h.update(tail)
        digest = h.hexdigest()
    os.replace(tmp, path)
    return {"md5": digest, "offsets": {e.name.decode().split("/data/", 1)[1]: e.data_offset for e in entries
                                       if b"/data/" in e.name}, "bytes": nbytes}


def write_raw(path: str, chunks, fsync: bool = True, direct: Optional[bool] = None, observer=None, epilogue=None) -> int:
    """Stream bytes-likes to `<path>.tmp` and rename (a rank's shard file): O_DIRECT for the
    4096-aligned bulk of every aligned buffer, buffered writes otherwise.  Returns the byte count.
    observer(file_offset, memoryview) sees every chunk as it goes out (checksums); epilogue(fd, end_offset)
    runs before the fsync and may pwrite anywhere in the file -- it returns the final file size."""
    path = str(path)
    tmp = path + ".tmp"
    fd = os.open(tmp, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
    dfd = -1
    if direct is not False and hasattr(os, "O_DIRECT"):
        try:
            dfd = os.open(tmp, os.O_WRONLY | os.O_DIRECT)
        except OSError:
            dfd = -1
    off = 0
    try:
        for chunk in chunks:
            mv = memoryview(chunk).cast("B")
            n = mv.nbytes
            if observer is not None:
                observer(off, mv)
            bulk = n - n % DIRECT_ALIGN
            if dfd >= 0 and bulk and off % DIRECT_ALIGN == 0 and _buffer_address(mv) % DIRECT_ALIGN == 0:
                try:
                    _pwrite_all(dfd, mv[:bulk], off)
                except OSError:
                    os.close(dfd)
                    dfd, bulk = -1, 0
                if bulk < n:
                    _pwrite_all(fd, mv[bulk:], off + bulk)
            else:
                _pwrite_all(fd, mv, off)
            off += n
        if epilogue is not None:
            off = epilogue(fd, off)
        if fsync:
            os.fsync(fd)
    except BaseException:
        for d in (fd, dfd):
            if d >= 0:
                os.close(d)
        try:
            os.remove(tmp)
        except OSError:
            pass
        raise
    os.close(fd)
    if dfd >= 0:
        os.close(dfd)
    os.replace(tmp, path)
    return off


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
