"""Native writer (and fast reader) of torch.distributed.checkpoint directories, fed from the pinned arena.

The reference's sharded checkpoint is `dist_cp.save(state_dict, checkpoint_id, FileSystemWriter)`
(pyrecover/checkpoint.py:254-274): a directory holding `.metadata` (a pickled
torch.distributed.checkpoint.metadata.Metadata) and one `__<rank>_0.distcp` per rank, in which every
saved item is a self-contained `torch.save` archive at a recorded (offset, length).  Its loader
(:300-368, `dist_cp.load` + FileSystemReader) slices an item out of the file, `torch.load`s it and
narrows it to the requested box.

This module writes that format WITHOUT torch.distributed.checkpoint on the save path and without
re-replicating anything: rank r owns flat element ranges of the (p, m, v) arenas -- its reduce-scatter
slices -- and a flat range of a row-major tensor is at most 2 * ndim - 1 contiguous boxes ("chunks":
a ragged head, whole rows, a ragged tail).  DCP stores a plain tensor as a list of such chunks, each
written by whichever rank has it, so rank r frames every box it owns as a small torch archive around
bytes that the pack kernel drops straight at their final file offsets inside the staging image:

    __<r>_0.distcp = [archive of chunk 0][archive of chunk 1] ... [small items, rank 0 only]
    archive        = local headers + data.pkl (a tensor of the chunk's shape) + the chunk's bytes + directory

The image is copied D2H as a whole and streamed to the file; only the few hundred header bytes around
each payload are produced on the host.  `.metadata` is written by rank 0, which can lay out every
rank's file by itself (the layout is a pure function of the arena layout and the world size).

Keys are the reference's: "model.<name>" ("model.module.<name>" when the reference would have had a
DistributedDataParallel wrapper, i.e. world > 1), "optimizer.state.<i>.{step,exp_avg,exp_avg_sq}",
"optimizer.param_groups.<g>.<key>", "metadata.{epoch,step}", "lr_scheduler.<key>".  Additive keys under
"b200.*" carry what the fast loader needs (RNG state, PR-SUM64 per rank file).
"""
from __future__ import annotations

import io
import os
import pickle
import zlib
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
# Imported HERE, at start-up, and not inside the functions that need them: the first import of
# torch.distributed.checkpoint takes seconds (4.7 s from a cold page cache), and those functions run in rank 0's
# writer thread beside the training loop -- an import there holds the interpreter lock against the training thread of
# ONE rank for that long, and every other rank waits for it in the gradient exchange.
from torch.distributed.checkpoint._nested_dict import flatten_state_dict
from torch.distributed.checkpoint.filesystem import _StorageInfo
from torch.distributed.checkpoint.metadata import (BytesStorageMetadata, ChunkStorageMetadata, Metadata, MetadataIndex,
                                                   StorageMeta, TensorProperties, TensorStorageMetadata)

from . import ptfile as PT

ARCHIVE_STEM = "archive"          # the stem torch.save uses for a file object
BIG_ALIGN = 4096                  # archives with a large payload start on a 4096-byte file offset (O_DIRECT)
SMALL_ALIGN = 64


# ------------------------------------------------------------------------- boxes ----------
def flat_range_boxes(shape: Sequence[int], lo: int, hi: int) -> List[Tuple[Tuple[int, ...], Tuple[int, ...], int]]:
    """The flat element range [lo, hi) of a row-major tensor of `shape` as contiguous boxes, in order:
    [(offsets, sizes, first_flat_element)].  A box is contiguous in memory because it spans whole
    trailing dimensions, a run of one dimension, and single indices before it."""
    shape = tuple(int(d) for d in shape)
    out: List[Tuple[Tuple[int, ...], Tuple[int, ...], int]] = []
    if hi <= lo:
        return out
    if not shape:
        return [((), (), lo)]

    def rec(dim: int, prefix: Tuple[int, ...], base: int, a: int, b: int) -> None:
        # [a, b) is a range inside the sub-tensor shape[dim:] that starts at flat element `base`
        if dim == len(shape) - 1:
            out.append((prefix + (a,), (1,) * len(prefix) + (b - a,), base + a))
            return
        inner = 1
        for d in shape[dim + 1:]:
            inner *= d
        r0, r1 = a // inner, (b - 1) // inner          # first and last row touched
        if r0 == r1 and (a % inner or b % inner):       # inside one row: go one dimension down
            rec(dim + 1, prefix + (r0,), base + r0 * inner, a - r0 * inner, b - r0 * inner)
            return
        start = r0
        if a % inner:                                   # ragged head
            rec(dim + 1, prefix + (r0,), base + r0 * inner, a - r0 * inner, inner)
            start = r0 + 1
        full_end = b // inner                           # whole rows [start, full_end)
        if full_end > start:
            out.append((prefix + (start,) + (0,) * (len(shape) - dim - 1),
                        (1,) * len(prefix) + (full_end - start,) + shape[dim + 1:], base + start * inner))
        if b % inner:                                   # ragged tail
            rec(dim + 1, prefix + (full_end,), base + full_end * inner, 0, b - full_end * inner)

    rec(0, (), 0, lo, hi)
    return out


# ------------------------------------------------------------------------- framing --------
class ChunkArchive:
    """One item of a .distcp file: a torch archive holding a tensor of `sizes` whose bytes sit at
    `data_offset` (relative to the archive start).  CRC-32 fields are filled in by `headers(crc)`."""

    __slots__ = ("entries", "payload", "data_offset", "length", "cd_offset")

    def __init__(self, dtype: torch.dtype, sizes: Sequence[int], nbytes: int):
        numel = 1
        for d in sizes:
            numel *= int(d)
        stub = PT.TensorStub(PT.RecordRef("0", dtype, numel), 0, tuple(sizes))
        self.entries = [PT._Entry(f"{ARCHIVE_STEM}/data.pkl", PT.dumps_state(stub)),
                        PT._Entry(f"{ARCHIVE_STEM}/byteorder", b"little"),
                        PT._Entry(f"{ARCHIVE_STEM}/data/0", (nbytes, None)),
                        PT._Entry(f"{ARCHIVE_STEM}/version", b"3\n")]
        pos = 0
        for e in self.entries:
            pos = e.place(pos)
        self.payload = self.entries[2]
        self.data_offset = self.payload.data_offset
        self.cd_offset = pos
        for e in self.entries:                          # sizes of the directory do not depend on the CRCs
            e.crc = 0
        cd = b"".join(e.central_header() for e in self.entries)
        self.length = pos + len(cd) + len(PT._end_records(len(self.entries), pos, len(cd)))

    def pieces(self, payload_crc: int) -> List[Tuple[int, bytes]]:
        """(offset in the archive, bytes) of everything except the payload."""
        out = []
        for e in self.entries:
            e.crc = payload_crc if e is self.payload else (zlib.crc32(e.data) & 0xFFFFFFFF)
        for e in self.entries:
            out.append((e.header_offset, e.local_header()))
            if e.data is not None and e.size:
                out.append((e.data_offset, bytes(e.data)))
        cd = b"".join(e.central_header() for e in self.entries)
        out.append((self.cd_offset, cd + PT._end_records(len(self.entries), self.cd_offset, len(cd))))
        return out


def _align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


# ------------------------------------------------------------------------- plan -----------
_DTYPES = {"bfloat16": torch.bfloat16, "float32": torch.float32, "float16": torch.float16, "float64": torch.float64,
           "int64": torch.int64, "int32": torch.int32, "uint8": torch.uint8, "int16": torch.int16, "int8": torch.int8,
           "bool": torch.bool}


class Item:
    """One box of one tensor in one rank's file."""
    __slots__ = ("key", "offsets", "sizes", "dtype", "global_off", "nbytes", "file_off", "archive")

    def __init__(self, key, offsets, sizes, dtype, global_off, nbytes):
        self.key, self.offsets, self.sizes, self.dtype = key, tuple(offsets), tuple(sizes), dtype
        self.global_off, self.nbytes = int(global_off), int(nbytes)
        self.file_off = 0
        self.archive: Optional[ChunkArchive] = None

    @property
    def payload_off(self) -> int:            # where the box's bytes sit in the file (and in the staging image)
        return self.file_off + self.archive.data_offset


def tensor_table(layout: Dict[str, Any], module_prefix: str = "") -> List[Tuple[str, torch.dtype, Tuple[int, ...], int]]:
    """(DCP key, dtype, shape, byte offset in the full image) of every big tensor, from
    Snapshotter.tensor_layout(): the reference's flattened names (pyrecover/checkpoint.py:254-264)."""
    rows = []
    for name, ent in layout["model"].items():
        if ent is not None:
            rows.append((f"model.{module_prefix}{name}", _DTYPES[ent[0]], tuple(ent[2]), int(ent[1])))
    for idx, st in layout["optimizer"].items():
        for k in ("exp_avg", "exp_avg_sq"):
            if st.get(k) is not None:
                ent = st[k]
                rows.append((f"optimizer.state.{int(idx)}.{k}", _DTYPES[ent[0]], tuple(ent[2]), int(ent[1])))
    rows.sort(key=lambda r: r[3])
    return rows


def plan_rank(table, owned_runs: Sequence[Tuple[int, int]]) -> Tuple[List[Item], int]:
    """Items of one rank's file and the size of its tensor part.  `owned_runs`: (byte offset in the full
    image, nbytes) runs the rank owns (Snapshotter.shard_pieces), any order."""
    runs = sorted((int(o), int(o) + int(n)) for o, n in owned_runs if n)
    items: List[Item] = []
    for key, dtype, shape, off in table:
        esize = torch.empty((), dtype=dtype).element_size()
        numel = 1
        for d in shape:
            numel *= d
        end = off + numel * esize
        for a, b in runs:
            lo, hi = max(a, off), min(b, end)
            if hi <= lo:
                continue
            if (lo - off) % esize or (hi - off) % esize:
                raise RuntimeError(f"owned run [{a}, {b}) cuts an element of {key}")
            for offsets, sizes, first in flat_range_boxes(shape, (lo - off) // esize, (hi - off) // esize):
                n = 1
                for d in sizes:
                    n *= d
                items.append(Item(key, offsets, sizes, dtype, off + first * esize, n * esize))
    cursor = 0
    for it in items:
        it.archive = ChunkArchive(it.dtype, it.sizes, it.nbytes)
        it.file_off = _align(cursor, BIG_ALIGN if it.nbytes >= PT._BIG_RECORD else SMALL_ALIGN)
        cursor = it.file_off + it.archive.length
    return items, _align(cursor, SMALL_ALIGN)


def rank_file(rank: int) -> str:
    return f"__{rank}_0.distcp"


# ------------------------------------------------------------------------- small items ----
def small_state_items(small_state: Dict[str, Any], n_params: int) -> Tuple[Dict[str, Any], Dict[str, Any]]:
    """The python-side part of the reference's state (everything except the big tensors) as DCP's flat
    items: ({flat key: object or small tensor}, {flat key: path}).  `small_state` carries epoch, step,
    param_groups (with "params" index lists), opt_steps, lr_scheduler, sampler_state -- what
    pyrecover/checkpoint.py:254-264 hands to dist_cp.save besides the tensors."""
    groups = small_state["param_groups"]
    steps = small_state.get("opt_steps", [0] * len(groups))
    opt_state = {}
    for gi, g in enumerate(groups):
        for idx in g["params"]:
            opt_state[idx] = {"step": torch.tensor(float(steps[gi]), dtype=torch.float32)}
    nested: Dict[str, Any] = {"optimizer": {"state": opt_state, "param_groups": groups},
                              "metadata": {"epoch": small_state.get("epoch"), "step": small_state.get("step")}}
    if small_state.get("lr_scheduler") is not None:
        nested["lr_scheduler"] = small_state["lr_scheduler"]
    if small_state.get("sampler_state") is not None:
        nested["sampler"] = small_state["sampler_state"]
    flat, mapping = flatten_state_dict(nested)
    return flat, mapping


def tensor_paths(table) -> Dict[str, Tuple]:
    """planner_data entries of the big tensors: how the flat key nests in the reference's state dict."""
    out = {}
    for key, _, _, _ in table:
        if key.startswith("model."):
            out[key] = ("model", key[len("model."):])
        else:
            _, _, idx, leaf = key.split(".")
            out[key] = ("optimizer", "state", idx, leaf)       # DCP spells integer dict keys as strings
    return out


# ------------------------------------------------------------------------- .metadata ------
def build_metadata(table, plans: Dict[int, List[Item]], small_flat: Dict[str, Any], small_paths: Dict[str, Any],
                   small_storage: Dict[str, Tuple[str, int, int]]):
    """torch.distributed.checkpoint's Metadata for the directory: every tensor with the boxes all ranks
    wrote, every small item, and where each lives (file, offset, length)."""
    sdm: Dict[str, Any] = {}
    storage: Dict[Any, Any] = {}
    boxes: Dict[str, List[Tuple[Item, int]]] = {}
    for rank, items in plans.items():
        for it in items:
            boxes.setdefault(it.key, []).append((it, rank))
    for key, dtype, shape, _ in table:
        mine = sorted(boxes.get(key, []), key=lambda x: x[0].global_off)
        chunks = []
        for ci, (it, rank) in enumerate(mine):
            chunks.append(ChunkStorageMetadata(offsets=torch.Size(it.offsets), sizes=torch.Size(it.sizes)))
            storage[MetadataIndex(key, it.offsets, ci)] = _StorageInfo(rank_file(rank), it.file_off, it.archive.length)
        sdm[key] = TensorStorageMetadata(properties=TensorProperties(dtype=dtype), size=torch.Size(shape), chunks=chunks)
    for key, val in small_flat.items():
        path, off, length = small_storage[key]
        if torch.is_tensor(val):
            sdm[key] = TensorStorageMetadata(properties=TensorProperties(dtype=val.dtype), size=val.size(),
                                             chunks=[ChunkStorageMetadata(offsets=torch.Size([0] * val.dim()), sizes=val.size())])
            storage[MetadataIndex(key, [0] * val.dim(), 0)] = _StorageInfo(path, off, length)
        else:
            sdm[key] = BytesStorageMetadata()
            storage[MetadataIndex(key)] = _StorageInfo(path, off, length)
    planner = dict(tensor_paths(table))
    planner.update(small_paths)
    return Metadata(state_dict_metadata=sdm, planner_data=planner, storage_data=storage, storage_meta=StorageMeta())


# ------------------------------------------------------------------------- writing --------
class _ItemCrc:
    """CRC-32 of every item's payload, chained in file order on a helper thread while the writer thread
    keeps the disk busy (zlib releases the GIL).  PYRECOVER_B200_DCP_CRC=0 writes zeros instead:
    torch.load does not check the archive's CRC fields, other zip tools do."""

    def __init__(self, items: List[Item]):
        import queue
        import threading
        self.items = items
        self.crc = [0] * len(items)
        self.enabled = os.environ.get("PYRECOVER_B200_DCP_CRC", "1") != "0"
        self.cursor = 0
        self.err: Optional[BaseException] = None
        self.q: "queue.Queue" = queue.Queue()
        self.t = threading.Thread(target=self._run, name="pyrecover-b200-crc", daemon=True)
        if self.enabled:
            self.t.start()

    def _run(self):
        while True:
            job = self.q.get()
            if job is None:
                return
            i, mv = job
            try:
                self.crc[i] = zlib.crc32(mv, self.crc[i])
            except BaseException as e:      # noqa: BLE001 -- surfaced by finish()
                self.err = e

    def observe(self, off: int, mv: memoryview) -> None:
        """A chunk [off, off + len) of the file image is on its way out: feed the payload parts it holds."""
        if not self.enabled:
            return
        end = off + mv.nbytes
        i = self.cursor
        while i < len(self.items):
            it = self.items[i]
            a, b = it.payload_off, it.payload_off + it.nbytes
            if a >= end:
                break
            lo, hi = max(a, off), min(b, end)
            if hi > lo:
                self.q.put((i, mv[lo - off:hi - off]))
            if b <= end:
                self.cursor = i + 1
            i += 1

    def finish(self) -> List[int]:
        if self.enabled:
            self.q.put(None)
            self.t.join()
            if self.err is not None:
                raise self.err
        return [c & 0xFFFFFFFF for c in self.crc]


def write_rank_file(path: str, items: List[Item], image_bytes: int, chunks, small: Optional[Dict[str, Any]] = None,
                    fsync: bool = True) -> Dict[str, Tuple[str, int, int]]:
    """Write one rank's .distcp: `chunks` yields the staging image [0, image_bytes) in order (payloads
    already at their offsets; the bytes between them are whatever the arena held), then the archive framing
    is written around every payload, then -- rank 0 -- the small items are appended, each a torch.save
    blob.  Atomic (tmp + rename).  Returns {small key: (file name, offset, length)}."""
    crc = _ItemCrc(items)
    placed: Dict[str, Tuple[str, int, int]] = {}
    name = os.path.basename(path)

    def epilogue(fd: int, end: int) -> int:
        if end != image_bytes:
            raise IOError(f"{path}: streamed {end} bytes of a {image_bytes}-byte image")
        crcs = crc.finish()
        for it, c in zip(items, crcs):
            for off, blob in it.archive.pieces(c):
                PT._pwrite_all(fd, memoryview(blob), it.file_off + off)
        cur = end
        blobs: Dict[Any, Tuple[int, int]] = {}          # identical small items share one blob (147 equal step scalars)
        for key, obj in (small or {}).items():
            ident = ("t", obj.dtype, tuple(obj.shape), tuple(obj.reshape(-1).tolist())) if torch.is_tensor(obj) and obj.numel() <= 4 \
                else None
            if ident is not None and ident in blobs:
                placed[key] = (name,) + blobs[ident]
                continue
            buf = io.BytesIO()
            torch.save(obj, buf)
            blob = buf.getvalue()
            PT._pwrite_all(fd, memoryview(blob), cur)
            placed[key] = (name, cur, len(blob))
            if ident is not None:
                blobs[ident] = (cur, len(blob))
            cur = _align(cur + len(blob), 8)
        return cur

    PT.write_raw(path, chunks, fsync=fsync, observer=crc.observe, epilogue=epilogue)
    return placed


def write_metadata(dirpath: str, metadata) -> None:
    """`metadata`: a Metadata object or its pickled bytes (the caller may cache them: nothing in it changes from
    save to save as long as the small items keep their sizes)."""
    tmp = os.path.join(dirpath, ".metadata.tmp")
    with open(tmp, "wb") as f:
        f.write(metadata if isinstance(metadata, (bytes, bytearray)) else pickle.dumps(metadata))
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, os.path.join(dirpath, ".metadata"))
