"""Checkpoint snapshot engine: device-side pack + checksum -> pinned host arena -> writer thread.

Replaces the body of the reference's save path (pyrecover/checkpoint.py:58-84: state_dict ->
torch.save with one synchronous pageable D2H per storage -> re-read of the whole file for md5,
all on rank 0 while training is stopped; and :254-274, torch.distributed.checkpoint) and of its
load path (:182-199, :352-356).

Timeline of one save:

    training stream : ... optimizer.step | next forward + backward ...      | wait E1 | optimizer.step
    side stream     :          wait | K3 pack+checksum (HBM->HBM, ~ms) E1 | D2H image -> pinned arena E2
    writer thread   :                                             wait E2 | file write, fsync, rename, ...

`save_*` returns once the pack kernel and the copies are ENQUEUED on the side stream: the
snapshot is consistent (it is ordered after this step's optimizer update, and the NEXT optimizer
update is fenced on E1 -- forward and backward only read parameters), the training stream never
waits for it in practice, and durability is awaited by the next save / load / `wait()`.

Two images exist per (model, optimizer): the FULL image (vanilla .pt file: p | m | v arenas +
stray tensors, one record per dtype) and the SHARD image (sharded directory format: this rank's
1/W element slice of p, m and v).  Each owns its pack plan, HBM staging buffer and pinned arena,
created on first use.
"""
from __future__ import annotations

import copy
import json
import os
import re
import threading
import time
from collections import OrderedDict
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from . import dcpfile as DCP
from . import kernels as K
from . import ptfile as PT

ALIGN = 256
COPY_CHUNK = 64 << 20
_M64 = (1 << 64) - 1
SHARD_FORMAT = 2        # 2: torch.distributed.checkpoint directory written natively (dcpfile.py); 1: raw shard_<r>.bin (still read)
# Grid of the bulk-copy pack kernel.  It runs on a side stream beside the next forward + backward and only has
# to be done before the next optimizer step, so it is spread thin: measured on B200 (1 B model, 6 GB image,
# 25 ms step) the stall of the training stream is 1.2 ms with 8 CTAs (the pack then takes 18 ms), 2.8 ms with
# 16-32 and 3.5 ms with 64-148 -- a few CTAs ride in the SMs the GEMM waves leave idle, more displace them.
# One CTA moves ~85 GB/s (read + write).  PYRECOVER_B200_PACK_BLOCKS fixes the grid; otherwise it is sized so
# that the pack takes PACK_STEP_FRACTION of the step period the optimizer observes (15 ms when unknown).
PACK_BLOCKS = int(os.environ.get("PYRECOVER_B200_PACK_BLOCKS", "0"))
PACK_STEP_FRACTION = 0.7
PACK_CTA_GBS = 85.0


def pack_blocks_for(image_bytes: int, step_period_s: Optional[float]) -> int:
    if PACK_BLOCKS > 0:
        return PACK_BLOCKS
    target = PACK_STEP_FRACTION * step_period_s if step_period_s else 0.015
    blocks = int(2.0 * image_bytes / (PACK_CTA_GBS * 1e9 * max(target, 1e-4)) + 0.999)
    return max(4, min(148, blocks))


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class Piece:
    """A contiguous run of device bytes and where it sits in an image."""
    __slots__ = ("tensor", "img_off", "nbytes", "record", "rec_off", "global_off")

    def __init__(self, tensor: torch.Tensor, img_off: int, record: str = "", rec_off: int = 0,
                 global_off: int = 0):
        self.tensor = tensor
        self.img_off = img_off
        self.nbytes = tensor.numel() * tensor.element_size()
        self.record, self.rec_off, self.global_off = record, rec_off, global_off


class Image:
    """pack plan + staging + pinned arena + checksum words for one list of pieces."""

    def __init__(self, pieces: Sequence[Piece], image_bytes: int, device: torch.device, staged: bool):
        self.pieces = list(pieces)
        self.image_bytes = max(int(image_bytes), ALIGN)
        self.device = device
        self.plan = K.PackPlan([p.tensor for p in self.pieces], [p.img_off for p in self.pieces], self.image_bytes)
        self.staging = torch.zeros(self.image_bytes, dtype=torch.uint8, device=device) if staged else None
        self.host = K.HostArena(self.image_bytes)
        self.sum_dev = torch.zeros(2, dtype=torch.int64, device=device)
        self.sum_host = torch.zeros(2, dtype=torch.int64).pin_memory()
        self.e_copied: Optional[torch.cuda.Event] = None

    def stage(self, side: torch.cuda.Stream, step_period_s: Optional[float] = None):
        """Runs on `side`, ordered after the work already queued on the training stream: pack
        (+PR-SUM64) into the staging image.  Returns e_read: once it has fired the live tensors
        may be modified again (the caller fences the next optimizer step on it -- forward and
        backward never write them, so the training stream itself does not wait).

        Staged mode does NOT enqueue the D2H here: the writer thread drains the staging image
        chunk by chunk at the pace of the file write (`drain`).  A back-to-back 6 GB D2H would
        saturate PCIe for ~110 ms and, measured on B200, slow the training stream by half of that
        (kernel launches travel over the same link); the disk is 40x slower than PCIe anyway."""
        main = torch.cuda.current_stream(self.device)
        e_ready = torch.cuda.Event()
        e_ready.record(main)
        side.wait_event(e_ready)
        with torch.cuda.stream(side):
            self.sum_dev.zero_()
        e_read = torch.cuda.Event()
        if self.staging is not None:
            # the bulk-copy (TMA-staged) pack kernel on a bounded number of CTAs (pack_blocks_for)
            torch.cuda.nvtx.range_push("ckpt_pack")
            self.pack_blocks = pack_blocks_for(self.image_bytes, step_period_s)
            self.plan.pack(self.staging.data_ptr(), self.sum_dev, stream=side, max_blocks=self.pack_blocks)
            torch.cuda.nvtx.range_pop()
            e_read.record(side)
        else:                                    # direct: copy engines read the live tensors, all at once
            for p in self.pieces:
                if p.nbytes:
                    K.checksum_into(p.tensor.reshape(-1), self.sum_dev, p.img_off // 4, stream=side)
                    K.copy_d2h_async(self.host.ptr + p.img_off, p.tensor, p.nbytes, COPY_CHUNK, stream=side)
            e_read.record(side)
        with torch.cuda.stream(side):
            self.sum_host.copy_(self.sum_dev, non_blocking=True)
        e = torch.cuda.Event()
        e.record(side)
        self.e_copied = e            # direct: whole image on the host; staged: checksum on the host
        return e_read, e

    def drain(self, side: torch.cuda.Stream, lo: int = 0, hi: Optional[int] = None):
        """Generator over memoryviews of the pinned arena covering image bytes [lo, hi), in
        order.  Staged mode: each chunk is copied D2H on `side` just before it is handed out, with
        the NEXT chunk's copy already in flight (the consumer's file write overlaps it).  Called
        from the writer thread."""
        hi = self.image_bytes if hi is None else hi
        if self.staging is None:                 # direct mode: stage() already copied everything
            self.e_copied.synchronize()
            for o in range(lo, hi, COPY_CHUNK):
                yield self.host.view(o, min(COPY_CHUNK, hi - o))
            return
        torch.cuda.set_device(self.device)
        self.e_copied.synchronize()              # pack finished
        spans = [(o, min(COPY_CHUNK, hi - o)) for o in range(lo, hi, COPY_CHUNK)]

        def issue(span):
            o, n = span
            torch.cuda.nvtx.range_push("ckpt_d2h")
            K.copy_d2h_async(self.host.ptr + o, self.staging[o:o + n], n, 0, stream=side)
            torch.cuda.nvtx.range_pop()
            ev = torch.cuda.Event()
            ev.record(side)
            return ev
        pending = issue(spans[0]) if spans else None
        for i, (o, n) in enumerate(spans):
            ev = pending
            pending = issue(spans[i + 1]) if i + 1 < len(spans) else None
            ev.synchronize()
            yield self.host.view(o, n)

    def checksum(self) -> Tuple[int, int]:
        a, b = self.sum_host.tolist()
        return a & _M64, b & _M64

    def restore(self) -> Tuple[int, int]:
        """pinned arena -> HBM -> unpack into the live tensors; returns PR-SUM64 of the image."""
        main = torch.cuda.current_stream(self.device)
        image = self.staging if self.staging is not None else \
            torch.empty(self.image_bytes, dtype=torch.uint8, device=self.device)
        K.copy_h2d_async(image, self.host.ptr, self.image_bytes, COPY_CHUNK, stream=main)
        self.sum_dev.zero_()
        self.plan.unpack(image.data_ptr(), self.sum_dev, stream=main)
        self._last_image = image
        return K._sum_to_ints(self.sum_dev)

    def close(self) -> None:
        self.plan.close()
        self.host.close()


class Snapshotter:
    """Checkpoint engine for one (model, B200AdamW) pair; layouts are static for the job."""

    def __init__(self, model: torch.nn.Module, optimizer, mode: Optional[str] = None):
        self.module = model.module if hasattr(model, "module") else model
        self.optimizer = optimizer
        self.mode = mode or os.environ.get("PYRECOVER_B200_SNAPSHOT", "staged")
        if self.mode not in ("staged", "direct"):
            raise ValueError("snapshot mode must be 'staged' or 'direct'")
        self.arenas = optimizer.arenas
        self.device = self.arenas[0].device
        self.side = torch.cuda.Stream(device=self.device)
        self._full: Optional[Image] = None
        self._shard: Optional[Image] = None
        self._shard_key = None
        self._writer: Optional[threading.Thread] = None
        self._writer_error: Optional[BaseException] = None
        self._stub_cache: Dict[Any, PT.TensorStub] = {}
        self._layout: Optional[Dict[str, Any]] = None
        self._skeleton: Optional[Dict[str, Any]] = None
        self._dcp_plans: Dict[Any, Any] = {}
        self._dcp_metadata: Optional[Tuple[Any, bytes]] = None
        self.last_stats: Dict[str, Any] = {}
        # The writer thread runs Python between its (GIL-free) copies and writes.  With CPython's default 5 ms
        # switch interval a training thread that wants the interpreter back can wait that long each time -- in a
        # loop that is launch-bound on the host that is GPU idle time (measured: ~4 ms per checkpoint at 4-8 GPUs).
        import sys
        if sys.getswitchinterval() > 2e-4:
            sys.setswitchinterval(2e-4)
        self.durable_s: Optional[float] = None      # save_* call -> file renamed, of the last FINISHED write
        self._durable_reported = True
        # ---- full image layout: one record per dtype; arenas first (p | m | v), extras after ----
        by_dtype: "OrderedDict[torch.dtype, List[torch.Tensor]]" = OrderedDict()
        for a in self.arenas:
            by_dtype.setdefault(a.dtype, []).extend([a.p, a.m, a.v])
        self._arena_ranges = [(t.data_ptr(), t.data_ptr() + t.numel() * t.element_size())
                              for a in self.arenas for t in (a.p, a.m, a.v)]
        self.extras: List[torch.Tensor] = []
        for name, t in self.module.state_dict().items():
            # the image is packed from, and restored into, the LIVE tensors: a private contiguous or
            # device copy made here would go stale after the first optimizer step
            if not t.is_cuda:
                raise RuntimeError(f"pyrecover_b200 snapshot: state_dict entry '{name}' lives on {t.device}; every "
                                   "entry must be a CUDA tensor (keep host-side state out of state_dict or make "
                                   "the buffer non-persistent)")
            if not t.is_contiguous():
                raise RuntimeError(f"pyrecover_b200 snapshot: state_dict entry '{name}' is not contiguous")
            if not self._in_arena(t):
                by_dtype.setdefault(t.dtype, []).append(t)
                self.extras.append(t)
        self.records: "OrderedDict[str, dict]" = OrderedDict()
        self.full_pieces: List[Piece] = []
        img_off = 0
        for ri, (dtype, tensors) in enumerate(by_dtype.items()):
            key, rec_off, start = str(ri), 0, img_off
            for t in tensors:
                self.full_pieces.append(Piece(t, start + rec_off, key, rec_off, start + rec_off))
                rec_off += _round_up(t.numel() * t.element_size(), ALIGN)
            self.records[key] = {"dtype": dtype, "off": start, "nbytes": rec_off}
            img_off = start + rec_off
        self.full_bytes = max(img_off, ALIGN)

    # ------------------------------------------------------------------ layout helpers ----
    def _in_arena(self, t: torch.Tensor) -> bool:
        p = t.data_ptr()
        return any(lo <= p < hi for lo, hi in self._arena_ranges)

    def full(self) -> Image:
        if self._full is None:
            self._full = Image(self.full_pieces, self.full_bytes, self.device, self.mode == "staged")
        return self._full

    def shard_pieces(self, rank: int, world: int) -> List[Tuple[int, int]]:
        """(global image offset, nbytes) runs that rank `rank` of `world` owns: its element ranges
        of every arena tensor -- the reduce-scatter slices of K1 (GradSync.owned_ranges: one slice of
        the arena, or one per bucket in "overlap" mode), 16-byte granular -- so a sharded save writes
        the optimizer state where it already lives; rank 0 also owns the stray tensors."""
        gs = getattr(self.optimizer, "_grad_sync", None)
        runs = []
        for p in self.full_pieces:
            if self._in_arena(p.tensor):
                esize = p.tensor.element_size()
                if gs is not None and gs.world == world and gs.arena is not None and p.nbytes == gs.arena.layout.nbytes:
                    ranges = [(lo * esize, hi * esize) for lo, hi in gs.owned_ranges(rank, world)]
                else:
                    nvec = (p.nbytes + 15) // 16
                    per = (nvec + world - 1) // world
                    ranges = [(min(per * rank, nvec) * 16, min(min(per * (rank + 1), nvec) * 16, p.nbytes))]
                for lo, hi in ranges:
                    if hi > lo:
                        runs.append((p.global_off + lo, hi - lo))
            elif rank == 0 and p.nbytes:
                runs.append((p.global_off, p.nbytes))
        return runs

    def dcp_plan(self, rank: int, world: int):
        """(items, tensor-part bytes) of rank `rank`'s `__<rank>_0.distcp`: every box of every tensor the rank
        owns, framed as a torch archive at a fixed file offset (dcpfile.plan_rank).  Static for the job."""
        key = (rank, world)
        if key not in self._dcp_plans:
            # the reference's keys: its model is DDP-wrapped whenever world > 1 ("model.module.<name>")
            table = DCP.tensor_table(self.tensor_layout(), "module." if world > 1 else "")
            self._dcp_plans[key] = DCP.plan_rank(table, self.shard_pieces(rank, world))
        return self._dcp_plans[key]

    def shard(self, rank: int, world: int) -> Image:
        """This rank's image = the tensor part of its .distcp file: the pack kernel drops every owned box at
        the offset its bytes have in the file; the framing around them is written by the host."""
        if self._shard is None or self._shard_key != (rank, world):
            if self._shard is not None:
                self._shard.close()
            items, nbytes = self.dcp_plan(rank, world)
            flats = [(p.global_off, p.nbytes, p.tensor.reshape(-1).view(torch.uint8)) for p in self.full_pieces if p.nbytes]
            pieces, k = [], 0
            for it in items:                       # items and full_pieces are both ordered by image offset
                while not (flats[k][0] <= it.global_off < flats[k][0] + flats[k][1]):
                    k += 1
                base, _, flat = flats[k]
                lo = it.global_off - base
                pieces.append(Piece(flat[lo:lo + it.nbytes], it.payload_off, global_off=it.global_off))
            self._shard = Image(pieces, nbytes, self.device, self.mode == "staged")
            self._shard_key = (rank, world)
            if rank == 0:                          # rank 0 lays out every rank's file for .metadata: do it now, not in the
                for r in range(world):             # writer thread beside the training loop
                    self.dcp_plan(r, world)
        return self._shard

    def _locate(self, t: torch.Tensor) -> Optional[Tuple[str, int]]:
        p = t.data_ptr()
        for s in self.full_pieces:
            base = s.tensor.data_ptr()
            if base <= p < base + max(s.nbytes, 1) and t.dtype == s.tensor.dtype:
                return s.record, (s.rec_off + (p - base)) // t.element_size()
        return None

    _ATOMS = (int, float, str, bool, type(None), bytes, complex)

    def stubify(self, obj: Any, small: Dict[torch.dtype, list]) -> Any:
        """Replace every tensor of a nested state by a TensorStub.  CUDA tensors inside a piece
        point into the big records; anything else lands in a small per-dtype host record (step
        scalars, RNG state, ...).

        This runs on the TRAINING thread inside save_*: a training loop that is close to
        launch-bound stalls the GPU by however long it takes, so arena views are resolved through a
        cache (the layout is static for the job) and the rest is kept allocation-light."""
        if isinstance(obj, torch.Tensor):
            if obj.is_cuda:
                key = (obj.data_ptr(), obj.dtype, obj.shape, obj.stride())
                stub = self._stub_cache.get(key)
                if stub is not None:
                    return stub
                loc = self._locate(obj) if obj.is_contiguous() else None
                if loc is not None:
                    rec = PT.RecordRef(loc[0], obj.dtype, self.records[loc[0]]["nbytes"] // obj.element_size())
                    stub = PT.TensorStub(rec, loc[1], obj.shape)
                    self._stub_cache[key] = stub
                    return stub
            host = obj.detach().reshape(-1).to("cpu", copy=True)      # strays: a private host copy
            bucket = small.get(host.dtype)
            if bucket is None:
                bucket = small[host.dtype] = [[], 0]
            off = bucket[1]
            bucket[0].append(host)
            bucket[1] = off + host.numel()
            return _Small(host.dtype, off, tuple(obj.shape))
        if type(obj) in self._ATOMS:
            return obj
        if isinstance(obj, dict):
            return type(obj)((k, self.stubify(v, small)) for k, v in obj.items())
        if isinstance(obj, list):
            return [self.stubify(v, small) for v in obj]
        if isinstance(obj, tuple):
            return tuple(self.stubify(v, small) for v in obj)
        return copy.deepcopy(obj)

    def tensor_layout(self) -> Dict[str, Any]:
        """Where every tensor of model.state_dict() and of the optimizer state sits in the full image:
        {"model": {name: [dtype, byte offset, shape]}, "optimizer": {index: {"exp_avg": [...],
        "exp_avg_sq": [...]}}}.  Static for the job; stored in a sharded directory's meta.pt so that
        the directory can be read back WITHOUT this library or a GPU (pyrecover/convert.py)."""
        if self._layout is None:
            def where(t: torch.Tensor):
                loc = self._locate(t) if t.is_cuda and t.is_contiguous() else None
                if loc is None:
                    return None
                rec = self.records[loc[0]]
                return [str(t.dtype).replace("torch.", ""), rec["off"] + loc[1] * t.element_size(), list(t.shape)]
            model = {k: where(t) for k, t in self.module.state_dict().items()}
            opt: Dict[int, Any] = {}
            idx = 0
            for group in self.optimizer.param_groups:
                for q in group["params"]:
                    st = self.optimizer.state.get(q, {})
                    opt[idx] = {k: where(st[k]) for k in ("exp_avg", "exp_avg_sq") if k in st}
                    idx += 1
            self._layout = {"model": model, "optimizer": opt, "image_bytes": self.full_bytes}
        return self._layout

    # ------------------------------------------------------------------ writer plumbing ---
    def wait(self) -> None:
        """Join the writer of the previous save (durability point); re-raise its failure."""
        if self._writer is not None:
            self._writer.join()
            self._writer = None
        if self._writer_error is not None:
            err, self._writer_error = self._writer_error, None
            raise RuntimeError(f"checkpoint writer failed: {err!r}") from err

    def _spawn(self, work) -> None:
        def run():
            try:
                work()
            except BaseException as e:       # surfaced by wait()
                self._writer_error = e
        self._writer = threading.Thread(target=run, name="pyrecover-b200-writer")
        self._writer.start()

    def _fence_optimizer(self, e: torch.cuda.Event) -> None:
        # the live arenas must not change before the snapshot has read them: the next
        # optimizer.step() waits for this event (a whole forward+backward later)
        self.optimizer._snapshot_fence = e

    def take_durable_time(self) -> Optional[float]:
        """Seconds from the save_* call to the renamed file of the most recent write that has FINISHED,
        once per write (None when nothing new finished).  Never blocks: the time-aware policy feeds on
        it (reference train.py:334-337 measures the same span around its synchronous save)."""
        if self._durable_reported or (self._writer is not None and self._writer.is_alive()):
            return None
        self._durable_reported = True
        return self.durable_s

    def _check_exchange(self) -> None:
        """A gradient exchange whose barrier timed out leaves this rank's parameters untouched but
        the job broken; nothing may be published as a checkpoint from then on."""
        gs = getattr(self.optimizer, "_grad_sync", None)
        if gs is not None and gs.comm.timed_out():
            raise RuntimeError("refusing to write a checkpoint: the gradient exchange timed out waiting for a peer")

    def skeleton(self) -> Dict[str, Any]:
        """The part of the saved state that never changes during the job, stubified once: every tensor of
        model.state_dict() and the exp_avg / exp_avg_sq of every parameter, as views into the image
        records.  A save then costs the training thread only the handful of python scalars that do
        change (step counts, lr, scheduler counters, RNG state) -- in a loop that is launch-bound on the
        host, whatever save_* spends on the training thread is GPU idle time."""
        if self._skeleton is None:
            small: Dict[torch.dtype, list] = {}
            model = self.stubify(self.module.state_dict(), small)
            opt, index_of, idx = {}, {}, 0
            for group in self.optimizer.param_groups:
                for q in group["params"]:
                    st = self.optimizer.state[q]
                    opt[idx] = {"exp_avg": self.stubify(st["exp_avg"], small),
                                "exp_avg_sq": self.stubify(st["exp_avg_sq"], small)}
                    index_of[id(q)] = idx
                    idx += 1
            if small:
                raise RuntimeError("pyrecover_b200 snapshot: a state tensor lies outside the packed image")
            self._skeleton = {"model": model, "opt": opt, "index_of": index_of, "n_params": idx}
        return self._skeleton

    def dynamic_state(self, lr_scheduler, sampler, step, epoch, rng_state) -> Dict[str, Any]:
        """What changes from save to save, copied on the training thread (a few dozen python objects)."""
        sk = self.skeleton()
        groups, gsteps = [], []
        for gi, group in enumerate(self.optimizer.param_groups):
            g = {k: v for k, v in group.items() if k != "params"}
            g["params"] = [sk["index_of"][id(q)] for q in group["params"]]
            groups.append(g)
            gsteps.append(float(self.optimizer._steps[gi]))
        dyn = {"epoch": epoch, "step": step, "param_groups": copy.deepcopy(groups), "group_steps": gsteps,
               "rng_state": rng_state}
        if lr_scheduler is not None:
            dyn["lr_scheduler"] = copy.deepcopy(lr_scheduler.state_dict())
        if sampler and hasattr(sampler, "set_state"):
            dyn["sampler_state"] = copy.deepcopy(sampler.state_dict())
        return dyn

    def _assemble(self, dyn: Dict[str, Any], small: Dict[torch.dtype, list]) -> Dict[str, Any]:
        """Writer thread: skeleton + dynamic part -> the reference's state dict (same keys, same order as
        pyrecover/checkpoint.py:60-73; optimizer entry as torch.optim.Optimizer.state_dict() lays it out),
        tensors as stubs.  Every parameter gets its own step scalar, as torch.optim.AdamW hands them out."""
        sk = self.skeleton()
        state: Dict[str, Any] = {"epoch": dyn["epoch"], "step": dyn["step"], "model": sk["model"]}
        opt_state = {}
        for gi, g in enumerate(dyn["param_groups"]):
            steps = torch.full((len(g["params"]),), dyn["group_steps"][gi], dtype=torch.float32)
            for k, idx in enumerate(g["params"]):
                opt_state[idx] = {"step": self.stubify(steps[k], small), **sk["opt"][idx]}
        state["optimizer"] = {"state": {i: opt_state[i] for i in sorted(opt_state)}, "param_groups": dyn["param_groups"]}
        for key in ("lr_scheduler", "sampler_state", "rng_state"):
            if key in dyn:
                state[key] = self.stubify(dyn[key], small)
        return state

    # ------------------------------------------------------------------ vanilla (.pt) -----
    def save_vanilla(self, path: str, dyn: Dict[str, Any], verify: bool, max_keep: int,
                     blocking: bool = False) -> None:
        """`dyn` = dynamic_state(...): the python-side values of this save; tensor bytes are captured by
        the pack, and the state dict the file holds is put together in the writer thread."""
        t0 = time.perf_counter()
        self._check_exchange()
        self.wait()                          # the pinned arena is about to be overwritten
        img = self.full()
        e_read, e_copied = img.stage(self.side, getattr(self.optimizer, "step_period_s", None))
        self._fence_optimizer(e_read)
        path = str(path)
        self.last_stats = {"enqueue_s": time.perf_counter() - t0, "image_bytes": img.image_bytes,
                           "pack_blocks": getattr(img, "pack_blocks", None)}
        self._durable_reported = False

        def work():
            t1 = time.perf_counter()
            e_copied.synchronize()
            self._check_exchange()           # everything up to the pack has run: a timed-out exchange shows by now
            t2 = time.perf_counter()
            small: Dict[torch.dtype, list] = {}
            frozen = self._assemble(dyn, small)
            small_bufs = {dt: (torch.cat(parts) if len(parts) > 1 else parts[0]) for dt, (parts, _) in small.items()}
            records = [(key, (r["nbytes"], img.drain(self.side, r["off"], r["off"] + r["nbytes"])))
                       for key, r in self.records.items()]
            refs = {}
            for i, (dt, buf) in enumerate(small_bufs.items()):
                key = f"s{i}"
                refs[dt] = PT.RecordRef(key, dt, buf.numel())
                records.append((key, buf.numpy().view("uint8").data if buf.numel() else b""))
            final = _finish_small(frozen, refs)
            A, B = img.checksum()
            final["b200"] = {"format": 1, "sum64": [A, B], "image_bytes": img.image_bytes,
                             "records": {k: [str(r["dtype"]), r["off"], r["nbytes"]] for k, r in self.records.items()}}
            info = PT.write_archive(path, PT.dumps_state(final), records, md5=verify)
            if verify:
                with open(path + ".md5", "w") as f:      # reference side-car, checkpoint.py:81-83
                    f.write(info["md5"])
            with open(path + ".b200sum", "w") as f:
                json.dump({"sum64": [A, B], "image_bytes": img.image_bytes, "file_bytes": info["bytes"],
                           "record_offsets": info["offsets"]}, f)
            self.durable_s = time.perf_counter() - t0
            rotate_files(Path(path).parent, max_keep)
            self.last_stats.update(d2h_wait_s=t2 - t1, write_s=time.perf_counter() - t2, file_bytes=info["bytes"],
                                   durable_s=self.durable_s)

        self._spawn(work)
        if blocking:
            self.wait()

    def load_vanilla_image(self, path: str, meta: Dict[str, Any], verify: bool) -> None:
        """Fast restore of a file written by save_vanilla: the records are contiguous in the file
        in image order, so the whole image is read with large reads into the pinned arena."""
        self.wait()
        img = self.full()
        if int(meta["image_bytes"]) != img.image_bytes:
            raise RuntimeError("checkpoint image layout differs from the live model")
        offs = PT.record_offsets(path)
        mv = img.host.view()
        with open(path, "rb", buffering=0) as f:
            for key, r in self.records.items():
                off, size = offs[f"data/{key}"]
                if size != r["nbytes"]:
                    raise RuntimeError(f"record {key}: {size} bytes on disk, {r['nbytes']} expected")
                f.seek(off)
                _read_exact(f, mv[r["off"]:r["off"] + size], path)
        got = img.restore()
        if verify and tuple(meta["sum64"]) != got:
            raise RuntimeError(f"Checksum mismatch for checkpoint {path}: stored {tuple(meta['sum64'])}, "
                               f"computed on GPU {got}")

    # ------------------------------------------------------------------ sharded directory -
    def save_sharded(self, dirpath: str, small_state: Dict[str, Any], rank: int, world: int,
                     max_keep: int, blocking: bool = False) -> None:
        """Every rank writes `__<rank>_0.distcp` -- its boxes of p, m, v as torch archives, straight from
        the pinned image -- and a side-car `b200_<rank>.json` (PR-SUM64, where each box lies); rank 0 adds the
        python-side items to its file, writes `b200_meta.pt` and, once every rank's side-car is there,
        `.metadata`: the directory is then a torch.distributed.checkpoint checkpoint that stock
        `dist_cp.load` and the reference's load_ckpt_distributed (pyrecover/checkpoint.py:300-368) open, and it
        is complete exactly when `.metadata` exists (as with dist_cp.save)."""
        t0 = time.perf_counter()
        self._check_exchange()
        self.wait()
        img = self.shard(rank, world)
        items, nbytes = self.dcp_plan(rank, world)
        e_read, e_copied = img.stage(self.side, getattr(self.optimizer, "step_period_s", None))
        self._fence_optimizer(e_read)
        self._durable_reported = False
        small_state = _to_cpu(small_state)
        n_params = sum(len(g["params"]) for g in self.optimizer.param_groups)
        if rank == 0:
            small_state["layout"] = self.tensor_layout()
            nxt = 0
            for g, live in zip(small_state["param_groups"], self.optimizer.param_groups):
                g["params"] = list(range(nxt, nxt + len(live["params"])))
                nxt += len(live["params"])
        d = Path(dirpath)
        self.last_stats = {"enqueue_s": time.perf_counter() - t0, "image_bytes": img.image_bytes,
                           "pack_blocks": getattr(img, "pack_blocks", None)}

        def work():
            t1 = time.perf_counter()
            d.mkdir(parents=True, exist_ok=True)
            e_copied.synchronize()
            self._check_exchange()
            t2 = time.perf_counter()
            small_flat = small_paths = None
            if rank == 0:
                small_flat, small_paths = DCP.small_state_items(small_state, n_params)
            placed = DCP.write_rank_file(str(d / DCP.rank_file(rank)), items, nbytes, img.drain(self.side), small_flat)
            A, B = img.checksum()
            side = {"format": SHARD_FORMAT, "rank": rank, "world": world, "sum64": [A, B], "step": small_state.get("step"),
                    "image_bytes": img.image_bytes, "full_image_bytes": self.full_bytes, "file": DCP.rank_file(rank),
                    "pieces": [[p.global_off, p.nbytes, p.img_off] for p in img.pieces]}
            if rank == 0:
                torch.save(small_state, str(d / "b200_meta.pt.tmp"))
                os.replace(str(d / "b200_meta.pt.tmp"), str(d / "b200_meta.pt"))
            else:
                torch.save({"rng_state": small_state.get("rng_state")}, str(d / f"rng_{rank:05d}.pt"))
            _write_text_atomic(d / f"b200_{rank:05d}.json", json.dumps(side))
            if rank == 0:
                _await_sidecars(d, world, small_state.get("step"))
                # .metadata only changes when a small item changes size: keep the pickled bytes (building ~1000
                # metadata objects every save would hold the interpreter lock against the training thread)
                sig = (world, tuple(sorted(placed.items())))
                if self._dcp_metadata is None or self._dcp_metadata[0] != sig:
                    plans = {r: self.dcp_plan(r, world)[0] for r in range(world)}
                    table = DCP.tensor_table(self.tensor_layout(), "module." if world > 1 else "")
                    import pickle
                    self._dcp_metadata = (sig, pickle.dumps(DCP.build_metadata(table, plans, small_flat, small_paths, placed)))
                DCP.write_metadata(str(d), self._dcp_metadata[1])
            self.durable_s = time.perf_counter() - t0
            if rank == 0:
                rotate_dirs(d.parent, max_keep)
            self.last_stats.update(d2h_wait_s=t2 - t1, write_s=time.perf_counter() - t2, file_bytes=img.image_bytes,
                                   durable_s=self.durable_s)

        self._spawn(work)
        if blocking:
            self.wait()

    def load_sharded(self, dirpath: str, verify: bool = True, rank: int = 0, world: int = 1,
                     grad_sync=None) -> Dict[str, Any]:
        """Restore from a sharded directory and return meta.pt's dict.

        Same world size as the save and a GradSync at hand (the usual requeue): every rank reads
        ONLY its own shard (1/W of the bytes) -> pinned -> H2D -> unpack + PR-SUM64 on the GPU, and
        the slices are re-replicated over NVLink by K1's all-gather -- instead of the reference's W
        full reads staggered by 3 s x rank (pyrecover/checkpoint.py:140-141, :352-356).
        Any other world size: read every shard into the full pinned image, upload, unpack, verify
        each shard's PR-SUM64."""
        self.wait()
        d = Path(dirpath)
        sides = sorted(d.glob("b200_*.json"))
        native_dcp = bool(sides)
        if not sides:
            sides = sorted(d.glob("shard_*.json"))          # format 1 (raw shard files), still readable
        if not sides:
            raise RuntimeError(f"No checkpoint shards found in {d}")
        infos = [json.loads(s.read_text()) for s in sides]
        saved_world = infos[0]["world"]
        complete = (d / ".metadata").exists() if native_dcp else \
            all((d / f"COMPLETE_{i['rank']:05d}").exists() for i in infos)
        if len(infos) != saved_world or not complete:
            raise RuntimeError(f"Sharded checkpoint {d} is incomplete ({len(infos)} of {saved_world} shards)")
        if infos[0]["full_image_bytes"] != self.full_bytes:
            raise RuntimeError("checkpoint image layout differs from the live model")
        meta_path = d / ("b200_meta.pt" if native_dcp else "meta.pt")

        def shard_file(info):
            return d / (info["file"] if native_dcp else f"shard_{info['rank']:05d}.bin")
        if native_dcp and grad_sync is not None and world > 1 and saved_world == world and grad_sync.world == world \
                and os.environ.get("PYRECOVER_B200_SLICE_LOAD", "1") != "0":
            self._load_own_shard(d, next(i for i in infos if i["rank"] == rank), rank, world, grad_sync, verify)
            return torch.load(str(meta_path), map_location="cpu", weights_only=False)
        img = self.full()
        mv = img.host.view()
        for info in infos:
            with open(shard_file(info), "rb", buffering=0) as f:
                for goff, n, loff in info["pieces"]:
                    f.seek(loff)
                    _read_exact(f, mv[goff:goff + n], str(d))
        img.restore()
        if verify:
            image = img._last_image
            for info in infos:
                s = torch.zeros(2, dtype=torch.int64, device=self.device)
                for goff, n, loff in info["pieces"]:
                    K.checksum_into(image[goff:goff + n], s, loff // 4)
                if K._sum_to_ints(s) != tuple(info["sum64"]):
                    raise RuntimeError(f"Checksum mismatch for shard {info['rank']} of {d}")
        return torch.load(str(meta_path), map_location="cpu", weights_only=False)

    def _load_own_shard(self, d: Path, info: Dict[str, Any], rank: int, world: int, grad_sync, verify: bool) -> None:
        import torch.distributed as dist
        img = self.shard(rank, world)
        if info["image_bytes"] != img.image_bytes or \
                [list(x) for x in info["pieces"]] != [[p.global_off, p.nbytes, p.img_off] for p in img.pieces]:
            raise RuntimeError(f"shard {rank} of {d} does not match this rank's slice layout")
        with open(d / info["file"], "rb", buffering=0) as f:
            _read_exact(f, img.host.view(0, img.image_bytes), str(d))
        got = img.restore()                        # H2D + unpack into the live slices + PR-SUM64
        if verify and tuple(info["sum64"]) != got:
            raise RuntimeError(f"Checksum mismatch for shard {rank} of {d}: stored {tuple(info['sum64'])}, "
                               f"computed on GPU {got}")
        for a in self.arenas:                      # slices -> every rank, over NVLink
            for name, t in (("p", a.p), ("m", a.m), ("v", a.v)):
                grad_sync.allgather_arena(name, t)
        for t in self.extras:                      # stray tensors live in rank 0's shard
            dist.broadcast(t, src=dist.get_global_rank(grad_sync.pg, 0) if grad_sync.pg is not None else 0,
                           group=grad_sync.pg)
        self.last_stats = {"load": "own shard + NVLink all-gather", "bytes_read": img.image_bytes}

    def close(self) -> None:
        self.wait()
        torch.cuda.synchronize(self.device)
        for im in (self._full, self._shard):
            if im is not None:
                im.close()
        self._full = self._shard = None


# ---------------------------------------------------------------------- small helpers -----
class _Small:
    __slots__ = ("dtype", "off", "shape")

    def __init__(self, dtype, off, shape):
        self.dtype, self.off, self.shape = dtype, off, shape


def _finish_small(obj: Any, refs: Dict[torch.dtype, PT.RecordRef]) -> Any:
    if isinstance(obj, _Small):
        return PT.TensorStub(refs[obj.dtype], obj.off, obj.shape)
    if isinstance(obj, dict):
        return type(obj)((k, _finish_small(v, refs)) for k, v in obj.items())
    if isinstance(obj, list):
        return [_finish_small(v, refs) for v in obj]
    if isinstance(obj, tuple):
        return tuple(_finish_small(v, refs) for v in obj)
    return obj


def _to_cpu(obj: Any) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.detach().to("cpu").clone()
    if isinstance(obj, dict):
        return type(obj)((k, _to_cpu(v)) for k, v in obj.items())
    if isinstance(obj, list):
        return [_to_cpu(v) for v in obj]
    if isinstance(obj, tuple):
        return tuple(_to_cpu(v) for v in obj)
    return copy.deepcopy(obj)


def _read_exact(f, mv: memoryview, what: str) -> None:
    got, n = 0, mv.nbytes
    while got < n:
        k = f.readinto(mv[got:got + COPY_CHUNK])
        if not k:
            raise RuntimeError(f"{what}: truncated checkpoint ({got} of {n} bytes)")
        got += k


def _await_sidecars(d: Path, world: int, step, timeout_s: Optional[float] = None) -> None:
    """Rank 0's writer thread, before it writes `.metadata`: every rank's file is durable (each rank writes
    its side-car after its .distcp has been renamed into place).  Ranks save the same step within a few
    milliseconds of each other, so this normally returns at once."""
    timeout_s = float(os.environ.get("PYRECOVER_B200_PEER_WRITE_TIMEOUT_S", "1800")) if timeout_s is None else timeout_s
    deadline = time.monotonic() + timeout_s
    missing = list(range(world))
    while missing:
        still = []
        for r in missing:
            try:
                if json.loads((d / f"b200_{r:05d}.json").read_text()).get("step") != step:
                    still.append(r)
            except (OSError, ValueError):
                still.append(r)
        missing = still
        if missing:
            if time.monotonic() > deadline:
                raise RuntimeError(f"sharded checkpoint {d}: ranks {missing} did not finish writing within {timeout_s:.0f} s")
            time.sleep(0.02)


def _write_text_atomic(path: Path, text: str) -> None:
    tmp = str(path) + ".tmp"
    with open(tmp, "w") as f:
        f.write(text)
    os.replace(tmp, str(path))


# ---------------------------------------------------------------------- rotation ----------
_STEP_RE = re.compile(r"ckpt_(\d+)")


def _step_key(p: Path):
    m = _STEP_RE.search(p.name)
    return (int(m.group(1)) if m else -1, "_final" in p.name, p.name)


def rotate_files(base: Path, max_keep: int) -> None:
    """Keep the newest `max_keep` vanilla checkpoints (reference pyrecover/checkpoint.py:87-101),
    ordered by STEP NUMBER rather than lexicographically (ckpt_1000.pt sorts after ckpt_200.pt)."""
    if max_keep <= 0:
        return
    # only files named like checkpoints: a sharded directory's b200_meta.pt / rng_<r>.pt living under the same
    # experiment directory must never be "rotated" away (the reference removes any old *.pt)
    files = sorted((x for x in base.glob("**/*.pt") if x.is_file() and _STEP_RE.match(x.name)), key=_step_key)
    for f in files[:-max_keep] if len(files) > max_keep else []:
        for side in ("", ".md5", ".b200sum"):
            try:
                os.remove(str(f) + side)
            except FileNotFoundError:
                pass


def rotate_dirs(base: Path, max_keep: int) -> None:
    """Same for sharded checkpoint directories; only `ckpt_*` directories are candidates (the
    reference deletes ANY old directory, pyrecover/checkpoint.py:277-292)."""
    if max_keep <= 0:
        return
    import shutil
    dirs = sorted((x for x in base.glob("ckpt_*") if x.is_dir()), key=_step_key)
    now = time.time()
    for x in dirs[:-max_keep] if len(dirs) > max_keep else []:
        # ranks are only loosely in step while their writer threads run: never pull a directory
        # from under a peer that is still writing its shard (incomplete and recently touched)
        if _sharded_dir_in_progress(x, now):
            continue
        shutil.rmtree(x, ignore_errors=True)


def _sharded_dir_in_progress(d: Path, now: float, grace_s: float = 900.0) -> bool:
    sides = list(d.glob("shard_*.json")) + list(d.glob("shard_*.bin.tmp")) + list(d.glob("shard_*.bin"))
    if not sides and not (d / ".metadata").exists():
        try:
            return now - d.stat().st_mtime < grace_s          # just created, nothing written yet
        except OSError:
            return False
    marks = list(d.glob("COMPLETE_*"))
    if marks:
        try:
            if len(marks) >= int(marks[0].read_text()):
                return False                                   # complete
        except (ValueError, OSError):
            pass
    elif (d / ".metadata").exists():
        return False                                           # a torch.distributed.checkpoint directory
    try:
        newest = max(f.stat().st_mtime for f in d.iterdir())
    except (OSError, ValueError):
        return False
    return now - newest < grace_s
