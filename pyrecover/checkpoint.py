"""pyrecover.checkpoint -- the reference's save/load/resume API on the B200-native snapshot path.

Same five callables, same positional order, same return values as the reference module
(pyrecover/checkpoint.py:25 save_ckpt_vanilla, :108 load_ckpt_vanilla, :218
save_ckpt_distributed, :300 load_ckpt_distributed, :371 get_latest_checkpoint), so the
reference's train.py (:199-209, :319-331, :358-370) and tests/check_weights_equality.py (:16,
:95, :114) run against this package unmodified.

Two paths sit behind each function, selected by WHERE THE STATE LIVES, not by availability:

  * CUDA model + pyrecover_b200.B200AdamW  ->  B200 path: device-side pack/checksum kernel,
    pinned arena, async D2H on a side stream, writer thread (pyrecover_b200/snapshot.py).  If the
    CUDA library is missing this path raises; it never degrades to the host path.
  * anything else (CPU model on gloo, a stock torch optimizer)  ->  host path: what the reference
    does (torch.save / torch.load / torch.distributed.checkpoint), minus its defects (non-atomic
    write, whole-file md5 read into RAM, 3 s x rank stagger, `.device` assert, lexicographic
    rotation).  BASELINE config 1 (tiny model, CPU/gloo) runs here.
"""
from __future__ import annotations

import hashlib
import os
import threading
import weakref
from pathlib import Path
from typing import Any, Dict, Optional, Tuple

import torch
import torch.distributed as dist

_MD5_CHUNK = 64 << 20
# one snapshot engine per live optimizer; keyed by a weak reference (an id() may be reused by a later
# object once the optimizer is gone, and the entry must not keep the arenas alive)
_SNAPSHOTTERS: "weakref.WeakKeyDictionary[Any, Any]" = weakref.WeakKeyDictionary()


# ------------------------------------------------------------------------- helpers --------
def _unwrap(model: torch.nn.Module) -> torch.nn.Module:
    return model.module if hasattr(model, "module") else model


def _is_b200(optimizer) -> bool:
    try:
        from pyrecover_b200.optim import B200AdamW
    except Exception:
        return False
    return isinstance(optimizer, B200AdamW)


def _snapshotter(model, optimizer):
    from pyrecover_b200.snapshot import Snapshotter
    snap = _SNAPSHOTTERS.get(optimizer)
    if snap is None:
        snap = Snapshotter(model, optimizer)
        _SNAPSHOTTERS[optimizer] = snap
    return snap


def wait_for_checkpoints() -> None:
    """Block until every asynchronous checkpoint write has reached the disk (additive API:
    call before reading a just-written checkpoint from another process, or at job end)."""
    for snap in list(_SNAPSHOTTERS.values()):
        snap.wait()


def take_durable_checkpoint_time() -> Optional[float]:
    """Seconds the most recently FINISHED asynchronous checkpoint took from the save call to the renamed
    file -- reported once per write, None when no write has finished since the last call; never blocks.
    This, not the milliseconds a save_* call takes to enqueue the snapshot, is what the time-aware
    policy has to learn from (reference train.py:334-337 times its synchronous save)."""
    best = None
    for snap in list(_SNAPSHOTTERS.values()):
        t = snap.take_durable_time()
        if t is not None and (best is None or t > best):
            best = t
    return best


def release_checkpoint_engines() -> None:
    """Free staging buffers / pinned arenas (called from dist_utils.maybe_cleanup_distributed)."""
    for opt in list(_SNAPSHOTTERS.keys()):
        snap = _SNAPSHOTTERS.pop(opt, None)
        if snap is not None:
            snap.close()


def _save_barrier(fast: bool) -> None:
    """The barrier the reference places before and after a save (pyrecover/checkpoint.py:56,103).
    Host path: dist.barrier(), as the reference.  B200 path: nothing.  The save is asynchronous and
    stream-ordered after this rank's own optimizer step; what every rank writes is state that is
    identical on all of them (replicated) or its own slice of it (sharded), and the steps themselves
    are kept in lock-step by the begin/end barriers inside every K1 launch -- a second rendezvous
    would only import the slowest rank's skew into everybody's training stream (measured in round 1:
    up to 10 ms per checkpoint at 8 GPUs)."""
    if not fast:
        dist.barrier()


def _pre_save_checks(optimizer) -> None:
    """B200 path: never publish state that a failed step may have touched."""
    check = getattr(optimizer, "check_finite", None)
    if check is not None:
        check()
    gs = getattr(optimizer, "_grad_sync", None)
    if gs is not None and gs.comm.timed_out():
        raise RuntimeError("refusing to write a checkpoint: the gradient exchange timed out waiting for a peer")


def _rng_state() -> Dict[str, torch.Tensor]:
    st = {"cpu": torch.get_rng_state()}
    if torch.cuda.is_available():
        st["cuda"] = torch.cuda.get_rng_state()
    return st


def _set_rng_state(st: Optional[Dict[str, torch.Tensor]]) -> None:
    if not st:
        return
    torch.set_rng_state(st["cpu"].cpu())
    if "cuda" in st and torch.cuda.is_available():
        torch.cuda.set_rng_state(st["cuda"].cpu())


def _file_md5(path: str) -> str:
    h = hashlib.md5()
    with open(path, "rb") as f:
        while True:
            chunk = f.read(_MD5_CHUNK)
            if not chunk:
                return h.hexdigest()
            h.update(chunk)


def _collect_state(model, optimizer, lr_scheduler, sampler, step, epoch) -> Dict[str, Any]:
    state = {"epoch": epoch, "step": step, "model": _unwrap(model).state_dict(),
             "optimizer": optimizer.state_dict()}
    if lr_scheduler is not None:
        state["lr_scheduler"] = lr_scheduler.state_dict()
    if sampler and hasattr(sampler, "set_state"):
        state["sampler_state"] = sampler.state_dict()
    state["rng_state"] = _rng_state()
    return state


# ------------------------------------------------------------------------- vanilla --------
def save_ckpt_vanilla(
    model: torch.nn.Module,
    optimizer: torch.optim.Optimizer,
    lr_scheduler=None,
    sampler=None,
    step: int = 0,
    epoch: Optional[int] = None,
    checkpoint_path: str = "/iopstores/scratch",
    max_keep: int = 3,
    verify: bool = True,
    is_distributed: bool = False,
    rank: int = 0,
) -> str:
    """Rank 0 writes one torch.load-able file (top-level keys epoch, step, model, optimizer,
    lr_scheduler[, sampler_state], rng_state) + `<path>.md5` when verify; older files beyond
    max_keep are removed; all ranks meet at a barrier before and after.

    B200 path: returns when the snapshot is staged (pack kernel + async D2H enqueued); the file
    becomes durable in the writer thread -- `wait_for_checkpoints()`, the next save and every
    load join it."""
    fast = _is_b200(optimizer)
    if is_distributed:
        _save_barrier(fast)
    if fast:
        _pre_save_checks(optimizer)
        if getattr(optimizer, "_grad_sync", None) is not None:
            optimizer._grad_sync.replicate_moments()      # collective: every rank takes part
    if rank == 0 or not is_distributed:
        if fast:
            snap = _snapshotter(model, optimizer)
            dyn = snap.dynamic_state(lr_scheduler, sampler, step, epoch, _rng_state())
            snap.save_vanilla(str(checkpoint_path), dyn, verify, max_keep)
        else:
            state = _collect_state(model, optimizer, lr_scheduler, sampler, step, epoch)
            _host_save_vanilla(state, str(checkpoint_path), verify, max_keep)
    if is_distributed:
        _save_barrier(fast)
    return checkpoint_path


def _host_save_vanilla(state, path: str, verify: bool, max_keep: int) -> None:
    from pyrecover_b200.snapshot import rotate_files
    tmp = path + ".tmp"
    torch.save(state, tmp)
    if verify:
        digest = _file_md5(tmp)
    os.replace(tmp, path)
    if verify:
        with open(path + ".md5", "w") as f:
            f.write(digest)
    rotate_files(Path(path).parent, max_keep)


def load_ckpt_vanilla(
    model: torch.nn.Module,
    optimizer: torch.optim.Optimizer,
    lr_scheduler=None,
    sampler=None,
    checkpoint_path: str = "latest",
    experiment_dir: str = "/iopstores/scratch",
    verify: bool = True,
    is_distributed: bool = False,
    rank: int = 0,
) -> Tuple[int, int]:
    """Restore model / optimizer / scheduler / sampler / RNG in place; returns (epoch, step)."""
    if is_distributed:
        dist.barrier()
    wait_for_checkpoints()
    if checkpoint_path == "latest":
        checkpoint_path = get_latest_checkpoint(experiment_dir)
        if checkpoint_path is None:
            raise RuntimeError(f"No checkpoint found in {experiment_dir}")
    checkpoint_path = str(checkpoint_path)

    ck = torch.load(checkpoint_path, map_location="cpu", mmap=True)
    meta = ck.get("b200")
    fast = meta is not None and _is_b200(optimizer)
    md5_thread, md5_result = None, {}
    if fast:
        # the whole image is read once into the pinned arena, uploaded, scattered by the unpack
        # kernel and verified with the on-GPU checksum -- on every rank, no stagger
        _snapshotter(model, optimizer).load_vanilla_image(checkpoint_path, meta, verify)
        optimizer.load_state_dict(ck["optimizer"], tensors_restored=True)
    else:
        if verify and rank == 0 and os.path.exists(checkpoint_path + ".md5"):
            def _verify():
                try:
                    md5_result["ok"] = _file_md5(checkpoint_path) == open(checkpoint_path + ".md5").read().strip()
                except Exception as e:  # noqa: BLE001
                    md5_result["ok"], md5_result["err"] = False, str(e)
            md5_thread = threading.Thread(target=_verify)
            md5_thread.start()
        _unwrap(model).load_state_dict(ck["model"])
        optimizer.load_state_dict(ck["optimizer"])
    if lr_scheduler is not None and "lr_scheduler" in ck:
        lr_scheduler.load_state_dict(ck["lr_scheduler"])
    if sampler and "sampler_state" in ck:
        sampler.load_state_dict(ck["sampler_state"])
    _set_rng_state(ck.get("rng_state"))
    epoch, step = ck.get("epoch", 0), ck.get("step", 0)
    if md5_thread is not None:
        md5_thread.join()
        if not md5_result.get("ok", False):
            raise RuntimeError(md5_result.get("err") or f"Checksum mismatch for checkpoint {checkpoint_path}")
    if is_distributed:
        dist.barrier()
    print(f"Checkpoint loaded from {checkpoint_path} (epoch {epoch}, step {step})")
    return epoch, step


# ------------------------------------------------------------------------- sharded --------
def save_ckpt_distributed(
    model: torch.nn.Module,
    optimizer: torch.optim.Optimizer,
    lr_scheduler=None,
    sampler=None,
    step: int = 0,
    epoch: Optional[int] = None,
    checkpoint_path: str = "/iopstores/scratch",
    max_keep: int = 3,
    verify: bool = True,
    is_distributed: bool = False,
    rank: int = 0,
) -> str:
    """Sharded save into the directory `checkpoint_path`: every rank stores 1/W of the state.

    B200 path: rank r packs its element slices of (p, m, v) -- the reduce-scatter slices of the
    gradient exchange -- into `__<r>_0.distcp`, framed box by box as torch.distributed.checkpoint items,
    and rank 0 writes `.metadata` last: the directory is what the reference's dist_cp.save call
    (pyrecover/checkpoint.py:254-274) produces, written without re-replicating or re-serialising anything
    (pyrecover_b200/dcpfile.py).  Host path (CPU/gloo): the reference's torch.distributed.checkpoint call."""
    fast = _is_b200(optimizer)
    if is_distributed:
        _save_barrier(fast)
    os.makedirs(os.path.dirname(str(checkpoint_path)) or ".", exist_ok=True)
    if fast:
        _pre_save_checks(optimizer)
        world = dist.get_world_size() if is_distributed else 1
        groups = [{k: v for k, v in g.items() if k != "params"} for g in optimizer.param_groups]
        small = {"epoch": epoch, "step": step, "param_groups": groups,
                 "opt_steps": list(optimizer._steps), "rng_state": _rng_state()}
        if lr_scheduler is not None:
            small["lr_scheduler"] = lr_scheduler.state_dict()
        if sampler and hasattr(sampler, "set_state"):
            small["sampler_state"] = sampler.state_dict()
        _snapshotter(model, optimizer).save_sharded(str(checkpoint_path), small, rank if is_distributed else 0,
                                                    world, max_keep)
    else:
        _host_save_distributed(model, optimizer, lr_scheduler, sampler, step, epoch, str(checkpoint_path),
                               max_keep, rank)
    if is_distributed:
        _save_barrier(fast)
    return checkpoint_path


def _materialize_adam_state(optimizer) -> None:
    """torch creates Adam's moments lazily at the first step (torch/optim/adam.py:168-184) and
    torch.distributed.checkpoint uses `optimizer.state_dict()` as its load template, so a resume
    into a FRESH optimizer silently drops exp_avg / exp_avg_sq (reference defect, SURVEY.md
    section 5).  Give every parameter a zero state first so the template has the keys."""
    if not isinstance(optimizer, (torch.optim.Adam, torch.optim.AdamW)) or len(optimizer.state) > 0:
        return
    for group in optimizer.param_groups:
        on_device = bool(group.get("fused") or group.get("capturable"))
        for p in group["params"]:
            if not p.requires_grad:
                continue
            step = torch.zeros((), dtype=torch.float32, device=p.device if on_device else "cpu")
            optimizer.state[p] = {"step": step, "exp_avg": torch.zeros_like(p, memory_format=torch.preserve_format),
                                  "exp_avg_sq": torch.zeros_like(p, memory_format=torch.preserve_format)}


def _dcp_objects(model, optimizer, lr_scheduler, sampler, epoch, step) -> Dict[str, Any]:
    objs = {"model": model, "optimizer": optimizer, "metadata": {"epoch": epoch, "step": step}}
    if lr_scheduler is not None:
        objs["lr_scheduler"] = lr_scheduler
    if sampler and hasattr(sampler, "set_state"):
        objs["sampler"] = sampler
    return objs


def _host_save_distributed(model, optimizer, lr_scheduler, sampler, step, epoch, path, max_keep, rank):
    import torch.distributed.checkpoint as dist_cp
    from pyrecover_b200.snapshot import rotate_dirs
    dist_cp.save(state_dict=_dcp_objects(model, optimizer, lr_scheduler, sampler, epoch, step),
                 checkpoint_id=path, storage_writer=dist_cp.FileSystemWriter(path))
    if rank == 0:
        rotate_dirs(Path(path).parent, max_keep)


def load_ckpt_distributed(
    model: torch.nn.Module,
    optimizer: torch.optim.Optimizer,
    lr_scheduler=None,
    sampler=None,
    checkpoint_path: str = "latest",
    experiment_dir: str = "/iopstores/scratch",
    verify: bool = True,
    is_distributed: bool = False,
    rank: int = 0,
) -> Tuple[int, int]:
    """Restore from a sharded checkpoint directory; returns (epoch, step)."""
    if is_distributed:
        dist.barrier()
    wait_for_checkpoints()
    if checkpoint_path == "latest":
        checkpoint_path = get_latest_checkpoint(experiment_dir, distributed=True)
        if checkpoint_path is None:
            raise RuntimeError(f"No checkpoint found in {experiment_dir}")
    checkpoint_path = str(checkpoint_path)
    native = any(Path(checkpoint_path).glob("b200_*.json")) or any(Path(checkpoint_path).glob("shard_*.json"))
    if native and not _is_b200(optimizer) and (Path(checkpoint_path) / ".metadata").exists():
        native = False      # a stock optimizer: our directories are torch.distributed.checkpoint directories too
    if native:
        if not _is_b200(optimizer):
            raise RuntimeError(f"{checkpoint_path} is a pyrecover_b200 sharded checkpoint; it loads into a "
                               "B200AdamW optimizer on a CUDA device")
        world = dist.get_world_size() if is_distributed else 1
        small = _snapshotter(model, optimizer).load_sharded(
            checkpoint_path, verify, rank=dist.get_rank() if is_distributed else 0, world=world,
            grad_sync=getattr(optimizer, "_grad_sync", None))
        for g, saved in zip(optimizer.param_groups, small["param_groups"]):
            g.update({k: v for k, v in saved.items() if k != "params"})
        for gi, s in enumerate(small["opt_steps"]):
            optimizer.set_step(gi, s)
        if getattr(optimizer, "_grad_sync", None) is not None:
            optimizer._grad_sync.moments_loaded()
        if lr_scheduler is not None and "lr_scheduler" in small:
            lr_scheduler.load_state_dict(small["lr_scheduler"])
        if sampler and "sampler_state" in small:
            sampler.load_state_dict(small["sampler_state"])
        rng = small.get("rng_state")
        own = Path(checkpoint_path) / f"rng_{rank:05d}.pt"
        if rank != 0 and own.exists():
            rng = torch.load(str(own), map_location="cpu")["rng_state"]
        _set_rng_state(rng)
        epoch, step = small["epoch"], small["step"]
    else:
        import torch.distributed.checkpoint as dist_cp
        _materialize_adam_state(optimizer)
        objs = _dcp_objects(model, optimizer, lr_scheduler, sampler, 0, 0)
        dist_cp.load(state_dict=objs, checkpoint_id=checkpoint_path,
                     storage_reader=dist_cp.FileSystemReader(checkpoint_path))
        epoch, step = objs["metadata"]["epoch"], objs["metadata"]["step"]
    if is_distributed:
        dist.barrier()
    print(f"Distributed checkpoint loaded from {checkpoint_path} (epoch {epoch}, step {step})")
    return epoch, step


# ------------------------------------------------------------------------- lookup ---------
def _dir_complete(d: Path) -> bool:
    sides = list(d.glob("shard_*.json"))
    if sides:
        marks = list(d.glob("COMPLETE_*"))
        try:
            world = int(marks[0].read_text()) if marks else -1
        except (ValueError, OSError):
            return False
        return len(marks) == world and (d / "meta.pt").exists()
    return (d / ".metadata").exists()


def get_latest_checkpoint(checkpoint_dir: str, distributed: bool = False) -> Optional[str]:
    """Newest checkpoint by modification time: `ckpt_*` directories when distributed, `**/*.pt`
    files otherwise; incomplete directories and in-flight `.tmp` files are never returned."""
    base = Path(checkpoint_dir)
    if distributed:
        items = [d for d in base.glob("ckpt_*") if d.is_dir() and _dir_complete(d)]
    else:
        items = [f for f in base.glob("**/*.pt") if f.is_file() and f.name not in ("meta.pt", "b200_meta.pt")
                 and not f.name.startswith("rng_")]
    if not items:
        return None
    items.sort(key=lambda x: x.stat().st_mtime)
    return str(items[-1])
