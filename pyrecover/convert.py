"""Offline checkpoint conversion (CPU only, no GPU and no CUDA library needed).

Keeps checkpoints written by the B200 path portable to stock tooling, and the other way round:

    python -m pyrecover.convert SRC DST --to vanilla|dcp

SRC is a vanilla `.pt` file (ours or one written by the reference's save_ckpt_vanilla,
pyrecover/checkpoint.py:25-105) or one of our sharded directories (`__<r>_0.distcp` + `b200_<r>.json` +
`b200_meta.pt`; the older `shard_<r>.bin` + `meta.pt` layout is still read).  Our sharded directories ARE
torch.distributed.checkpoint directories already (stock `dist_cp.load` opens them), so `--to dcp` matters
for vanilla sources.  A DCP directory written by stock torch is not a source here: DCP can only be loaded
into an existing state dict, which `load_ckpt_distributed` does.
  --to vanilla : one torch.load-able file with the reference's top-level keys
                 (epoch, step, model, optimizer, lr_scheduler[, sampler_state], rng_state)
  --to dcp     : a torch.distributed.checkpoint directory with exactly the keys the reference's
                 save_ckpt_distributed produces ("model.<name>", "optimizer.state.<i>.exp_avg",
                 "optimizer.param_groups...", "metadata.epoch/step", "lr_scheduler...."), so that the
                 reference's load_ckpt_distributed (:300-368) -- or any DCP reader -- opens it.
"""
from __future__ import annotations

import argparse
import json
import os
from collections import OrderedDict
from pathlib import Path
from typing import Any, Dict

import numpy as np
import torch

_DTYPES = {"bfloat16": torch.bfloat16, "float32": torch.float32, "float16": torch.float16, "float64": torch.float64,
           "int64": torch.int64, "int32": torch.int32, "uint8": torch.uint8, "int16": torch.int16, "int8": torch.int8,
           "bool": torch.bool}


def _read_sharded(d: Path) -> Dict[str, Any]:
    """Assemble the full image from every shard (any world size) and rebuild the reference's state
    dict from the layout table rank 0 stored in meta.pt."""
    sides = sorted(d.glob("b200_*.json"))
    native_dcp = bool(sides)              # format 2: the shard files are torch.distributed.checkpoint files
    if not sides:
        sides = sorted(d.glob("shard_*.json"))
    infos = [json.loads(p.read_text()) for p in sides]
    if not infos:
        raise RuntimeError(f"No checkpoint shards found in {d}")
    world = infos[0]["world"]
    complete = (d / ".metadata").exists() if native_dcp else all((d / f"COMPLETE_{i['rank']:05d}").exists() for i in infos)
    if len(infos) != world or not complete:
        raise RuntimeError(f"Sharded checkpoint {d} is incomplete ({len(infos)} of {world} shards)")
    meta = torch.load(str(d / ("b200_meta.pt" if native_dcp else "meta.pt")), map_location="cpu", weights_only=False)
    layout = meta.get("layout")
    if layout is None:
        raise RuntimeError(f"{d}/meta.pt carries no tensor layout (written by an older build): load it on a GPU "
                           "with load_ckpt_distributed and re-save")
    image = np.zeros(int(infos[0]["full_image_bytes"]), dtype=np.uint8)
    for info in infos:
        with open(d / (info["file"] if native_dcp else f"shard_{info['rank']:05d}.bin"), "rb") as f:
            for goff, n, loff in info["pieces"]:
                f.seek(loff)
                got = f.readinto(memoryview(image)[goff:goff + n])
                if got != n:
                    raise RuntimeError(f"{d}: shard {info['rank']} is truncated")
    buf = torch.from_numpy(image)

    def view(entry):
        dtype, off, shape = entry
        dt = _DTYPES[dtype]
        numel = int(np.prod(shape)) if len(shape) else 1
        nbytes = numel * torch.empty((), dtype=dt).element_size()
        return buf[off:off + nbytes].view(dt).reshape(shape).clone()

    missing = [k for k, e in layout["model"].items() if e is None]
    if missing:          # written by an older build that allowed state outside the image: nothing here could restore it
        raise RuntimeError(f"{d}: model entries without a place in the checkpoint image: {missing[:5]}")
    model = OrderedDict((k, view(e)) for k, e in layout["model"].items())
    steps = meta.get("opt_steps", [0])
    opt_state = {}
    for i, ent in layout["optimizer"].items():
        if ent:
            opt_state[int(i)] = {"step": torch.tensor(float(steps[0]), dtype=torch.float32),
                                 **{k: view(e) for k, e in ent.items() if e is not None}}
    groups = []
    nxt = 0
    for g in meta["param_groups"]:
        n = len(layout["optimizer"]) if len(meta["param_groups"]) == 1 else g.get("n_params", 0)
        groups.append({**g, "params": list(range(nxt, nxt + n))})
        nxt += n
    state = {"epoch": meta.get("epoch"), "step": meta.get("step"), "model": model,
             "optimizer": {"state": opt_state, "param_groups": groups}}
    for k in ("lr_scheduler", "sampler_state", "rng_state"):
        if k in meta:
            state[k] = meta[k]
    return state


def load_state_cpu(src: str) -> Dict[str, Any]:
    """The reference's checkpoint dict (top-level keys epoch, step, model, optimizer, ...) on the CPU."""
    p = Path(src)
    if p.is_file():
        return torch.load(str(p), map_location="cpu", weights_only=False)
    if any(p.glob("b200_*.json")) or any(p.glob("shard_*.json")):
        return _read_sharded(p)
    raise RuntimeError(f"{src}: neither a checkpoint file nor a pyrecover_b200 sharded directory")


def save_vanilla(state: Dict[str, Any], dst: str) -> str:
    tmp = str(dst) + ".tmp"
    torch.save({k: v for k, v in state.items() if k != "b200"}, tmp)
    os.replace(tmp, str(dst))
    return str(dst)


def save_dcp(state: Dict[str, Any], dst: str) -> str:
    """Same flattened keys as the reference's save_ckpt_distributed, which hands DCP the live
    objects (Stateful) -- i.e. their state_dict()s -- under "model", "optimizer", "lr_scheduler" and
    a plain {"epoch", "step"} dict under "metadata" (pyrecover/checkpoint.py:254-274)."""
    import torch.distributed.checkpoint as dist_cp
    objs: Dict[str, Any] = {"model": dict(state["model"]), "optimizer": state["optimizer"],
                            "metadata": {"epoch": state.get("epoch"), "step": state.get("step")}}
    if state.get("lr_scheduler") is not None:
        objs["lr_scheduler"] = state["lr_scheduler"]
    os.makedirs(os.path.dirname(os.path.abspath(dst)) or ".", exist_ok=True)
    dist_cp.save(state_dict=objs, checkpoint_id=str(dst), storage_writer=dist_cp.FileSystemWriter(str(dst)),
                 no_dist=True)
    return str(dst)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("src")
    ap.add_argument("dst")
    ap.add_argument("--to", choices=["vanilla", "dcp"], required=True)
    a = ap.parse_args(argv)
    state = load_state_cpu(a.src)
    out = save_vanilla(state, a.dst) if a.to == "vanilla" else save_dcp(state, a.dst)
    print(f"wrote {out} (epoch {state.get('epoch')}, step {state.get('step')}, {len(state['model'])} model tensors)")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
