"""Recipe: stage the reference's own Python sources for the timed CPU arm -- TEST / BASELINE INFRASTRUCTURE.

The reference (Shaswat-G/PyRecover) is pure Python, so "compiling it from where it lies" is a copy:
the three files its CPU training step needs go from /root/reference into oracle/_ref/, which is
git-ignored (no reference source enters the history) but travels to the GPU box with the snapshot,
like a built .so.  `__graft_entry__.build()` runs this whenever /root/reference is present; on the
GPU box only the staged files are used.  bench.py's `--impl reference` arm and `cpu_baseline` leg then
run the reference's OWN model.py, utils.build_lr_scheduler and pyrecover/checkpoint.py:save_ckpt_vanilla
(kind "reference"); without the staged files they fall back to the port in cpu_reference_port.py.

    python oracle/stage_ref.py [/root/reference]
"""
from __future__ import annotations

import importlib.util
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"
FILES = ("model.py", "utils.py", "pyrecover/checkpoint.py")


def stage(ref_root: str = "/root/reference") -> bool:
    """Copy FILES from the reference checkout; returns False (and stages nothing) when it is absent."""
    root = Path(ref_root)
    if not all((root / f).is_file() for f in FILES):
        return False
    for f in FILES:
        dst = REF_DIR / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(root / f, dst)
    return True


def available() -> bool:
    return all((REF_DIR / f).is_file() for f in FILES)


def load(name: str):
    """Import one staged file by path under a private module name (the reference's package __init__
    imports modules it does not ship, so `import pyrecover` from there fails -- SURVEY.md section 8c)."""
    path = REF_DIR / {"model": "model.py", "utils": "utils.py", "checkpoint": "pyrecover/checkpoint.py"}[name]
    mod_name = f"_pyrecover_reference_{name}"
    if mod_name in sys.modules:
        return sys.modules[mod_name]
    spec = importlib.util.spec_from_file_location(mod_name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[mod_name] = mod
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    ok = stage(*(sys.argv[1:2]))
    print("staged" if ok else "reference checkout not found; nothing staged", REF_DIR)
