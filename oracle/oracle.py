"""ctypes front-end of the CPU oracle (oracle/pr_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs as the checker -- never by pyrecover_b200/ or pyrecover/.

bf16 arrays travel as numpy uint16 (bit patterns); use `to_bits` / `from_bits` to convert
torch tensors.  Each function cites the reference lines it restates in pr_oracle.c.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "libpr_oracle.so"


def build(force: bool = False) -> Path:
    """Compile pr_oracle.c with gcc (seconds).  Building the checker is not using it."""
    src = _HERE / "pr_oracle.c"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-C", str(_HERE), "-s"])
    return _LIB_PATH


class AdamWConsts(ctypes.Structure):
    _fields_ = [(k, ctypes.c_float) for k in
                ("decay", "b1", "omb1", "b2", "omb2", "step_size", "bc2_sqrt", "eps", "gscale")]


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(str(_LIB_PATH))
        vp, sz, i64, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int64, ctypes.c_uint64
        dbl, cp = ctypes.c_double, ctypes.POINTER(AdamWConsts)
        L.orc_adamw_prepare.argtypes = [dbl, dbl, dbl, dbl, dbl, i64, dbl, cp]
        L.orc_adamw_bf16.argtypes = [vp, vp, vp, vp, sz, cp, ctypes.c_int]
        L.orc_adamw_f32.argtypes = [vp, vp, vp, vp, sz, cp]
        L.orc_adamw_bf16_gf32.argtypes = [vp, vp, vp, vp, sz, cp]
        L.orc_allreduce_bf16.argtypes = [vp, ctypes.c_int, sz, ctypes.c_float, vp]
        L.orc_allreduce_bf16_to_f32.argtypes = [vp, ctypes.c_int, sz, ctypes.c_float, vp]
        L.orc_allreduce_bf16_ddp.argtypes = [vp, ctypes.c_int, sz, vp]
        L.orc_allreduce_f32.argtypes = [vp, ctypes.c_int, sz, ctypes.c_float, vp]
        L.orc_checksum64.argtypes = [vp, sz, u64, vp]
        L.orc_pack.argtypes = [vp, vp, vp, ctypes.c_int, vp, u64, vp]
        L.orc_unpack.argtypes = [vp, vp, vp, ctypes.c_int, vp]
        L.orc_grad_norm_bf16.argtypes = [vp, sz, ctypes.c_float, ctypes.c_float, vp]
        L.orc_grad_norm_f32.argtypes = [vp, sz, ctypes.c_float, ctypes.c_float, vp]
        L.orc_scale_mul.argtypes = [cp, ctypes.c_float]
        for f in ("orc_grad_norm_bf16", "orc_grad_norm_f32", "orc_scale_mul", "orc_adamw_prepare", "orc_adamw_bf16", "orc_adamw_f32", "orc_adamw_bf16_gf32",
                  "orc_allreduce_bf16", "orc_allreduce_bf16_to_f32", "orc_allreduce_bf16_ddp",
                  "orc_allreduce_f32", "orc_checksum64", "orc_pack", "orc_unpack"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


# --------------------------------------------------------------------------- helpers ----

def to_bits(t) -> np.ndarray:
    """torch bf16/fp32 tensor (any device) -> contiguous numpy array (uint16 bits / float32)."""
    import torch
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    if t.dtype == torch.float32:
        return t.numpy().copy()
    raise TypeError(t.dtype)


def from_bits(a: np.ndarray):
    import torch
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)
    return torch.from_numpy(a.copy())


def _ptr(a: np.ndarray) -> ctypes.c_void_p:
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def adamw_consts(lr, beta1, beta2, eps, wd, step, gscale=1.0, coef=None) -> AdamWConsts:
    """coef: clip coefficient (binary32) multiplied into the gradient scale, as K2 does with
    pr_adamw_args.grad_scale_dev."""
    c = AdamWConsts()
    lib().orc_adamw_prepare(lr, beta1, beta2, eps, wd, int(step), gscale, ctypes.byref(c))
    if coef is not None:
        lib().orc_scale_mul(ctypes.byref(c), float(coef))
    return c


def grad_norm(g: np.ndarray, max_norm: float, prescale: float = 1.0):
    """(norm, coefficient, nonfinite) of a flat gradient array (uint16 bf16 bits or float32) --
    reference utils.py:84-89 (clip_grad_norm_)."""
    out = np.zeros(3, dtype=np.float32)
    g = np.ascontiguousarray(g)
    if g.dtype == np.uint16:
        lib().orc_grad_norm_bf16(_ptr(g), g.size, prescale, max_norm, _ptr(out))
    else:
        assert g.dtype == np.float32
        lib().orc_grad_norm_f32(_ptr(g), g.size, prescale, max_norm, _ptr(out))
    return float(out[0]), float(out[1]), bool(out[2])


# --------------------------------------------------------------------------- AdamW ------

def adamw_bf16(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.01, step=1,
               gscale=1.0, mode=0, coef=None):
    """In-place on uint16 arrays p, m, v.  mode 0 = fp32 op-math (fused class),
    mode 1 = bf16 rounding after every torch op (_single_tensor_adam class)."""
    c = adamw_consts(lr, beta1, beta2, eps, wd, step, gscale, coef)
    for a in (p, g, m, v):
        assert a.dtype == np.uint16 and a.size == p.size
    lib().orc_adamw_bf16(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.size, ctypes.byref(c), mode)


def adamw_f32(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.01, step=1, gscale=1.0):
    c = adamw_consts(lr, beta1, beta2, eps, wd, step, gscale)
    for a in (p, g, m, v):
        assert a.dtype == np.float32 and a.size == p.size
    lib().orc_adamw_f32(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.size, ctypes.byref(c))


def adamw_bf16_gf32(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.01, step=1,
                    gscale=1.0):
    c = adamw_consts(lr, beta1, beta2, eps, wd, step, gscale)
    assert g.dtype == np.float32
    lib().orc_adamw_bf16_gf32(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.size, ctypes.byref(c))


# --------------------------------------------------------------------------- allreduce --

def _ptr_table(bufs):
    arr = (ctypes.c_void_p * len(bufs))(*[b.ctypes.data for b in bufs])
    return arr


def allreduce_bf16(bufs, scale) -> np.ndarray:
    n = bufs[0].size
    out = np.empty(n, dtype=np.uint16)
    lib().orc_allreduce_bf16(ctypes.cast(_ptr_table(bufs), ctypes.c_void_p), len(bufs), n,
                             scale, _ptr(out))
    return out


def allreduce_bf16_to_f32(bufs, scale) -> np.ndarray:
    n = bufs[0].size
    out = np.empty(n, dtype=np.float32)
    lib().orc_allreduce_bf16_to_f32(ctypes.cast(_ptr_table(bufs), ctypes.c_void_p), len(bufs), n,
                                    scale, _ptr(out))
    return out


def allreduce_bf16_ddp(bufs) -> np.ndarray:
    n = bufs[0].size
    out = np.empty(n, dtype=np.uint16)
    lib().orc_allreduce_bf16_ddp(ctypes.cast(_ptr_table(bufs), ctypes.c_void_p), len(bufs), n,
                                 _ptr(out))
    return out


def allreduce_f32(bufs, scale) -> np.ndarray:
    n = bufs[0].size
    out = np.empty(n, dtype=np.float32)
    lib().orc_allreduce_f32(ctypes.cast(_ptr_table(bufs), ctypes.c_void_p), len(bufs), n,
                            scale, _ptr(out))
    return out


# --------------------------------------------------------------------------- checksum ---

def checksum64(data, word_base: int = 0):
    """PR-SUM64 of a bytes-like / numpy array -> (A, B) python ints."""
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else \
        np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    out = np.zeros(2, dtype=np.uint64)
    lib().orc_checksum64(_ptr(a) if a.size else None, a.size, word_base, _ptr(out))
    return int(out[0]), int(out[1])


def checksum64_numpy(data, word_base: int = 0):
    """Independent numpy restatement (vectorised, no C) used to cross-check the C loop."""
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else \
        np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    pad = (-a.size) % 4
    if pad:
        a = np.concatenate([a, np.zeros(pad, dtype=np.uint8)])
    w = a.view("<u4").astype(np.uint64)
    idx = np.arange(1, w.size + 1, dtype=np.uint64) + np.uint64(word_base)
    with np.errstate(over="ignore"):
        A = int(w.sum(dtype=np.uint64)) if w.size else 0
        B = int((w * idx).sum(dtype=np.uint64)) if w.size else 0
    return A & (2**64 - 1), B & (2**64 - 1)


def combine(parts):
    """Sum of partial (A, B) pairs mod 2^64 (each computed with its own word_base)."""
    A = sum(p[0] for p in parts) & (2**64 - 1)
    B = sum(p[1] for p in parts) & (2**64 - 1)
    return A, B


# --------------------------------------------------------------------------- pack -------

def pack(segments, dst_offsets, total_bytes):
    """segments: list of numpy arrays; returns (packed uint8 array, (A, B))."""
    segs = [np.ascontiguousarray(s).view(np.uint8).reshape(-1) for s in segments]
    n = len(segs)
    src = (ctypes.c_void_p * n)(*[s.ctypes.data for s in segs])
    nb = np.array([s.size for s in segs], dtype=np.uint64)
    off = np.array(dst_offsets, dtype=np.uint64)
    dst = np.empty(total_bytes, dtype=np.uint8)
    out = np.zeros(2, dtype=np.uint64)
    lib().orc_pack(ctypes.cast(src, ctypes.c_void_p), _ptr(nb), _ptr(off), n, _ptr(dst),
                   total_bytes, _ptr(out))
    return dst, (int(out[0]), int(out[1]))


def unpack(packed, nbytes, src_offsets):
    n = len(nbytes)
    outs = [np.empty(int(b), dtype=np.uint8) for b in nbytes]
    dst = (ctypes.c_void_p * n)(*[o.ctypes.data for o in outs])
    nb = np.array(nbytes, dtype=np.uint64)
    off = np.array(src_offsets, dtype=np.uint64)
    lib().orc_unpack(_ptr(packed), _ptr(nb), _ptr(off), n, ctypes.cast(dst, ctypes.c_void_p))
    return outs
