"""ctypes binding of libpyrecover_b200.so (C ABI in include/pyrecover_b200.h).

There is no fallback: if the shared library is missing, `lib()` raises -- build it with
`python -c "import __graft_entry__ as g; g.build()"` or `make -C pyrecover_b200/csrc`.
"""
from __future__ import annotations

import ctypes
import subprocess
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libpyrecover_b200.so"

PR_BF16, PR_F32 = 0, 1
PR_AR_AUTO, PR_AR_ONE_SHOT, PR_AR_TWO_SHOT, PR_AR_MULTIMEM = 0, 1, 2, 3
PR_ABI_VERSION = 3
PR_PACK_AUTO, PR_PACK_LDG, PR_PACK_BULK = 0, 1, 2
PR_GRAD_NORM_SCRATCH = 2048
PR_MAX_WORLD = 8
PR_MAX_REGIONS = 4
PR_EXPORT_BLOB_BYTES = 512


class AdamWArgs(ctypes.Structure):
    """struct pr_adamw_args"""
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("grad_scale", c_double), ("step", c_int64),
                ("numerics", c_int32), ("grad_dtype", c_int32), ("grad_scale_dev", c_void_p)]


_SIGNATURES = {
    # name: (restype, argtypes)
    "pr_last_error": (c_char_p, []),
    "pr_abi_version": (c_int, []),
    "pr_device_info": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "pr_launch_count": (c_uint64, []),
    "pr_adamw": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_uint64, c_int,
                         POINTER(AdamWArgs), c_void_p]),
    "pr_adamw_plan_create": (c_int, [c_void_p, c_void_p, c_int, c_uint64, POINTER(c_void_p)]),
    "pr_adamw_plan_destroy": (c_int, [c_void_p]),
    "pr_adamw_scattered": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(AdamWArgs), c_void_p]),
    "pr_gather_grads": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p]),
    "pr_grad_norm": (c_int, [c_void_p, c_uint64, c_int, c_float, c_float, c_void_p, c_void_p, c_void_p]),
    "pr_grad_norm_scattered": (c_int, [c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p, c_void_p]),
    "pr_checksum": (c_int, [c_void_p, c_uint64, c_uint64, c_void_p, c_void_p]),
    "pr_pack_plan_create": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_uint64, POINTER(c_void_p)]),
    "pr_pack_plan_destroy": (c_int, [c_void_p]),
    "pr_ckpt_pack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "pr_ckpt_unpack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "pr_ckpt_pack_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pr_ckpt_unpack_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pr_ckpt_pack_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_uint64, c_void_p, c_void_p]),
    "pr_host_arena_alloc": (c_int, [c_uint64, POINTER(c_void_p)]),
    "pr_host_arena_free": (c_int, [c_void_p]),
    "pr_host_arena_device_ptr": (c_int, [c_void_p, POINTER(c_void_p)]),
    "pr_copy_d2h_async": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64, c_void_p]),
    "pr_copy_h2d_async": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64, c_void_p]),
    "pr_file_write_atomic": (c_int, [c_char_p, c_void_p, c_void_p, c_int, c_int]),
    "pr_comm_create": (c_int, [c_int, c_int, POINTER(c_void_p)]),
    "pr_comm_create_emulated": (c_int, [c_int, POINTER(c_void_p)]),
    "pr_comm_destroy": (c_int, [c_void_p]),
    "pr_comm_region_alloc": (c_int, [c_void_p, c_int, c_uint64]),
    "pr_comm_region_ptr": (c_int, [c_void_p, c_int, c_int, POINTER(c_void_p)]),
    "pr_comm_export": (c_int, [c_void_p, c_void_p]),
    "pr_comm_import": (c_int, [c_void_p, c_void_p]),
    "pr_comm_error": (c_int, [c_void_p, POINTER(c_int)]),
    "pr_comm_multicast": (c_int, [c_void_p, c_int, POINTER(c_int), POINTER(c_int)]),
    "pr_comm_stats": (c_int, [c_void_p, c_void_p, c_int]),
    "pr_comm_signal_post": (c_int, [c_void_p, ctypes.c_uint32, c_void_p]),
    "pr_comm_signal_read": (c_int, [c_void_p, POINTER(ctypes.c_uint32), POINTER(ctypes.c_uint32)]),
    "pr_comm_seq": (c_int, [c_void_p, POINTER(ctypes.c_uint32)]),
    "pr_comm_test_absent_rank": (c_int, [c_void_p, c_int]),
    "pr_allreduce": (c_int, [c_void_p, c_int, c_uint64, c_uint64, c_int, c_float, c_int, c_int, c_void_p]),
    "pr_allreduce_adamw": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_uint64, c_uint64, c_int,
                                   POINTER(AdamWArgs), c_int, c_void_p]),
    "pr_allgather_slices": (c_int, [c_void_p, c_int, c_uint64, c_uint64, c_int, c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def build(verbose: bool = False) -> Path:
    """Compile the CUDA sources for sm_100a with nvcc (cross-compiles without a GPU)."""
    cmd = ["make", "-C", str(PKG_DIR / "csrc"), "-j4"]
    if not verbose:
        cmd.append("-s")
    subprocess.check_call(cmd)
    if not LIB_PATH.exists():
        raise RuntimeError(f"build finished but {LIB_PATH} is missing")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    """The loaded library; raises RuntimeError (never falls back) when it is not built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} not found: the sm_100a CUDA extension is not built. "
                "Run `make -C pyrecover_b200/csrc` (or __graft_entry__.build()). "
                "pyrecover_b200 has no CPU fallback.")
        L = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)   # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        if L.pr_abi_version() != PR_ABI_VERSION:
            raise RuntimeError(f"ABI version mismatch: library reports {L.pr_abi_version()}")
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    """Turn a negative pr_status into the reference's error convention (RuntimeError)."""
    if rc != 0:
        msg = lib().pr_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what or 'pyrecover_b200'} failed ({rc}): {msg}")


def stream_ptr(stream=None) -> c_void_p:
    """cudaStream_t of a torch stream (default: the current stream) as void*."""
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return c_void_p(s.cuda_stream)


def dtype_code(dtype) -> int:
    import torch
    if dtype == torch.bfloat16:
        return PR_BF16
    if dtype == torch.float32:
        return PR_F32
    raise TypeError(f"pyrecover_b200 supports bf16 and fp32 arenas, not {dtype}")
