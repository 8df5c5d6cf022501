"""Tensor-level wrappers over the C ABI (include/pyrecover_b200.h).

Everything here takes torch CUDA tensors only for their device pointers and the current
stream; the arithmetic is in the sm_100a kernels.  CPU tensors are rejected: there is no
CPU fallback on the product path.
"""
from __future__ import annotations

import ctypes
from ctypes import c_int, c_uint64, c_void_p
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import AdamWArgs, check, dtype_code, lib, stream_ptr


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("pyrecover_b200 kernels need CUDA tensors (no CPU fallback)")
        if not t.is_contiguous():
            raise RuntimeError("pyrecover_b200 kernels need contiguous tensors")


def device_info():
    sm, major, minor = c_int(), c_int(), c_int()
    check(lib().pr_device_info(ctypes.byref(sm), ctypes.byref(major), ctypes.byref(minor)), "pr_device_info")
    return sm.value, major.value, minor.value


def launch_count() -> int:
    """Kernels launched by libpyrecover_b200 in this process so far."""
    return int(lib().pr_launch_count())


# ------------------------------------------------------------------------- K2 AdamW -----

def adamw_args(lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0, numerics=0,
               grad_dtype=_lib.PR_BF16, grad_scale_dev: Optional[torch.Tensor] = None) -> AdamWArgs:
    """grad_scale_dev: a 1-element fp32 CUDA tensor (e.g. GradNorm.coef) multiplied into grad_scale on
    the device when the kernel starts; the caller keeps it alive until the launch has run."""
    ptr = None
    if grad_scale_dev is not None:
        if not (grad_scale_dev.is_cuda and grad_scale_dev.dtype == torch.float32 and grad_scale_dev.numel() == 1):
            raise TypeError("grad_scale_dev must be a 1-element fp32 CUDA tensor")
        ptr = grad_scale_dev.data_ptr()
    return AdamWArgs(float(lr), float(beta1), float(beta2), float(eps), float(weight_decay),
                     float(grad_scale), int(step), int(numerics), int(grad_dtype), ptr)


def adamw_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, *, lr: float,
           beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.01,
           step: int = 1, grad_scale: float = 1.0, numerics: int = 0, stream=None,
           grad_scale_dev: Optional[torch.Tensor] = None) -> None:
    """One flat AdamW update, in place on p, m, v (reference train.py:274)."""
    _require_cuda(p, g, m, v)
    n = p.numel()
    if not (g.numel() == n and m.numel() == n and v.numel() == n):
        raise ValueError("p, g, m, v must have the same number of elements")
    if m.dtype != p.dtype or v.dtype != p.dtype:
        raise TypeError("m and v must have the dtype of p")
    a = adamw_args(lr, beta1, beta2, eps, weight_decay, step, grad_scale, numerics, dtype_code(g.dtype), grad_scale_dev)
    check(lib().pr_adamw(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n, dtype_code(p.dtype),
                         ctypes.byref(a), stream_ptr(stream)), "pr_adamw")


class GradNorm:
    """Device-side result of pr_grad_norm: out = [total norm, clip coefficient, non-finite flag]
    (fp32) plus the caller-owned scratch the two launches need.  Nothing here synchronises."""

    def __init__(self, device):
        self.out = torch.zeros(4, dtype=torch.float32, device=device)
        self.scratch = torch.zeros(_lib.PR_GRAD_NORM_SCRATCH, dtype=torch.float64, device=device)

    @property
    def norm(self) -> torch.Tensor:
        return self.out[0:1]

    @property
    def coef(self) -> torch.Tensor:
        return self.out[1:2]

    @property
    def nonfinite(self) -> torch.Tensor:
        return self.out[2:3]

    def compute(self, g: torch.Tensor, max_norm: float, prescale: float = 1.0, stream=None) -> None:
        """Flat gradient buffer (pr_grad_norm; reference utils.py:84-89)."""
        _require_cuda(g)
        check(lib().pr_grad_norm(g.data_ptr(), g.numel(), dtype_code(g.dtype), float(prescale), float(max_norm),
                                 self.out.data_ptr(), self.scratch.data_ptr(), stream_ptr(stream)), "pr_grad_norm")


class AdamWPlan:
    """Layout of a flat arena for the scattered-gradient K2 (pr_adamw_plan)."""

    def __init__(self, elem_offsets: Sequence[int], numels: Sequence[int], arena_elems: int):
        n = len(numels)
        self.n = n
        offs = (c_uint64 * n)(*[int(o) for o in elem_offsets])
        lens = (c_uint64 * n)(*[int(x) for x in numels])
        self._h = c_void_p()
        check(lib().pr_adamw_plan_create(offs, lens, n, int(arena_elems), ctypes.byref(self._h)), "pr_adamw_plan_create")
        self._ptrs = (c_void_p * n)()

    def grad_norm(self, grads: Sequence[torch.Tensor], norm: "GradNorm", max_norm: float, prescale: float = 1.0,
                  stream=None) -> None:
        """Global L2 norm + clip coefficient of the gradients where autograd left them (pr_grad_norm_scattered)."""
        for i, g in enumerate(grads):
            self._ptrs[i] = g.data_ptr()
        check(lib().pr_grad_norm_scattered(self._h, self._ptrs, float(prescale), float(max_norm), norm.out.data_ptr(),
                                           norm.scratch.data_ptr(), stream_ptr(stream)), "pr_grad_norm_scattered")

    def step(self, p: torch.Tensor, grads: Sequence[torch.Tensor], m: torch.Tensor, v: torch.Tensor, *, lr, beta1=0.9,
             beta2=0.999, eps=1e-8, weight_decay=0.01, step=1, grad_scale=1.0, numerics=0, stream=None,
             grad_scale_dev: Optional[torch.Tensor] = None) -> None:
        """grads[i]: contiguous, 16-B aligned bf16 CUDA tensor holding tensor i's gradient."""
        _require_cuda(p, m, v)
        for i, g in enumerate(grads):
            self._ptrs[i] = g.data_ptr()
        a = adamw_args(lr, beta1, beta2, eps, weight_decay, step, grad_scale, numerics, _lib.PR_BF16, grad_scale_dev)
        check(lib().pr_adamw_scattered(self._h, p.data_ptr(), self._ptrs, m.data_ptr(), v.data_ptr(), ctypes.byref(a),
                                       stream_ptr(stream)), "pr_adamw_scattered")

    def gather(self, grads: Sequence[Optional[torch.Tensor]], g_arena: torch.Tensor, seg_lo: int = 0,
               seg_hi: Optional[int] = None, max_blocks: int = 0, stream=None) -> None:
        """Copy grads[seg_lo:seg_hi] (where autograd left them) into their slots of the flat gradient arena
        (pr_gather_grads).  `grads` is indexed by tensor; entries outside the range are not looked at."""
        seg_hi = self.n if seg_hi is None else seg_hi
        for i in range(seg_lo, seg_hi):
            self._ptrs[i] = grads[i].data_ptr()
        check(lib().pr_gather_grads(self._h, self._ptrs, seg_lo, seg_hi, g_arena.data_ptr(), max_blocks,
                                    stream_ptr(stream)), "pr_gather_grads")

    def close(self) -> None:
        if self._h:
            lib().pr_adamw_plan_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------- K3 checksum --

def checksum_into(t: torch.Tensor, sum_dev: torch.Tensor, word_base: int = 0, stream=None) -> None:
    """sum_dev (2 x int64 on the device, zeroed by the caller) += PR-SUM64(t's bytes)."""
    _require_cuda(t, sum_dev)
    assert sum_dev.dtype == torch.int64 and sum_dev.numel() == 2
    check(lib().pr_checksum(t.data_ptr(), t.numel() * t.element_size(), word_base, sum_dev.data_ptr(),
                            stream_ptr(stream)), "pr_checksum")


def _sum_to_ints(sum_dev: torch.Tensor):
    a, b = sum_dev.tolist()
    return a & (2**64 - 1), b & (2**64 - 1)


def checksum(t: torch.Tensor, word_base: int = 0):
    """PR-SUM64 of a CUDA tensor's bytes -> (A, B) python ints (synchronises)."""
    s = torch.zeros(2, dtype=torch.int64, device=t.device)
    checksum_into(t, s, word_base)
    return _sum_to_ints(s)


class PackPlan:
    """Static layout: n CUDA tensors -> one packed image (pr_pack_plan)."""

    def __init__(self, tensors: Sequence[torch.Tensor], offsets: Sequence[int], image_bytes: int):
        _require_cuda(*tensors)
        self.tensors = list(tensors)          # keeps the memory alive
        self.offsets = [int(o) for o in offsets]
        self.nbytes = [t.numel() * t.element_size() for t in tensors]
        self.image_bytes = int(image_bytes)
        n = len(tensors)
        ptrs = (c_void_p * n)(*[t.data_ptr() for t in tensors])
        lens = (c_uint64 * n)(*self.nbytes)
        offs = (c_uint64 * n)(*self.offsets)
        self._h = c_void_p()
        check(lib().pr_pack_plan_create(ptrs, lens, offs, n, self.image_bytes, ctypes.byref(self._h)),
              "pr_pack_plan_create")

    def pack(self, image_ptr: int, sum_dev: Optional[torch.Tensor] = None, stream=None,
             algo: int = _lib.PR_PACK_AUTO, max_blocks: int = 0) -> None:
        """algo: PR_PACK_AUTO (bulk-copy kernel when every tensor is 16-B aligned and the image is device
        memory, else the LDG/STG kernel), PR_PACK_LDG, PR_PACK_BULK; max_blocks bounds the grid (0 = default)."""
        check(lib().pr_ckpt_pack_ex(self._h, image_ptr, sum_dev.data_ptr() if sum_dev is not None else None,
                                    algo, max_blocks, stream_ptr(stream)), "pr_ckpt_pack")

    def unpack(self, image_ptr: int, sum_dev: Optional[torch.Tensor] = None, stream=None,
               algo: int = _lib.PR_PACK_AUTO, max_blocks: int = 0) -> None:
        check(lib().pr_ckpt_unpack_ex(self._h, image_ptr, sum_dev.data_ptr() if sum_dev is not None else None,
                                      algo, max_blocks, stream_ptr(stream)), "pr_ckpt_unpack")

    def close(self) -> None:
        if self._h:
            lib().pr_pack_plan_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostArena:
    """Pinned, device-mapped, zero-filled host buffer (pr_host_arena_*)."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        p = c_void_p()
        check(lib().pr_host_arena_alloc(self.nbytes, ctypes.byref(p)), "pr_host_arena_alloc")
        self.ptr = p.value or 0
        self._buf = (ctypes.c_char * self.nbytes).from_address(self.ptr) if self.nbytes else None

    def view(self, offset: int = 0, nbytes: Optional[int] = None) -> memoryview:
        nbytes = self.nbytes - offset if nbytes is None else nbytes
        return memoryview(self._buf).cast("B")[offset:offset + nbytes]

    def device_ptr(self) -> int:
        d = c_void_p()
        check(lib().pr_host_arena_device_ptr(self.ptr, ctypes.byref(d)), "pr_host_arena_device_ptr")
        return d.value

    def close(self) -> None:
        if self.ptr:
            self._buf = None
            check(lib().pr_host_arena_free(self.ptr), "pr_host_arena_free")
            self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def copy_d2h_async(dst_host_ptr: int, src: torch.Tensor, nbytes: int, chunk_bytes: int = 0, stream=None):
    check(lib().pr_copy_d2h_async(dst_host_ptr, src.data_ptr(), nbytes, chunk_bytes, stream_ptr(stream)),
          "pr_copy_d2h_async")


def copy_h2d_async(dst: torch.Tensor, src_host_ptr: int, nbytes: int, chunk_bytes: int = 0, stream=None):
    check(lib().pr_copy_h2d_async(dst.data_ptr(), src_host_ptr, nbytes, chunk_bytes, stream_ptr(stream)),
          "pr_copy_h2d_async")


def file_write_atomic(path: str, pieces: Sequence[tuple], fsync: bool = True) -> None:
    """pieces: (host_ptr, nbytes) pairs written back to back to `path` (tmp + rename)."""
    n = len(pieces)
    ptrs = (c_void_p * n)(*[p for p, _ in pieces])
    lens = (c_uint64 * n)(*[b for _, b in pieces])
    check(lib().pr_file_write_atomic(str(path).encode(), ptrs, lens, n, 1 if fsync else 0),
          "pr_file_write_atomic")


# ------------------------------------------------------------------------- K1 comm ------

class _DevMem:
    """__cuda_array_interface__ holder so torch can view library-owned device memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, nbytes: int, dtype: torch.dtype, device, owner=None) -> torch.Tensor:
    t = torch.as_tensor(_DevMem(ptr, nbytes, owner), device=device)
    return t.view(dtype)


class Comm:
    """P2P communicator (pr_comm).  `pg` is a torch.distributed group used ONLY to exchange
    the IPC export blobs (bootstrap); emulated=True runs all ranks on the current device."""

    def __init__(self, rank: int = 0, world: int = 1, emulated: bool = False):
        self.rank, self.world, self.emulated = rank, world, emulated
        self._h = c_void_p()
        if emulated:
            check(lib().pr_comm_create_emulated(world, ctypes.byref(self._h)), "pr_comm_create_emulated")
        else:
            check(lib().pr_comm_create(rank, world, ctypes.byref(self._h)), "pr_comm_create")
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.region_bytes = {}

    def alloc_region(self, region: int, nbytes: int) -> None:
        check(lib().pr_comm_region_alloc(self._h, region, nbytes), "pr_comm_region_alloc")
        self.region_bytes[region] = nbytes

    def region_tensor(self, region: int, dtype: torch.dtype, rank: int = 0) -> torch.Tensor:
        p = c_void_p()
        check(lib().pr_comm_region_ptr(self._h, region, rank, ctypes.byref(p)), "pr_comm_region_ptr")
        return tensor_from_ptr(p.value, self.region_bytes[region], dtype, self.device, owner=self)

    def connect(self, pg=None) -> None:
        """Exchange export blobs over torch.distributed (bootstrap only) and map the peers."""
        if self.emulated or self.world == 1:
            if not self.emulated:
                blob = ctypes.create_string_buffer(_lib.PR_EXPORT_BLOB_BYTES)
                check(lib().pr_comm_export(self._h, blob), "pr_comm_export")
                check(lib().pr_comm_import(self._h, blob), "pr_comm_import")
            return
        import torch.distributed as dist
        blob = ctypes.create_string_buffer(_lib.PR_EXPORT_BLOB_BYTES)
        check(lib().pr_comm_export(self._h, blob), "pr_comm_export")
        mine = torch.frombuffer(bytearray(blob.raw), dtype=torch.uint8)
        backend = dist.get_backend(pg)
        dev = self.device if "nccl" in str(backend) else torch.device("cpu")
        mine = mine.to(dev)
        gathered = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(gathered, mine, group=pg)
        allb = b"".join(bytes(t.cpu().numpy().tobytes()) for t in gathered)
        check(lib().pr_comm_import(self._h, allb), "pr_comm_import")

    def allreduce(self, region: int, offset: int, nbytes: int, dtype: torch.dtype, scale: float,
                  algo: int = _lib.PR_AR_AUTO, max_blocks: int = 0, stream=None) -> None:
        check(lib().pr_allreduce(self._h, region, offset, nbytes, dtype_code(dtype), scale, algo,
                                 max_blocks, stream_ptr(stream)), "pr_allreduce")

    def allreduce_adamw(self, grad_region: int, param_region: int, m_region: int, v_region: int,
                        elem_offset: int, n: int, args: AdamWArgs, max_blocks: int = 0, stream=None) -> None:
        check(lib().pr_allreduce_adamw(self._h, grad_region, param_region, m_region, v_region, elem_offset,
                                       n, _lib.PR_BF16, ctypes.byref(args), max_blocks, stream_ptr(stream)),
              "pr_allreduce_adamw")

    def allgather_slices(self, region: int, offset: int, nbytes: int, max_blocks: int = 0, stream=None):
        check(lib().pr_allgather_slices(self._h, region, offset, nbytes, max_blocks, stream_ptr(stream)),
              "pr_allgather_slices")

    def multicast(self, policy: int = -2):
        """NVSwitch multicast state -> (available, in_use).  policy -2 only queries; -1 automatic
        (world >= 4), 0 off, 1 on whenever available (pr_comm_multicast; same choice on every rank)."""
        av, use = c_int(), c_int()
        check(lib().pr_comm_multicast(self._h, policy, ctypes.byref(av), ctypes.byref(use)), "pr_comm_multicast")
        return bool(av.value), bool(use.value)

    def stats(self, reset: bool = False):
        """(ns CTA 0 waited at begin barriers, launches, ns CTA 0 ran) -- synchronises (pr_comm_stats)."""
        out = (c_uint64 * 3)()
        check(lib().pr_comm_stats(self._h, out, 1 if reset else 0), "pr_comm_stats")
        return int(out[0]), int(out[1]), int(out[2])

    def signal_post(self, value: int, stream=None) -> None:
        check(lib().pr_comm_signal_post(self._h, int(value) & 0xffffffff, stream_ptr(stream)), "pr_comm_signal_post")

    def signal_read(self):
        """(launch sequence number, rank 0's word) as last forwarded by a collective's begin barrier; no GPU sync."""
        seq, val = ctypes.c_uint32(), ctypes.c_uint32()
        check(lib().pr_comm_signal_read(self._h, ctypes.byref(seq), ctypes.byref(val)), "pr_comm_signal_read")
        return seq.value, val.value

    def seq(self) -> int:
        s = ctypes.c_uint32()
        check(lib().pr_comm_seq(self._h, ctypes.byref(s)), "pr_comm_seq")
        return s.value

    def test_absent_rank(self, rank: int) -> None:
        check(lib().pr_comm_test_absent_rank(self._h, rank), "pr_comm_test_absent_rank")

    def timed_out(self) -> bool:
        e = c_int()
        check(lib().pr_comm_error(self._h, ctypes.byref(e)), "pr_comm_error")
        return bool(e.value)

    def close(self) -> None:
        if self._h:
            lib().pr_comm_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
