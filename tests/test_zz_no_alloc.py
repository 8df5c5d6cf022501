"""pr_ckpt_pack_noalloc -- the pack entry point for callers that must not allocate device memory
(an emergency checkpoint from the walltime / preemption handler of a job that has filled HBM, or a
process whose memory manager forbids cudaMalloc after start-up).

It has to give exactly what the allocating pair pr_pack_plan_create + pr_ckpt_pack gives: the same
image bytes and the same PR-SUM64 -- and what the CPU oracle gives for the same layout.  The property
it exists for is asserted too: the driver's count of free device memory is the same after a call as
before it (once the first call on the device has loaded the kernel image and the built-in array).
Because its tile table lives in one array shared by every call
and still holding the previous call's entries, the check is made on two calls in a row with
different layouts -- the second one smaller than the first, so stale entries lie right behind the
range it uses.  Collected last on purpose (file name), and it must also pass alone in a fresh
process: nothing here depends on state left by other tests.
"""
import ctypes
from ctypes import c_uint64, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _layout(sizes):
    offs, cur = [], 0
    for s in sizes:
        offs.append(cur)
        cur += (s + 255) // 256 * 256
    return offs, max(cur, 256)


def _tensors(seed, sizes):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, 256, (s,), dtype=torch.uint8, generator=g).cuda() for s in sizes]


def _pack_with_plan(tensors, offs, image_bytes):
    from pyrecover_b200 import kernels as K
    plan = K.PackPlan(tensors, offs, image_bytes)
    image = torch.zeros(image_bytes, dtype=torch.uint8, device="cuda")
    s = torch.zeros(2, dtype=torch.int64, device="cuda")
    plan.pack(image.data_ptr(), s)
    torch.cuda.synchronize()
    plan.close()
    return image, K._sum_to_ints(s)


def _pack_noalloc(lib, tensors, offs, image_bytes, with_sum=True, free_bytes=None):
    """free_bytes: a list that receives cudaMemGetInfo's free byte count right before and right after
    the call (everything torch needs is allocated beforehand)."""
    from pyrecover_b200 import kernels as K
    from pyrecover_b200._lib import stream_ptr
    n = len(tensors)
    ptrs = (c_void_p * n)(*[t.data_ptr() for t in tensors])
    lens = (c_uint64 * n)(*[t.numel() for t in tensors])
    dsts = (c_uint64 * n)(*offs)
    image = torch.zeros(image_bytes, dtype=torch.uint8, device="cuda")
    s = torch.zeros(2, dtype=torch.int64, device="cuda")
    stream = stream_ptr()
    torch.cuda.synchronize()
    if free_bytes is not None:
        free_bytes.append(torch.cuda.mem_get_info()[0])
    rc = lib.pr_ckpt_pack_noalloc(ptrs, lens, dsts, n, image.data_ptr(), image_bytes,
                                  s.data_ptr() if with_sum else None, stream)
    if rc != 0:
        raise RuntimeError(lib.pr_last_error().decode())
    torch.cuda.synchronize()
    if free_bytes is not None:
        free_bytes.append(torch.cuda.mem_get_info()[0])
    return image, K._sum_to_ints(s)


def test_noalloc_pack_matches_plan_pack_twice_in_a_row(cuda_lib, oracle):
    first = [70001, 65536, 17, 0, 3 * 65536 + 40, 1 << 20]      # 6 tensors, ragged tails
    second = [4096, 131072 + 5, 9]                              # fewer tensors and tiles than `first`
    # the warm-up call a caller makes at start-up: loads the kernel image, materialises the array
    warm = _tensors(10, [64])
    _pack_noalloc(cuda_lib, warm, [0], 256)
    for seed, sizes in ((11, first), (12, second)):
        tensors = _tensors(seed, sizes)
        offs, image_bytes = _layout(sizes)
        want_image, want_sum = _pack_with_plan(tensors, offs, image_bytes)
        free = []
        got_image, got_sum = _pack_noalloc(cuda_lib, tensors, offs, image_bytes, free_bytes=free)
        assert torch.equal(got_image, want_image), f"image differs for layout {sizes}"
        assert got_sum == want_sum, f"PR-SUM64 differs for layout {sizes}"
        ref_image, ref_sum = oracle.pack([t.cpu().numpy() for t in tensors], offs, image_bytes)
        assert (got_image.cpu().numpy() == ref_image).all(), f"image differs from the oracle for layout {sizes}"
        assert got_sum == ref_sum, f"PR-SUM64 differs from the oracle for layout {sizes}"
        assert free[0] == free[1], f"device memory changed across the call: {free[0]} -> {free[1]} bytes free"
    # checksum is optional, exactly as in pr_ckpt_pack
    tensors = _tensors(13, second)
    offs, image_bytes = _layout(second)
    img, s = _pack_noalloc(cuda_lib, tensors, offs, image_bytes, with_sum=False)
    assert torch.equal(img, _pack_with_plan(tensors, offs, image_bytes)[0]) and s == (0, 0)


def test_noalloc_pack_rejects_what_does_not_fit(cuda_lib):
    """Beyond the built-in table the call must refuse with a message naming the limit -- it may
    not quietly allocate."""
    small = torch.zeros(16, dtype=torch.uint8, device="cuda")
    n = 1025
    offs, image_bytes = _layout([16] * n)
    with pytest.raises(RuntimeError, match="max 1024"):
        _pack_noalloc(cuda_lib, [small] * n, offs, image_bytes)
    # 65537 tiles: one tensor of 65536 full tiles plus one more byte's worth
    huge = 65536 * 65536 + 1
    ptrs = (c_void_p * 1)(small.data_ptr())          # never dereferenced: the call fails on validation
    lens = (c_uint64 * 1)(huge)
    dsts = (c_uint64 * 1)(0)
    from pyrecover_b200._lib import stream_ptr
    rc = cuda_lib.pr_ckpt_pack_noalloc(ptrs, lens, dsts, 1, small.data_ptr(), huge + 256, None, stream_ptr())
    assert rc != 0 and b"65536 tiles" in cuda_lib.pr_last_error()
    # and the usual layout checks still apply
    with pytest.raises(RuntimeError, match="multiple of 16"):
        _pack_noalloc(cuda_lib, [small], [8], 256)
    # an empty tensor between two overlapping ones must not hide the overlap; a wrapping offset is caught
    big = torch.zeros(100, dtype=torch.uint8, device="cuda")
    empty = torch.zeros(0, dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="overlap"):
        _pack_noalloc(cuda_lib, [big, empty, big], [0, 16, 32], 512)
    ptrs = (c_void_p * 1)(small.data_ptr())
    lens = (c_uint64 * 1)(16)
    dsts = (c_uint64 * 1)(2 ** 64 - 16)
    rc = cuda_lib.pr_ckpt_pack_noalloc(ptrs, lens, dsts, 1, small.data_ptr(), 256, None, stream_ptr())
    assert rc != 0 and b"overruns" in cuda_lib.pr_last_error()
