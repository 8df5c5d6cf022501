#!/usr/bin/env python
"""Small-size pass over the kernels with hand-rolled synchronisation, for compute-sanitizer:

    compute-sanitizer --tool memcheck  python scripts/sanitize_k1.py
    compute-sanitizer --tool racecheck python scripts/sanitize_k1.py

K1's flag protocol runs on an EMULATED communicator (all W ranks as blockIdx.y of one cooperative launch on
one GPU: the same kernels, barriers and flag words as over NVLink); the bulk-copy pack kernel exercises its
mbarrier / cp.async.bulk pipeline (shared memory: what racecheck looks at).  Results are checked against torch.
"""
import os
import sys
from pathlib import Path

import torch

os.environ.setdefault("PR_COMM_TIMEOUT_S", "300")          # kernels run 10-100x slower under the sanitizer
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pyrecover_b200 import _lib  # noqa: E402
from pyrecover_b200 import kernels as K  # noqa: E402

torch.manual_seed(0)
for world in (2, 4):
    n = 8 * 3000
    c = K.Comm(world=world, emulated=True)
    for r in range(4):
        c.alloc_region(r, n * 2)
    c.connect()
    G = [c.region_tensor(0, torch.bfloat16, r) for r in range(world)]
    P = [c.region_tensor(1, torch.bfloat16, r) for r in range(world)]
    loc = [(torch.randn(n) * 1e-3).bfloat16().cuda() for _ in range(world)]
    want = (torch.stack([x.float() for x in loc]).sum(0) / world).bfloat16()
    for algo in (_lib.PR_AR_TWO_SHOT, _lib.PR_AR_ONE_SHOT):
        for g, x in zip(G, loc):
            g.copy_(x)
        c.allreduce(0, 0, n * 2, torch.bfloat16, 1.0 / world, algo=algo, max_blocks=4)
        torch.cuda.synchronize()
        assert all((g.float() - want.float()).abs().max() <= 1e-5 for g in G), "allreduce result off"
    for g, x in zip(G, loc):
        g.copy_(x)
    for p in P:
        p.fill_(0.5)
    args = K.adamw_args(1e-3, 0.9, 0.999, 1e-8, 0.01, step=1, grad_scale=1.0 / world)
    c.signal_post(7)
    c.allreduce_adamw(0, 1, 2, 3, 0, n, args, max_blocks=4)
    c.allgather_slices(2, 0, n * 2, max_blocks=4)
    torch.cuda.synchronize()
    assert all(torch.equal(P[0], p) for p in P) and not c.timed_out() and c.signal_read()[1] == 7
    c.close()
sizes = [70001, 32768, 65536 * 3 + 16, 5, (1 << 20) + 48]
ten = [torch.randint(0, 256, (s,), dtype=torch.uint8).cuda() for s in sizes]
offs, cur = [], 0
for s in sizes:
    offs.append(cur)
    cur += (s + 255) // 256 * 256
plan = K.PackPlan(ten, offs, cur)
img = torch.zeros(cur, dtype=torch.uint8, device="cuda")
img2 = torch.zeros(cur, dtype=torch.uint8, device="cuda")
s1 = torch.zeros(2, dtype=torch.int64, device="cuda")
s2 = torch.zeros(2, dtype=torch.int64, device="cuda")
plan.pack(img.data_ptr(), s1, algo=_lib.PR_PACK_BULK, max_blocks=2)
plan.pack(img2.data_ptr(), s2, algo=_lib.PR_PACK_LDG)
torch.cuda.synchronize()
assert torch.equal(img, img2) and torch.equal(s1, s2)
fresh = [torch.zeros_like(t) for t in ten]
K.PackPlan(fresh, offs, cur).unpack(img.data_ptr(), None, algo=_lib.PR_PACK_BULK, max_blocks=3)
torch.cuda.synchronize()
assert all(torch.equal(a, b) for a, b in zip(fresh, ten))
print("sanitize_k1 ok:", K.launch_count(), "launches")
