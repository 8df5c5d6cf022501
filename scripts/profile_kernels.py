#!/usr/bin/env python
"""Launch every single-GPU kernel of libpyrecover_b200 a few times at the BASELINE configs[1] size
(1 006 700 544 bf16 elements per arena), for an `ncu --set full` capture:

    ncu --set full --clock-control none --import-source on -o gpurun_out/k \\
        -k regex:'adamw|pack_kernel|pack_bulk|gather_grads|checksum|sumsq|norm_finish' python scripts/profile_kernels.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pyrecover_b200 import _lib  # noqa: E402
from pyrecover_b200 import kernels as K  # noqa: E402

n = 1_006_700_544
dev = "cuda"
p = (torch.randn(n, device=dev) * 0.02).bfloat16()
g = (torch.randn(n, device=dev) * 1e-3).bfloat16()
m = torch.zeros_like(p)
v = torch.zeros_like(p)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
gn = K.GradNorm(dev)
s = torch.zeros(2, dtype=torch.int64, device=dev)
image = torch.zeros(3 * n * 2, dtype=torch.uint8, device=dev)
plan = K.PackPlan([p, m, v], [0, 2 * n, 4 * n], 6 * n)
chunk = n // 147 // 8 * 8                 # 16-byte aligned tensor starts, like the flat arena's slots
splan = K.AdamWPlan([i * chunk for i in range(147)], [chunk] * 147, n)
grads = [g[i * chunk:(i + 1) * chunk] for i in range(147)]
garena = torch.zeros(n, dtype=torch.bfloat16, device=dev)
for step in range(1, reps + 1):
    K.adamw_(p, g, m, v, lr=1e-5, step=step)                       # adamw_bf16_kernel<0,false>
    gn.compute(g, 1.0)                                              # sumsq_flat_kernel + norm_finish_kernel
    K.adamw_(p, g, m, v, lr=1e-5, step=step, grad_scale_dev=gn.coef)
    splan.grad_norm(grads, gn, 1.0)                                 # sumsq_scattered_kernel
    splan.step(p, grads, m, v, lr=1e-5, step=step)                  # adamw_bf16_scattered_kernel
    plan.pack(image.data_ptr(), s, algo=_lib.PR_PACK_LDG)           # pack_kernel<0,1>
    plan.unpack(image.data_ptr(), s, algo=_lib.PR_PACK_LDG)         # pack_kernel<1,1>
    plan.pack(image.data_ptr(), s, algo=_lib.PR_PACK_BULK, max_blocks=148)   # pack_bulk_kernel<0,1>, all SMs (run-alone roofline)
    plan.pack(image.data_ptr(), s, algo=_lib.PR_PACK_BULK, max_blocks=8)     # the grid the snapshot uses beside a 25 ms step
    plan.unpack(image.data_ptr(), s, algo=_lib.PR_PACK_BULK, max_blocks=148)
    splan.gather(grads, garena)                                     # gather_grads_kernel
    K.checksum_into(p, s)                                           # checksum_kernel
torch.cuda.synchronize()
print("ok", K.launch_count(), "launches")
