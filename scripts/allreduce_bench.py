#!/usr/bin/env python
"""Multi-GPU micro-benchmark + parity check of K1 (P2P allreduce) and the fused K1+K2 kernel
against NCCL (BASELINE.json config 5: bucket sweep at 2/4/8 GPUs vs the reference's DDP/NCCL).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 scripts/allreduce_bench.py [--max-log2 30] [--json out.json]

Timing: CUDA events on the launching stream, 5 warm-ups + 20 timed launches, median, MAX over
ranks.  bus GB/s = N_bytes / t * 2 (W-1) / W (nccl-tests convention); the roofline is the measured
770 GB/s peer copy (900 GB/s nominal) per direction per GPU.
Parity: every rank's result must be bit-identical across ranks and equal to the fp32 rank-order
sum rounded once (computed here with torch in fp32 from all_gathered inputs); NCCL's own result
is compared within the stated bf16 bound.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pyrecover_b200 import _lib  # noqa: E402
from pyrecover_b200 import kernels as K  # noqa: E402


def timed(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class NvlinkCounters:
    """NVLink data-byte counters of one GPU, summed over its links, read through NVML
    (NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX / _RX, KiB).  A link counter the kernels cannot influence: what
    `ncu --section NvLink` would show, without putting a profiler around a multi-rank launch."""

    def __init__(self, index: int):
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.read()
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def read(self):
        nv = self.nv
        vals = nv.nvmlDeviceGetFieldValues(self.h, [(nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF),
                                                    (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                raise RuntimeError(f"NVML field {v.fieldId}: return {v.nvmlReturn}")
            out.append(int(v.value.ullVal) * 1024)
        return out


def link_bytes(counters, fn, launches=10):
    """(tx, rx) bytes per launch of fn on this rank's GPU."""
    torch.cuda.synchronize()
    dist.barrier()
    time.sleep(0.3)                       # the counters are sampled by the driver, not instantaneous
    t0, r0 = counters.read()
    for _ in range(launches):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    time.sleep(0.3)
    t1, r1 = counters.read()
    return (t1 - t0) / launches, (r1 - r0) / launches


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-log2", type=int, default=20)
    ap.add_argument("--max-log2", type=int, default=30)
    ap.add_argument("--json", type=str, default=None)
    ap.add_argument("--max-blocks", type=int, default=0)
    ap.add_argument("--fused-max-log2", type=int, default=31, help="largest size for the fused AR+AdamW rows (4 regions)")
    ap.add_argument("--mc-blocks", type=str, default="32,64,148", help="grid sizes tried for the multicast kernels")
    ap.add_argument("--nvlink-counters", action="store_true", help="NVML link-byte counters around each K1 kernel at 1 GiB")
    ap.add_argument("--skip-nccl-above-log2", type=int, default=40, help="skip the NCCL rows above this size (time)")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    maxb = 1 << a.max_log2
    fmax = 1 << min(a.max_log2, a.fused_max_log2)
    comm = K.Comm(rank, world)
    comm.alloc_region(0, maxb)
    for region in range(1, 4):
        comm.alloc_region(region, fmax)
    comm.connect()
    mc_avail, mc_auto = comm.multicast()
    mc_blocks = [int(x) for x in a.mc_blocks.split(",") if x]
    if rank == 0:
        print(f"world {world}: NVSwitch multicast available={mc_avail}, automatic choice uses it={mc_auto}", flush=True)
    G = comm.region_tensor(0, torch.bfloat16)
    P = comm.region_tensor(1, torch.bfloat16)
    M = comm.region_tensor(2, torch.bfloat16)
    V = comm.region_tensor(3, torch.bfloat16)
    results = {"world": world, "multicast_available": mc_avail, "multicast_auto": mc_auto, "rows": []}
    scale = 1.0 / world

    # ---------------- parity at 8 MiB -----------------------------------------------------------
    n = (8 << 20) // 2
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    local_g = (torch.randn(n, generator=g, device=dev) * 1e-3).bfloat16()
    gathered = [torch.empty_like(local_g) for _ in range(world)]
    dist.all_gather(gathered, local_g)
    acc = torch.zeros(n, dtype=torch.float32, device=dev)
    for t in gathered:                       # rank order 0..W-1, fp32
        acc = acc + t.float()
    expect = (acc * scale).bfloat16()
    for algo, name in ((_lib.PR_AR_TWO_SHOT, "two_shot"), (_lib.PR_AR_ONE_SHOT, "one_shot")):
        nb = n * 2 if algo == _lib.PR_AR_TWO_SHOT else 1 << 20
        G[:n].copy_(local_g)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, 0, nb, torch.bfloat16, scale, algo=algo, max_blocks=a.max_blocks)
        torch.cuda.synchronize()
        ok = torch.equal(G[:nb // 2], expect[:nb // 2])
        flag = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        results[f"parity_{name}"] = bool(flag.item())
        assert flag.item() == 1, f"{name} allreduce differs from the fp32 rank-order sum on some rank"
        dist.barrier()
    S = torch.stack([t.float().abs() for t in gathered]).sum(0) / world
    bound = world * 2.0 ** -9 * S + 1e-30
    if mc_avail:
        # the switch adds the W contributions in its own order and rounds the bf16 sum before the
        # scale: identical on every rank, within the summation-order bound of the rank-order sum
        G[:n].copy_(local_g)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, 0, n * 2, torch.bfloat16, scale, algo=_lib.PR_AR_MULTIMEM)
        torch.cuda.synchronize()
        got = G[:n].clone()
        # bound: one bf16 ulp of the result (the switch's final rounding is not specified to be
        # round-to-nearest-even) + the fp32 summation-order term
        exact = acc * scale
        mc_bound = 2.0 ** -7 * exact.abs() + world * 2.0 ** -23 * S + 1e-30
        err = (got.float() - exact).abs()
        within = bool((err <= mc_bound).all())
        exact = float((got == expect).float().mean())
        ulp = (got.view(torch.int16).int() - expect.view(torch.int16).int())
        toward_zero = float(((got.float().abs() <= (acc * scale).abs()) | (got == expect)).float().mean())
        results["multimem_ulp_diff_hist"] = {str(k): int((ulp == k).sum()) for k in (-2, -1, 0, 1, 2)}
        results["multimem_fraction_not_above_exact_magnitude"] = toward_zero
        if rank == 0:
            bad = (got != expect).nonzero().flatten()[:6].tolist()
            for i in bad:
                print(f"  mismatch @{i}: inputs {[hex(t.view(torch.int16)[i].item() & 0xffff) for t in gathered]} "
                      f"fp32 mean {float(acc[i] * scale):.9e} got {hex(got.view(torch.int16)[i].item() & 0xffff)} "
                      f"rne {hex(expect.view(torch.int16)[i].item() & 0xffff)}", flush=True)
            print(f"  ulp-diff histogram {results['multimem_ulp_diff_hist']}, |got| <= |exact| fraction {toward_zero:.4f}", flush=True)
        allg = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(allg, got)
        same = all(torch.equal(allg[0], t) for t in allg)
        flag = torch.tensor([int(within and same)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        results["parity_multimem_within_bound"] = bool(flag.item())
        results["multimem_identical_across_ranks"] = same
        results["multimem_fraction_bit_equal_to_rank_order_sum"] = exact
        if rank == 0:
            print(f"multimem allreduce: within bound {within}, identical across ranks {same}, bit-equal fraction {exact:.6f}", flush=True)
        assert flag.item() == 1, "multimem allreduce outside the summation-order bound or differing across ranks"
        # run-to-run determinism (resume equality depends on it)
        G[:n].copy_(local_g)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, 0, n * 2, torch.bfloat16, scale, algo=_lib.PR_AR_MULTIMEM, max_blocks=17)
        torch.cuda.synchronize()
        results["multimem_deterministic"] = bool(torch.equal(G[:n], got))
        # fp32 payload
        Gf = comm.region_tensor(0, torch.float32)
        nf = n // 2
        lf = local_g[:nf].float()
        gf = [torch.empty_like(lf) for _ in range(world)]
        dist.all_gather(gf, lf)
        Gf[:nf].copy_(lf)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, 0, nf * 4, torch.float32, scale, algo=_lib.PR_AR_MULTIMEM)
        torch.cuda.synchronize()
        ef = torch.zeros(nf, dtype=torch.float32, device=dev)
        for t in gf:
            ef = ef + t
        ef = ef * scale
        okf = bool(((Gf[:nf] - ef).abs() <= 2.0 ** -22 * world * torch.stack([t.abs() for t in gf]).sum(0) + 1e-30).all())
        results["parity_multimem_f32"] = okf
        assert okf, "fp32 multimem allreduce outside the bound"
        dist.barrier()
    nccl = local_g.clone()
    nccl.mul_(scale)
    dist.all_reduce(nccl)
    within = bool(((nccl.float() - expect.float()).abs() <= bound).all())
    results["nccl_within_bound"] = within      # reported, not asserted: NCCL's accuracy is the reference's, not ours
    if rank == 0 and not within:
        print("note: NCCL's own result is outside W * 2^-9 * mean|g| of the fp32 rank-order sum on this box", flush=True)
    # fused reduce-scatter -> AdamW -> all-gather vs (allreduce -> pr_adamw) on a private copy
    p0 = (torch.randn(n, generator=torch.Generator(device="cuda").manual_seed(7), device=dev) * 0.02).bfloat16()
    G[:n].copy_(local_g); P[:n].copy_(p0); M[:n].zero_(); V[:n].zero_()
    torch.cuda.synchronize()
    dist.barrier()
    args = K.adamw_args(1e-4, 0.9, 0.999, 1e-8, 0.01, step=1, grad_scale=scale)
    comm.multicast(0)                               # P2P kernels: bit-exact against the rank-order sum
    comm.allreduce_adamw(0, 1, 2, 3, 0, n, args, max_blocks=a.max_blocks)
    comm.allgather_slices(2, 0, n * 2)
    comm.allgather_slices(3, 0, n * 2)
    torch.cuda.synchronize()
    pr, mr, vr = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    K.adamw_(pr, acc, mr, vr, lr=1e-4, step=1, grad_scale=scale)     # fp32 rank-order sum as gradient
    ok = torch.equal(P[:n], pr) and torch.equal(M[:n], mr) and torch.equal(V[:n], vr)
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    results["parity_fused"] = bool(flag.item())
    assert flag.item() == 1, "fused allreduce+AdamW differs from allreduce -> AdamW"
    dist.barrier()
    if mc_avail:
        # multicast fused kernel == (multimem allreduce -> pr_adamw on the result), bit for bit
        comm.multicast(1)
        G[:n].copy_(local_g)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, 0, n * 2, torch.bfloat16, 1.0, algo=_lib.PR_AR_MULTIMEM)
        torch.cuda.synchronize()
        gsum = G[:n].clone()
        pr, mr, vr = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
        K.adamw_(pr, gsum, mr, vr, lr=1e-4, step=1, grad_scale=scale)
        G[:n].copy_(local_g); P[:n].copy_(p0); M[:n].zero_(); V[:n].zero_()
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_adamw(0, 1, 2, 3, 0, n, args, max_blocks=a.max_blocks)
        comm.allgather_slices(2, 0, n * 2)
        comm.allgather_slices(3, 0, n * 2)
        torch.cuda.synchronize()
        ok = torch.equal(P[:n], pr) and torch.equal(M[:n], mr) and torch.equal(V[:n], vr)
        flag = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        results["parity_fused_multimem"] = bool(flag.item())
        if rank == 0:
            print(f"fused multimem AR+AdamW == multimem AR -> AdamW: {bool(flag.item())}", flush=True)
        assert flag.item() == 1, "multicast fused allreduce+AdamW differs from multimem allreduce -> AdamW"
        dist.barrier()
    comm.multicast(-1)

    # ---------------- NVLink byte counters (NVML) around each kernel ---------------------------------
    if a.nvlink_counters:
        ctr = NvlinkCounters(local)
        nb = min(1 << 30, maxb, fmax)
        ne_ = nb // 2
        rows = {}
        if ctr.ok:
            kernels = {"allreduce_twoshot_kernel (P2P)": (lambda: comm.allreduce(0, 0, nb, torch.bfloat16, scale, algo=_lib.PR_AR_TWO_SHOT),
                                                          "2 N (W-1)/W per direction", 2 * nb * (world - 1) / world)}
            if mc_avail:
                kernels["allreduce_mc_kernel (multicast)"] = (lambda: comm.allreduce(0, 0, nb, torch.bfloat16, scale, algo=_lib.PR_AR_MULTIMEM),
                                                              "N (1 + 1/W) per direction (reads of the own slice from W-1 peers arrive "
                                                              "reduced; one multicast store leaves)", nb * (1 + 1 / world))
                st = [0]

                def fused_mc():
                    st[0] += 1
                    comm.allreduce_adamw(0, 1, 2, 3, 0, ne_, K.adamw_args(1e-5, 0.9, 0.999, 1e-8, 0.01, st[0], scale), max_blocks=32)
                comm.multicast(1)
                kernels["allreduce_adamw_mc_kernel (fused, multicast)"] = (fused_mc, "N (1 + 1/W) per direction", nb * (1 + 1 / world))
            nccl_buf = torch.empty(ne_, dtype=torch.bfloat16, device=dev).normal_(0, 1e-3)
            kernels["ncclAllReduce (reference's DDP)"] = (lambda: dist.all_reduce(nccl_buf), "algorithm-dependent (ring: 2 N (W-1)/W; NVLS: ~N (1+1/W))", None)
            for name, (fn, formula, predicted) in kernels.items():
                for _ in range(3):
                    fn()
                tx, rx = link_bytes(ctr, fn)
                rows[name] = {"bytes_allreduced": nb, "tx_bytes_per_launch": tx, "rx_bytes_per_launch": rx, "expected": formula,
                              "expected_bytes_per_direction": predicted}
                if rank == 0:
                    print(f"NVLink counters, {name}: tx {tx / 1e9:.3f} GB rx {rx / 1e9:.3f} GB per launch of {nb / 1e9:.3f} GB "
                          f"(expected per direction: {predicted / 1e9 if predicted else float('nan'):.3f} GB)", flush=True)
            comm.multicast(-1)
            del nccl_buf
        else:
            rows = {"unavailable": getattr(ctr, "err", "NVML NVLink counters not readable")}
            if rank == 0:
                print("NVLink counters unavailable:", rows["unavailable"], flush=True)
        results["nvlink_counters_rank0"] = rows

    # ---------------- bandwidth sweep -----------------------------------------------------------
    for lg in range(a.min_log2, a.max_log2 + 1):
        nbytes = 1 << lg
        ne = nbytes // 2
        buf = torch.empty(ne, dtype=torch.bfloat16, device=dev).normal_(0, 1e-3)
        row = {"bytes": nbytes}
        t = timed(lambda: comm.allreduce(0, 0, nbytes, torch.bfloat16, scale, algo=_lib.PR_AR_TWO_SHOT, max_blocks=a.max_blocks))
        row["p2p_ms"], row["p2p_busbw"] = t, nbytes / t / 1e6 * 2 * (world - 1) / world
        if mc_avail:
            for nb in mc_blocks:
                t = timed(lambda: comm.allreduce(0, 0, nbytes, torch.bfloat16, scale, algo=_lib.PR_AR_MULTIMEM, max_blocks=nb))
                row[f"mc{nb}_ms"], row[f"mc{nb}_busbw"] = t, nbytes / t / 1e6 * 2 * (world - 1) / world
            best = min(mc_blocks, key=lambda nb: row[f"mc{nb}_ms"])
            row["mc_ms"], row["mc_busbw"], row["mc_best_blocks"] = row[f"mc{best}_ms"], row[f"mc{best}_busbw"], best
        if nbytes <= (4 << 20):
            t = timed(lambda: comm.allreduce(0, 0, nbytes, torch.bfloat16, scale, algo=_lib.PR_AR_ONE_SHOT))
            row["oneshot_ms"] = t
        if lg <= a.skip_nccl_above_log2:
            t = timed(lambda: dist.all_reduce(buf))
        else:
            t = float("nan")
        row["nccl_ms"], row["nccl_busbw"] = t, nbytes / t / 1e6 * 2 * (world - 1) / world
        step = [0]
        if nbytes > fmax:
            results["rows"].append(row)
            if rank == 0:
                print(f"{nbytes >> 20:6d} MiB  p2p {row['p2p_ms']:8.3f} ms {row['p2p_busbw']:7.1f} GB/s | "
                      + (f"multimem {row['mc_ms']:8.3f} ms {row['mc_busbw']:7.1f} GB/s ({row['mc_best_blocks']} CTAs) | " if mc_avail else "")
                      + f"nccl {row['nccl_ms']:8.3f} ms {row['nccl_busbw']:7.1f} GB/s", flush=True)
            del buf
            continue

        def fused():
            step[0] += 1
            comm.allreduce_adamw(0, 1, 2, 3, 0, ne, K.adamw_args(1e-5, 0.9, 0.999, 1e-8, 0.01, step[0], scale),
                                 max_blocks=a.max_blocks)
        comm.multicast(0)
        t = timed(fused)
        row["fused_ms"], row["fused_busbw"] = t, nbytes / t / 1e6 * 2 * (world - 1) / world
        if mc_avail:
            comm.multicast(1)
            t = timed(fused)
            row["fused_mc_ms"], row["fused_mc_busbw"] = t, nbytes / t / 1e6 * 2 * (world - 1) / world
        comm.multicast(-1)
        pl, ml, vl = (torch.zeros(ne, dtype=torch.bfloat16, device=dev) for _ in range(3))

        def unfused():
            step[0] += 1
            dist.all_reduce(buf)
            K.adamw_(pl, buf, ml, vl, lr=1e-5, step=step[0], grad_scale=scale)
        t = timed(unfused)
        row["nccl_plus_adamw_ms"] = t
        results["rows"].append(row)
        if rank == 0:
            print(f"{nbytes >> 20:6d} MiB  p2p {row['p2p_ms']:8.3f} ms {row['p2p_busbw']:7.1f} GB/s | "
                  + (f"multimem {row['mc_ms']:8.3f} ms {row['mc_busbw']:7.1f} GB/s ({row['mc_best_blocks']} CTAs) | " if mc_avail else "")
                  + f"nccl {row['nccl_ms']:8.3f} ms {row['nccl_busbw']:7.1f} GB/s | fused AR+AdamW {row['fused_ms']:8.3f} ms | "
                  + (f"fused multimem {row['fused_mc_ms']:8.3f} ms | " if mc_avail else "")
                  + f"nccl+adamw {row['nccl_plus_adamw_ms']:8.3f} ms", flush=True)
        del buf, pl, ml, vl
    assert not comm.timed_out()
    if rank == 0 and a.json:
        Path(a.json).parent.mkdir(parents=True, exist_ok=True)
        Path(a.json).write_text(json.dumps(results, indent=1))
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
