#!/usr/bin/env python
"""bench.py -- tokens/s of the data-parallel training step on BASELINE.json's configs[1]
(~1 B-parameter model.py-style transformer, bf16, seq 2048, batch 1 / GPU, fused AdamW).

    python bench.py --gpus N --steps K --warmup W              # this repo's arm
    python bench.py --impl reference --gpus N --steps K ...    # the reference's CPU path

For N > 1 the driver launches it under torch.distributed.run, one rank per GPU.

A step = one pass of the hot path over one synthetic batch: forward + backward of the workload
model (library kernels: the dense model is out of scope), the gradient exchange (K1, N > 1), the
flat AdamW update (K2) and -- once, in the middle of the timed region -- a checkpoint snapshot
(K3 pack+checksum, async D2H and file write).  One untimed snapshot is taken after the warm-up steps, as
the warm-up of that path (first-use costs of a process belong to job start-up).
  value : whole-job tokens/s with the batches already resident in HBM (CUDA-event timed)
  e2e   : the same through Trainer.step_from_host (pinned-host batch -> H2D inside, loss D2H every
          step), wall-clock between device synchronisations
  roofline : N = 1: K2 against HBM (14 B per bf16 parameter / mean CUDA-event duration of the launches
          inside the timed region).  N > 1: the exchange+update launches against NVLink (bus bytes =
          N_bytes * 2(W-1)/W per step / the time CTA 0 of those launches ran, read from the library's own
          device-side clock -- pr_comm_stats -- which also separates the wait for the slowest rank)
  ckpt  : checkpoint stall = median over 5 blocks WITH a snapshot - median over 6 interleaved blocks without
  parity (N > 1): the real-NVLink kernels checked against the CPU oracle on seeded inputs, after timing
  incumbent : stock torch DDP (NCCL) + torch.optim.AdamW(fused=True) + torch.save on the same workload
  cpu_baseline : the reference's CPU path on this box's host cores, bounded token sample (N = 1 only)
"""
import argparse
import json
import os
import shutil
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "tokens/s"
WORKLOADS = {
    "1b": "~1B model.py transformer (dim 2048, 16 layers, vocab 32768), bf16, seq 2048, batch 1/GPU, fused AdamW [BASELINE configs[1]]",
    "8b": "default model.py transformer (dim 4096, 32 layers, vocab 131072, ~8B params), bf16, seq 2048, batch 1/GPU, fused AdamW [BASELINE configs[2]]",
    "tiny": "tiny model.py transformer (dim 256, 4 layers, vocab 1024), bf16, batch 1/GPU, fused AdamW [BASELINE configs[0] shape, on the GPU path]",
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    ap.add_argument("--preset", type=str, default="1b", choices=["tiny", "1b", "8b"])
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--grad-sync", type=str, default=None, choices=[None, "replicated", "fused", "overlap"],
                    help="default: fused below 4 GPUs (the P2P kernels need every SM to fill the links, so they run after "
                         "backward), overlap from 4 up (the NVSwitch-multicast kernels need a few dozen CTAs and run per "
                         "bucket under backward)")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around timed leg 1 (for ncu --profile-from-start off)")
    ap.add_argument("--no-ckpt", action="store_true", help="skip the in-region checkpoint snapshot and the stall measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-incumbent", action="store_true", help="skip the stock DDP/NCCL + torch AdamW + torch.save arm")
    ap.add_argument("--no-parity", action="store_true", help="skip the N > 1 oracle parity check of the NVLink kernels")
    ap.add_argument("--stall-blocks", type=int, default=5, help="blocks WITH a checkpoint in the stall measurement")
    ap.add_argument("--cpu-budget-s", type=float, default=8.0, help="target seconds per CPU baseline step")
    ap.add_argument("--ckpt-dir", type=str, default=None)
    ap.add_argument("--ckpt-format", type=str, default=None, choices=[None, "vanilla", "sharded"],
                    help="default: vanilla at N=1 (rank 0 writes one torch-loadable file), sharded at N>1 (every rank writes 1/N)")
    return ap.parse_args()


# ------------------------------------------------------------------------- clocks ---------
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv is not None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def _median(xs):
    s = sorted(xs)
    n = len(s)
    return 0.5 * (s[(n - 1) // 2] + s[n // 2]) if n else None


# ------------------------------------------------------------------------- reference arm --
def run_reference(a):
    """The reference's own CPU implementation of the path on this box's host cores (oracle/_ref when the
    reference's files were staged at build time, else the port), rank 0 only.  At N > 1: W gloo ranks running
    DistributedDataParallel, the cores split between them -- the same weak-scaled workload as our arm."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpu_reference_port as port
    world = max(1, a.gpus)
    steps, warmup = max(1, a.steps), (1 if world > 1 else max(1, min(a.warmup, 2)))
    tmp = a.ckpt_dir or tempfile.mkdtemp(prefix="pr_ref_")
    unavailable = None
    if world > 1:
        try:
            import psutil
            need = world * 12.0 * _param_count(a.preset)          # p, g, DDP bucket, m, v in bf16 + slack
            if psutil.virtual_memory().available < need * 1.3:
                unavailable = (f"{world} CPU DDP ranks of the {a.preset} model need ~{need / 1e9:.0f} GB of host memory, "
                               f"{psutil.virtual_memory().available / 1e9:.0f} GB available")
        except ImportError:
            pass
    if unavailable:
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": None, "unit": "tokens/s", "n_gpus": a.gpus,
                          "unavailable": unavailable, "config": {"workload": WORKLOADS[a.preset]}}), flush=True)
        return 0
    try:
        if world == 1:
            res = port.run(a.preset, a.seq, steps=steps, warmup=warmup, fused=True, budget_s_per_step=a.cpu_budget_s,
                           ckpt_path=os.path.join(tmp, "ckpt_ref.pt"))
        else:
            res = port.run_ddp(a.preset, a.seq, steps=min(steps, 8), warmup=1, world=world, fused=True,
                               budget_s_per_step=a.cpu_budget_s)      # every step carries a gloo all-reduce of the model: keep it to minutes
    except Exception as e:  # noqa: BLE001 -- the arm reports, it never takes the driver's run down
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": None, "unit": "tokens/s", "n_gpus": a.gpus,
                          "unavailable": f"{type(e).__name__}: {e}", "config": _config(a.preset, None, world, a.seq)}), flush=True)
        return 0
    line = {
        "impl": "reference", "metric": METRIC, "value": res["tokens_per_s"], "unit": "tokens/s", "n_gpus": a.gpus,
        "steps": res["steps"], "warmup": warmup, "ms_per_step": res["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": _config(a.preset, None, world, a.seq),
        "device": "host CPU",
        "cpu_baseline": {"value": res["tokens_per_s"], "unit": "tokens/s", "cores": res["cores"], "kind": res["kind"],
                         "sample": res["sample"]},
        "e2e": {"value": res["tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ckpt": {"store_s": res.get("ckpt_store_s"), "bytes": res.get("ckpt_bytes"),
                 "gbs": (res["ckpt_bytes"] / res["ckpt_store_s"] / 1e9) if res.get("ckpt_store_s") else None,
                 "note": "the reference's save_ckpt_vanilla (torch.save of {model, optimizer, lr_scheduler}), synchronous, "
                         "outside the timed steps"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


_PARAMS = {"tiny": 4_065_536, "1b": 1_006_700_544, "8b": 8_053_329_920}


def _param_count(preset):
    return _PARAMS[preset]


def _config(preset, n_params, world, seq):
    """The workload description both arms print (the driver compares it between them)."""
    return {"workload": WORKLOADS[preset], "params": n_params or _PARAMS[preset], "global_batch": world, "seq_len": seq,
            "parallelism": f"dp{world}"}


# ------------------------------------------------------------------------- parity (N > 1) -
def parity_check(trainer, world, rank, device):
    """Oracle check of the kernels that actually carry the gradients at this world size, on real NVLink:
    (a) pr_allreduce as PR_AR_AUTO picks it (P2P two-shot below W = 4, NVSwitch multicast from there),
    (b) the exchange+update exactly as GradSync launches it (same entry point, regions and grid).
    Inputs are seeded per rank on the host, so rank 0 can feed the CPU oracle without gathering them.
    P2P contract: bit-exact against the rank-order fp32 sum.  Multicast contract: within one bf16 ulp of the
    result plus the fp32 summation-order term, identical on every rank, identical on a second run; and the
    fused multicast kernel bit-equal to oracle AdamW fed with the switch's own sums."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from oracle import oracle as O
    from pyrecover_b200 import _lib
    from pyrecover_b200 import kernels as K
    from pyrecover_b200.gradsync import REGION_G, REGION_M, REGION_P, REGION_V
    gs = trainer.grad_sync
    comm, arena = gs.comm, gs.arena
    n = min(arena.numel, 16 << 20) // (128 * world) * (128 * world)      # elements checked (a prefix of the arenas)
    scale = 1.0 / world
    multicast = comm.multicast()[1]
    fused_mode = gs.mode in ("fused", "overlap")

    def host_grad(r):
        g = torch.Generator().manual_seed(1234 + r)
        return (torch.randn(n, generator=g) * 1e-3).bfloat16()
    g0 = torch.Generator().manual_seed(99)
    p0 = (torch.randn(n, generator=g0) * 0.02).bfloat16()
    m0 = (torch.randn(n, generator=g0) * 1e-3).bfloat16()
    v0 = (torch.randn(n, generator=g0) * 1e-3).square().bfloat16()
    mine = host_grad(rank).to(device)
    saved = [t[:n].clone() for t in (arena.p, arena.g, arena.m, arena.v)]

    def sync():
        torch.cuda.synchronize()
        dist.barrier()

    def same_on_all_ranks(t):
        a, b = K.checksum(t.contiguous().view(torch.int16))                 # two u64: carry them as four u32
        s = torch.tensor([a >> 32, a & 0xffffffff, b >> 32, b & 0xffffffff], dtype=torch.int64, device=device)
        allg = [torch.empty_like(s) for _ in range(world)]
        dist.all_gather(allg, s)
        return all(torch.equal(allg[0], x) for x in allg)

    out = {"elements": n, "algo": "multimem" if multicast else "p2p", "world": world}
    # ---- (a) allreduce, automatic choice ------------------------------------------------------------
    runs = []
    for blocks in (0, 17):                       # a second run on another grid: determinism
        arena.g[:n].copy_(mine)
        sync()
        comm.allreduce(REGION_G, 0, n * 2, torch.bfloat16, scale, algo=_lib.PR_AR_AUTO, max_blocks=blocks)
        sync()
        runs.append(arena.g[:n].clone())
    out["deterministic"] = bool(torch.equal(runs[0], runs[1]))
    out["identical_across_ranks"] = same_on_all_ranks(runs[0])
    sums_scale1 = None
    if multicast:                                # the switch's own bf16 sums (scale 1), the fused kernel's gradient
        arena.g[:n].copy_(mine)
        sync()
        comm.allreduce(REGION_G, 0, n * 2, torch.bfloat16, 1.0, algo=_lib.PR_AR_MULTIMEM)
        sync()
        sums_scale1 = arena.g[:n].clone()
    # ---- (b) exchange + update as GradSync launches it ----------------------------------------------
    lr, step = 3e-4, 3
    arena.g[:n].copy_(mine)
    arena.p[:n].copy_(p0.to(device))
    arena.m[:n].copy_(m0.to(device))
    arena.v[:n].copy_(v0.to(device))
    sync()
    if fused_mode:
        args = K.adamw_args(lr, 0.9, 0.999, 1e-8, 0.01, step, grad_scale=scale)
        blocks = gs.overlap_blocks if gs.mode == "overlap" else gs.max_blocks
        comm.allreduce_adamw(REGION_G, REGION_P, REGION_M, REGION_V, 0, n, args, max_blocks=blocks)
        comm.allgather_slices(REGION_M, 0, n * 2)
        comm.allgather_slices(REGION_V, 0, n * 2)
        kernel_b = "pr_allreduce_adamw (" + ("allreduce_adamw_mc_kernel" if multicast else f"allreduce_adamw_kernel<{world}>") + ")"
    else:
        comm.allreduce(REGION_G, 0, n * 2, torch.bfloat16, scale, max_blocks=gs.max_blocks)
        K.adamw_(arena.p[:n], arena.g[:n], arena.m[:n], arena.v[:n], lr=lr, step=step, numerics=0)
        kernel_b = "pr_allreduce + pr_adamw"
    sync()
    got_p, got_m, got_v = arena.p[:n].clone(), arena.m[:n].clone(), arena.v[:n].clone()
    out["identical_across_ranks"] = out["identical_across_ranks"] and same_on_all_ranks(got_p)
    out["kernel"] = ("allreduce_mc_kernel" if multicast else f"allreduce_twoshot_kernel<{world}>") + " ; " + kernel_b
    # ---- rank 0: the oracle -------------------------------------------------------------------------
    ok = True
    if rank == 0:
        verdict = parity_oracle(world, n, multicast, fused_mode, [O.to_bits(host_grad(r)) for r in range(world)],
                                O.to_bits(p0), O.to_bits(m0), O.to_bits(v0), lr, step, O.to_bits(runs[0]),
                                O.to_bits(sums_scale1) if sums_scale1 is not None else None,
                                O.to_bits(got_p), O.to_bits(got_m), O.to_bits(got_v))
        out.update(verdict)
        ok = verdict.pop("_ok") and out["deterministic"] and out["identical_across_ranks"]
        out.pop("_ok", None)
    flag = torch.tensor([int(ok)], device=device)
    dist.broadcast(flag, src=0)
    out["ok"] = bool(flag.item())
    for t, s in zip((arena.p, arena.g, arena.m, arena.v), saved):
        t[:n].copy_(s)
    sync()
    return out


def parity_oracle(world, n, multicast, fused_mode, grad_bits, p0, m0, v0, lr, step, got, sums_scale1, got_p, got_m, got_v):
    """Rank 0's half of parity_check: the kernels' outputs (numpy bf16 bit patterns) against the CPU oracle.
    Kept free of CUDA so that tests/ can run it on synthetic outputs."""
    import numpy as np

    from oracle import oracle as O
    scale = 1.0 / world
    res, ok = {}, True
    expect = O.allreduce_bf16(grad_bits, scale)                        # rank-order fp32 sum, scaled, rounded once
    diff = got.astype(np.int32) - expect.astype(np.int32)
    same_sign = (got >> 15) == (expect >> 15)
    ulp = np.where(same_sign, np.abs(diff), 1 << 16)
    ulp = np.where(((got & 0x7fff) == 0) & ((expect & 0x7fff) == 0), 0, ulp)      # +0 and -0 are the same value
    res["allreduce"] = {"max_ulp": int(ulp.max()), "fraction_bit_equal": float((ulp == 0).mean())}
    pb, mb, vb = p0.copy(), m0.copy(), v0.copy()
    if multicast:
        fsum = O.allreduce_bf16_to_f32(grad_bits, 1.0).astype(np.float64)
        S = np.sum([np.abs(O.from_bits(b).float().numpy().astype(np.float64)) for b in grad_bits], axis=0) / world
        exact = fsum * scale
        bound = 2.0 ** -7 * np.abs(exact) + world * 2.0 ** -23 * S + 1e-30
        err = np.abs(O.from_bits(got).float().numpy().astype(np.float64) - exact)
        res["allreduce"]["within_bound"] = bool((err <= bound).all())
        ok = ok and res["allreduce"]["within_bound"]
        # the update must be what oracle AdamW makes of the switch's own sums, bit for bit
        if fused_mode:
            O.adamw_bf16(pb, sums_scale1, mb, vb, lr=lr, step=step, mode=0, gscale=scale)
        else:
            O.adamw_bf16(pb, got, mb, vb, lr=lr, step=step, mode=0)
    else:
        res["allreduce"]["bit_exact"] = bool((ulp == 0).all())
        ok = ok and res["allreduce"]["bit_exact"]
        if fused_mode:                           # the fp32 rank-order sum feeds AdamW unrounded, 1/W in the gradient scale
            O.adamw_bf16_gf32(pb, O.allreduce_bf16_to_f32(grad_bits, 1.0), mb, vb, lr=lr, step=step, gscale=scale)
        else:
            O.adamw_bf16(pb, expect, mb, vb, lr=lr, step=step, mode=0)
    upd = {"p": bool(np.array_equal(got_p, pb)), "m": bool(np.array_equal(got_m, mb)), "v": bool(np.array_equal(got_v, vb))}
    res["update_bit_exact_vs_oracle"] = upd
    res["_ok"] = bool(ok and all(upd.values()))
    return res


# ------------------------------------------------------------------------- incumbent arm --
def incumbent_arm(a, device, world, rank, local_rank, data, ckpt_dir):
    """The thing to beat on the same box: the reference's GPU stack as written --
    DistributedDataParallel over NCCL (train.py:107-115), torch.optim.AdamW(fused=True) (train.py:120-122),
    and one reference-style checkpoint (torch.save of {model, optimizer, lr_scheduler} on rank 0 with the
    other ranks in a barrier, pyrecover/checkpoint.py:55-103) -- on the same workload, same step count."""
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    from pyrecover_b200.workload import Transformer, preset_args
    from utils import build_lr_scheduler
    torch.manual_seed(0)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(device):
        model = Transformer(preset_args(a.preset, a.seq))
    torch.set_default_dtype(torch.float32)
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank,
                                                    find_unused_parameters=False) if world > 1 else model
    opt = torch.optim.AdamW(net.parameters(), lr=1e-5, fused=True)
    sched = build_lr_scheduler(opt, 10)
    net.train()

    def step(x, y):
        num_items = y.ne(-100).sum()
        opt.zero_grad()
        logits = net(x)
        loss = F.cross_entropy(logits.flatten(0, 1).float(), y.flatten(0, 1), reduction="sum") / num_items
        del logits
        loss.backward()
        opt.step()
        sched.step()
        return loss
    total = a.warmup + a.steps
    batches = [tuple(t.to(device) for t in data.batch(7000 + s)) for s in range(total)]
    for s in range(a.warmup):
        step(*batches[s])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(a.warmup, total):
        step(*batches[s])
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    # reference-style checkpoint: everything stops, rank 0 serialises, the others wait in the barrier
    path = ckpt_dir / "incumbent.pt"
    t0 = time.perf_counter()
    if world > 1:
        dist.barrier()
    if rank == 0:
        torch.save({"epoch": 1, "step": total, "model": model.state_dict(), "optimizer": opt.state_dict(),
                    "lr_scheduler": sched.state_dict()}, str(path))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    store_s = time.perf_counter() - t0
    nbytes = path.stat().st_size if rank == 0 else 0
    if rank == 0:
        path.unlink(missing_ok=True)
    t = torch.tensor([ms, store_s], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, store_s = t.tolist()
    del net, model, opt, batches
    torch.cuda.empty_cache()
    return {"tokens_per_s": world * a.seq * a.steps / (ms * 1e-3), "ms_per_step": ms / a.steps, "ckpt_store_s": store_s,
            "ckpt_bytes": nbytes, "ckpt_stall_frac_of_step": store_s * 1e3 / (ms / a.steps),
            "impl": "torch DistributedDataParallel (NCCL) + torch.optim.AdamW(fused=True) + torch.save on rank 0, "
                    "training stopped while it writes" if world > 1 else
                    "torch.optim.AdamW(fused=True) + torch.save, training stopped while it writes"}


# ------------------------------------------------------------------------- own arm --------
def run_b200(a):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback); "
                         "use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    from pyrecover import checkpoint as ckpt
    from pyrecover_b200 import kernels as K
    from pyrecover_b200.trainer import Trainer
    from pyrecover_b200.workload import MODEL_PRESETS, SyntheticTokens, Transformer, preset_args

    grad_sync_mode = a.grad_sync or ("overlap" if world >= 4 else "fused")
    vocab = MODEL_PRESETS[a.preset][5]
    torch.manual_seed(0)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(device):                      # initialise on the GPU: an 8 B model takes minutes on the host
        model = Transformer(preset_args(a.preset, a.seq))
    torch.set_default_dtype(torch.float32)
    trainer = Trainer(model, device, lr=1e-5, fused=True, warmup_steps=10, world_size=world, local_rank=local_rank,
                      grad_sync_mode=grad_sync_mode)
    opt = trainer.optimizer
    n_params = sum(p.numel() for p in model.parameters())
    data = SyntheticTokens(vocab, a.seq, 1, seed=0, rank=rank, pin=True)
    tokens_per_step = world * 1 * a.seq
    ckpt_dir = Path(a.ckpt_dir or tempfile.mkdtemp(prefix="pr_bench_"))
    ckpt_at = None if a.no_ckpt else a.warmup + max(1, a.steps // 2)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    fmt = a.ckpt_format or ("vanilla" if world == 1 else "sharded")

    def save(tag, step_idx):
        # the reference-facing API, exactly as train.py calls it
        if fmt == "sharded":
            ckpt.save_ckpt_distributed(trainer.model, opt, trainer.lr_scheduler, None, step_idx, 1,
                                       ckpt_dir / tag / f"ckpt_{step_idx}", max_keep=1, verify=False,
                                       is_distributed=world > 1, rank=rank)
        else:
            ckpt.save_ckpt_vanilla(trainer.model, opt, trainer.lr_scheduler, None, step_idx, 1,
                                   ckpt_dir / tag / f"ckpt_{step_idx}.pt", max_keep=1, verify=False,
                                   is_distributed=world > 1, rank=rank)

    def maybe_ckpt(step_idx, tag):
        if ckpt_at is not None and step_idx == ckpt_at:
            save(tag, step_idx)

    clocks = ClockSampler(local_rank)
    if ckpt_at is not None:
        # staging image + pinned arena are allocated once per job (first checkpoint); do it before timing
        snap0 = ckpt._snapshotter(trainer.model, opt)
        if fmt == "sharded":
            snap0.shard(rank, world)
        elif rank == 0:
            snap0.full()
        snap0.skeleton()

    # ---------------- leg 1: device-resident inputs, CUDA-event timed -------------------------
    total = a.warmup + a.steps
    dev_batches = [tuple(t.to(device) for t in data.batch(s)) for s in range(total)]
    if world > 1:      # every rank must see the same directory
        names = [str(ckpt_dir)]
        dist.broadcast_object_list(names, src=0)
        ckpt_dir = Path(names[0])
    for d in (ckpt_dir / "warm", ckpt_dir / "dev", ckpt_dir / "e2e", ckpt_dir / "stall"):
        d.mkdir(parents=True, exist_ok=True)
    for s in range(a.warmup):
        trainer.step(*dev_batches[s])
    if ckpt_at is not None:
        # warm the checkpoint path like the steps above warm the kernels: the first save of a process pays one-time costs
        # (lazy loading of the pack kernel's module, first use of the writer thread's code paths) that belong to job
        # start-up, not to the steady state the timed region -- which still holds one full checkpoint -- measures
        save("warm", 0)
        ckpt.wait_for_checkpoints()
        barrier()
        if rank == 0:
            shutil.rmtree(ckpt_dir / "warm", ignore_errors=True)
    barrier()
    timed_owner = trainer.grad_sync if trainer.grad_sync is not None else opt
    timed_owner.kernel_events = []
    if trainer.grad_sync is not None:
        trainer.grad_sync.comm.stats(reset=True)
    launches_t0 = K.launch_count()
    if a.profile_range:
        torch.cuda.cudart().cudaProfilerStart()
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t_issue0 = time.perf_counter()
    for s in range(a.warmup, total):
        trainer.step(*dev_batches[s])
        maybe_ckpt(s, "dev")
    host_issue_ms = (time.perf_counter() - t_issue0) * 1e3 / a.steps      # host time to ENQUEUE a step
    e1.record()
    barrier()
    if a.profile_range:
        torch.cuda.cudart().cudaProfilerStop()
    clocks.stop()
    launches_timed = K.launch_count() - launches_t0
    k1_stats = trainer.grad_sync.comm.stats(reset=True) if trainer.grad_sync is not None else None
    ms = e0.elapsed_time(e1)
    k2_ms = [x.elapsed_time(y) for x, y in timed_owner.kernel_events]
    timed_owner.kernel_events = None
    ckpt.wait_for_checkpoints()
    barrier()
    if rank == 0:
        shutil.rmtree(ckpt_dir / "dev", ignore_errors=True)           # GBs per checkpoint: do not fill the disk
    del dev_batches

    # ---------------- leg 2: end to end through the public step (host batches) ----------------
    host_batches = [data.batch(1000 + s) for s in range(total)]
    for s in range(a.warmup):
        float(trainer.step_from_host(*host_batches[s]).detach())
    barrier()
    loss_host = torch.zeros(total, dtype=torch.float32).pin_memory()
    t0 = time.perf_counter()
    for s in range(a.warmup, total):
        loss = trainer.step_from_host(*host_batches[s])
        maybe_ckpt(s, "e2e")
        loss_host[s:s + 1].copy_(loss.detach().reshape(1), non_blocking=True)   # D2H of the step's result, every step
    barrier()                                                                # all losses are on the host here
    e2e_s = time.perf_counter() - t0
    loss_val = float(loss_host[total - 1])
    ckpt.wait_for_checkpoints()
    barrier()
    if rank == 0:
        shutil.rmtree(ckpt_dir / "e2e", ignore_errors=True)
    h2d = sum(t.numel() * t.element_size() for t in host_batches[0])

    # ---------------- max over ranks ------------------------------------------------------------
    times = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, e2e_ms = times.tolist()

    # ---------------- checkpoint stall, measured on its own (after the timed legs) -------------
    ck = None
    if ckpt_at is not None:
        # Blocks of 8 steps timed with CUDA events on the training stream, no host sync inside.  A = no
        # checkpoint, B = one checkpoint after the 4th step; A B A B ... A, the writer joined BETWEEN blocks
        # so that a block never waits for the previous block's file.  stall = median(B) - median(A): what a
        # snapshot costs the training stream, including the HBM / PCIe traffic and SMs it shares.
        dev_small = [tuple(t.to(device) for t in data.batch(5000 + s)) for s in range(4)]

        def block(with_save, idx):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            a0.record()
            for i in range(8):
                trainer.step(*dev_small[i % len(dev_small)])
                if with_save and i == 3:
                    save("stall", 900 + idx)
            a1.record()
            a1.synchronize()
            t = torch.tensor([a0.elapsed_time(a1)], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        block(False, 0)
        tA, tB, durable = [block(False, 0)], [], []
        for i in range(max(1, a.stall_blocks)):
            t_save0 = time.perf_counter()
            tB.append(block(True, i))
            ckpt.wait_for_checkpoints()
            durable.append(time.perf_counter() - t_save0)
            barrier()
            if rank == 0:                  # tens of GB per checkpoint at 8 B: never keep two on the disk
                shutil.rmtree(ckpt_dir / "stall", ignore_errors=True)
                (ckpt_dir / "stall").mkdir(parents=True, exist_ok=True)
            barrier()
            tA.append(block(False, 0))
        st = ckpt._snapshotter(trainer.model, opt).last_stats
        barrier()
        if rank == 0:
            shutil.rmtree(ckpt_dir / "stall", ignore_errors=True)
        step_ms = _median(tA) / 8
        stall_ms = max(0.0, _median(tB) - _median(tA))
        pair = [b - 0.5 * (x + y) for b, x, y in zip(tB, tA[:-1], tA[1:])]
        ck = {"stall_ms": stall_ms, "stall_frac_of_step": stall_ms / step_ms, "step_ms": step_ms,
              "block_ms": {"no_ckpt": tA, "with_ckpt": tB}, "stall_ms_per_pair": pair,
              "stall_ms_spread": [min(pair), max(pair)], "blocks": {"with_ckpt": len(tB), "no_ckpt": len(tA)},
              "durable_s": _median(durable), "enqueue_ms": 1e3 * st.get("enqueue_s", 0.0),
              "image_bytes": st.get("image_bytes"), "file_bytes": st.get("file_bytes"),
              "d2h_wait_s": st.get("d2h_wait_s"), "write_s": st.get("write_s"),
              "write_gbs": (st["file_bytes"] / st["write_s"] / 1e9) if st.get("write_s") else None,
              "mode": ckpt._snapshotter(trainer.model, opt).mode, "format": fmt,
              "pack": "bulk-copy kernel (cp.async.bulk through shared memory), bounded grid, side stream"}

    # ---------------- parity of the NVLink kernels against the oracle (N > 1) --------------------
    parity = None
    if world > 1 and not a.no_parity:
        parity = parity_check(trainer, world, rank, device)

    # ---------------- line ingredients that need the live trainer --------------------------------
    arena_elems = opt.arenas[0].numel
    arena_gb = opt.arenas[0].layout.nbytes / 1e9
    multicast = bool(trainer.grad_sync is not None and trainer.grad_sync.comm.multicast()[1])
    overlap_blocks = trainer.grad_sync.overlap_blocks if trainer.grad_sync is not None else None
    n_buckets = len(trainer.grad_sync._buckets) if trainer.grad_sync is not None else 0
    snapshot_mode = ckpt._snapshotter(trainer.model, opt).mode if ckpt_at is not None else None
    trainer.close()
    ckpt.release_checkpoint_engines()
    del trainer, opt, model
    torch.cuda.empty_cache()

    # ---------------- the GPU incumbent on the same workload --------------------------------------
    inc = None
    if not a.no_incumbent:
        try:
            inc = incumbent_arm(a, device, world, rank, local_rank, data, ckpt_dir)
        except Exception as e:  # noqa: BLE001 -- the incumbent failing must not void our own measurement
            inc = {"error": f"{type(e).__name__}: {e}"}

    # ---------------- cpu baseline (rank 0, N = 1) -----------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        from oracle import cpu_reference_port as port
        r = port.run(a.preset, a.seq, steps=2, warmup=1, fused=True, budget_s_per_step=a.cpu_budget_s)
        cpu = {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}

    if rank == 0:
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.exists():
            peaks = json.loads(pk.read_text())
        hbm_peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json (measured copy bandwidth)") if "hbm_gbs" in peaks \
            else (6650.0, "fallback of B200_PROFILING.md")
        value = tokens_per_step * a.steps / (ms * 1e-3)
        config = _config(a.preset, n_params, world, a.seq)
        config.update({
            "grad_sync": (grad_sync_mode + (" + NVSwitch multicast" if multicast else " (P2P)")) if world > 1 else "none (N=1)",
            "l2": f"inputs >> L2: four {arena_gb:.1f} GB arenas streamed once per step",
            "checkpoint": "none" if ckpt_at is None else f"1 {fmt} snapshot inside each timed region (bulk-copy pack on a side "
                          f"stream into an HBM image, D2H + file write in the writer thread; snapshot mode {snapshot_mode})"})
        if world == 1:
            k2_mean = sum(k2_ms) / max(1, len(k2_ms))
            k2_bytes = 14.0 * arena_elems
            achieved = k2_bytes / (k2_mean * 1e-3) / 1e9 if k2_ms else None
            traffic, traffic_src = None, None
            tf = ROOT / "profiles" / "k2_traffic.json"
            tj = json.loads(tf.read_text()) if tf.exists() else {}
            if tj.get("algorithmic_bytes_per_launch") == int(k2_bytes):      # the capture is of THIS workload's launch
                traffic = tj.get("dram_bytes_per_launch")
                traffic_src = ("profiles/k2_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` "
                               "capture of this kernel on this workload (" + str(tj.get("captured", "round 1")) + "); a "
                               "constant read from the profile, not measured by this run")
            roofline = {"kernel": "adamw_bf16_scattered_kernel<0> (K2, gradients read where autograd left them)",
                        "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                        "frac": (achieved / hbm_peak) if achieved else None, "traffic": traffic,
                        "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": k2_bytes,
                        "launch_ms_mean": k2_mean, "launches_timed": len(k2_ms), "share_of_step": k2_mean / (ms / a.steps)}
        else:
            # K1 (+K2 fused): NVLink is the binding roof.  bus bytes per step = gradient bytes * 2(W-1)/W
            # (nccl-tests convention); time = what CTA 0 of the exchange launches ran per step, from the
            # library's device-side clock (valid under overlap, where CUDA events cannot bracket the kernels),
            # and the same minus the begin-barrier wait for the slowest rank = the transfer itself.
            wait_ns, n_launch, run_ns = k1_stats
            run_ms, wait_ms = run_ns / 1e6 / a.steps, wait_ns / 1e6 / a.steps
            bus_bytes = 2.0 * arena_elems * 2 * (world - 1) / world
            achieved = bus_bytes / (run_ms * 1e-3) / 1e9 if run_ms > 0 else None
            transfer = bus_bytes / ((run_ms - wait_ms) * 1e-3) / 1e9 if run_ms > wait_ms else None
            exposed = (sum(k2_ms) / max(1, len(k2_ms))) if k2_ms else None
            hbm_bytes = 14.0 * arena_elems / world
            fusedk = grad_sync_mode in ("fused", "overlap")
            roofline = {"kernel": (("allreduce_adamw_mc_kernel (K1+K2 fused, NVSwitch multicast)" if multicast else
                                    f"allreduce_adamw_kernel<{world}> (K1+K2 fused, P2P)") if fusedk else
                                   ("allreduce_mc_kernel" if multicast else f"allreduce_twoshot_kernel<{world}>") + " then adamw_bf16_kernel"),
                        "bound": "nvlink", "achieved": achieved, "peak": 900.0, "unit": "GB/s",
                        "frac": (achieved / 900.0) if achieved else None, "traffic": None,
                        "peak_source": "900 GB/s per direction per GPU (NVLink 5 nominal); measured peer copy on this pool 770 GB/s",
                        "algorithmic_bytes_per_step": bus_bytes, "launches_per_step": n_launch / a.steps,
                        "kernel_ms_per_step": run_ms, "barrier_wait_ms_per_step": wait_ms,
                        "transfer_only": {"bus_gbs": transfer, "frac_of_nominal": (transfer / 900.0) if transfer else None,
                                          "note": "kernel time minus CTA 0's wait at the begin barrier for the slowest rank"},
                        "exposed_ms_after_backward": exposed,
                        "hbm_secondary": {"bytes_per_step": hbm_bytes, "gbs": hbm_bytes / (run_ms * 1e-3) / 1e9 if run_ms > 0 else None,
                                          "peak": hbm_peak, "note": "each rank updates 1/W of the elements (14 B each) in the same launches"},
                        "grid": {"overlap_blocks": overlap_blocks, "buckets": n_buckets},
                        "share_of_step": (exposed / (ms / a.steps)) if exposed is not None else None,
                        "timing": "device-side globaltimer of CTA 0 summed over the exchange launches of the timed region (pr_comm_stats)"}
        line = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config,
            "e2e": {"value": tokens_per_step * a.steps / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms / a.steps, "last_loss": loss_val},
            "gpu_launches": launches_timed,
            "host_issue_ms_per_step": host_issue_ms,     # close to ms_per_step = the loop is launch-bound on the host
            "clocks": clocks.summary(),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "ckpt": ck,
            "parity": parity,
            "incumbent": inc,
            "vs_incumbent": (value / inc["tokens_per_s"]) if inc and inc.get("tokens_per_s") else None,
        }
        print(json.dumps(line), flush=True)
    if rank == 0:
        shutil.rmtree(ckpt_dir, ignore_errors=True)
    ok = parity is None or parity.get("ok", False)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        print("bench.py: PARITY FAILED for the multi-GPU kernels (see the 'parity' object)", file=sys.stderr, flush=True)
        return 3
    return 0


if __name__ == "__main__":
    args = parse()
    sys.exit(run_reference(args) if args.impl == "reference" else run_b200(args))
