#!/usr/bin/env python
"""bench.py -- tokens/s of the data-parallel training step on BASELINE.json's configs[1]
(~1 B-parameter model.py-style transformer, bf16, seq 2048, batch 1 / GPU, fused AdamW).

    python bench.py --gpus N --steps K --warmup W              # this repo's arm
    python bench.py --impl reference --gpus N --steps K ...    # the reference's CPU path (port)

For N > 1 the driver launches it under torch.distributed.run, one rank per GPU.

A step = one pass of the hot path over one synthetic batch: forward + backward of the workload
model (library kernels: the dense model is out of scope), the gradient exchange (K1, N > 1), the
flat AdamW update (K2) and -- once, in the middle of the timed region -- a checkpoint snapshot
(K3 pack+checksum, async D2H and file write).
  value : whole-job tokens/s with the batches already resident in HBM (CUDA-event timed)
  e2e   : the same through Trainer.step_from_host (pinned-host batch -> H2D inside, loss .item()
          D2H every step), wall-clock between device synchronisations
  roofline : K2 (the dominant kernel of the hot path): 14 B per bf16 parameter / its mean duration
             from CUDA events recorded around every launch inside the timed region
  cpu_baseline : the reference's CPU path (oracle/cpu_reference_port.py) on this box's host cores,
             bounded token sample, rank 0 at N = 1 only
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
    ap.add_argument("--grad-sync", type=str, default="fused", choices=["replicated", "fused", "overlap"])
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around timed leg 1 (for ncu --profile-from-start off)")
    ap.add_argument("--no-ckpt", action="store_true", help="skip the in-region checkpoint snapshot")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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


# ------------------------------------------------------------------------- reference arm --
def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpu_reference_port as port
    tmp = a.ckpt_dir or tempfile.mkdtemp(prefix="pr_ref_")
    res = port.run(a.preset, a.seq, steps=max(1, a.steps), warmup=max(1, min(a.warmup, 2)), fused=True,
                   budget_s_per_step=a.cpu_budget_s, ckpt_path=os.path.join(tmp, "ckpt_ref.pt"))
    line = {
        "impl": "reference", "metric": METRIC, "value": res["tokens_per_s"], "unit": "tokens/s", "n_gpus": a.gpus,
        "steps": res["steps"], "warmup": max(1, min(a.warmup, 2)), "ms_per_step": res["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOADS[a.preset], "device": "host CPU", "note": "reference's CPU path restated call for call "
                   "(oracle/cpu_reference_port.py): torch fwd/bwd + torch.optim.AdamW(fused=True) on a bounded token sample"},
        "cpu_baseline": {"value": res["tokens_per_s"], "unit": "tokens/s", "cores": res["cores"], "kind": "port",
                         "sample": res["sample"]},
        "e2e": {"value": res["tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ckpt": {"store_s": res.get("ckpt_store_s"), "bytes": res.get("ckpt_bytes"),
                 "gbs": (res["ckpt_bytes"] / res["ckpt_store_s"] / 1e9) if res.get("ckpt_store_s") else None,
                 "note": "torch.save of {model, optimizer, lr_scheduler}, synchronous, outside the timed steps"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


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

    vocab = MODEL_PRESETS[a.preset][5]
    torch.manual_seed(0)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(device):                      # initialise on the GPU: an 8 B model takes minutes on the host
        model = Transformer(preset_args(a.preset, a.seq))
    torch.set_default_dtype(torch.float32)
    trainer = Trainer(model, device, lr=1e-5, fused=True, warmup_steps=10, world_size=world, local_rank=local_rank,
                      grad_sync_mode=a.grad_sync)
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
        # the reference-facing API, exactly as train.py calls it (barriers included)
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
        snap0.shard(rank, world) if (a.ckpt_format or ("vanilla" if world == 1 else "sharded")) == "sharded" else \
            (snap0.full() if rank == 0 else None)

    # ---------------- leg 1: device-resident inputs, CUDA-event timed -------------------------
    total = a.warmup + a.steps
    dev_batches = [tuple(t.to(device) for t in data.batch(s)) for s in range(total)]
    if world > 1:      # every rank must see the same directory
        names = [str(ckpt_dir)]
        dist.broadcast_object_list(names, src=0)
        ckpt_dir = Path(names[0])
    for d in (ckpt_dir / "dev", ckpt_dir / "e2e", ckpt_dir / "stall"):
        d.mkdir(parents=True, exist_ok=True)
    for s in range(a.warmup):
        trainer.step(*dev_batches[s])
    barrier()
    timed_owner = trainer.grad_sync if trainer.grad_sync is not None else opt
    timed_owner.kernel_events = []
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
        # blocks of 8 steps timed with CUDA events on the training stream, no host sync inside:
        # A = no checkpoint, B = one checkpoint after the 4th step, A again.  stall = t(B) - mean t(A):
        # what the snapshot costs the training stream, including the HBM / PCIe traffic it shares.
        dev_small = [tuple(t.to(device) for t in data.batch(5000 + s)) for s in range(4)]

        def block(with_save):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            a0.record()
            for i in range(8):
                trainer.step(*dev_small[i % len(dev_small)])
                if with_save and i == 3:
                    save("stall", 900)
            a1.record()
            a1.synchronize()
            return a0.elapsed_time(a1)
        block(False)
        tA1 = block(False)
        t_save0 = time.perf_counter()
        tB = block(True)
        ckpt.wait_for_checkpoints()
        durable_s = time.perf_counter() - t_save0
        tA2 = block(False)
        st = ckpt._snapshotter(trainer.model, opt).last_stats
        barrier()
        if rank == 0:
            shutil.rmtree(ckpt_dir / "stall", ignore_errors=True)
        step_ms = 0.5 * (tA1 + tA2) / 8
        stall_ms = max(0.0, tB - 0.5 * (tA1 + tA2))
        ck = {"stall_ms": stall_ms, "stall_frac_of_step": stall_ms / step_ms, "step_ms": step_ms,
              "block_ms": {"no_ckpt": [tA1, tA2], "with_ckpt": tB}, "durable_s": durable_s,
              "image_bytes": st.get("image_bytes"), "file_bytes": st.get("file_bytes"),
              "d2h_wait_s": st.get("d2h_wait_s"), "write_s": st.get("write_s"),
              "write_gbs": (st["file_bytes"] / st["write_s"] / 1e9) if st.get("write_s") else None,
              "mode": ckpt._snapshotter(trainer.model, opt).mode, "format": fmt}

    # ---------------- cpu baseline (rank 0, N = 1) -----------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        from oracle import cpu_reference_port as port
        r = port.run(a.preset, a.seq, steps=2, warmup=1, fused=True, budget_s_per_step=a.cpu_budget_s)
        cpu = {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["cores"], "kind": "port", "sample": r["sample"]}

    if rank == 0:
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.exists():
            peaks = json.loads(pk.read_text())
        hbm_peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        arena_elems = opt.arenas[0].numel
        k2_mean = sum(k2_ms) / max(1, len(k2_ms))
        fused_multi = world > 1 and a.grad_sync in ("fused", "overlap")
        multicast = bool(trainer.grad_sync is not None and trainer.grad_sync.comm.multicast()[1])
        # fused K1+K2: every rank updates 1/W of the elements (14 B each) -- reported as HBM GB/s --
        # and moves N * 2(W-1)/W bus bytes over NVLink in the same launch
        k2_bytes = 14.0 * arena_elems / (world if fused_multi else 1)
        achieved = k2_bytes / (k2_mean * 1e-3) / 1e9 if k2_ms else None
        nvlink = None
        overlapped = world > 1 and a.grad_sync == "overlap"
        if overlapped:
            # the per-bucket fused kernels run under backward on a side stream; the events bracket only
            # what is left AFTER backward (the exposed tail), which is not a kernel duration
            achieved = None
        if fused_multi and k2_ms and not overlapped:
            bus = 2.0 * arena_elems * 2 * (world - 1) / world / (k2_mean * 1e-3) / 1e9
            nvlink = {"bus_gbs": bus, "peak_measured_peer_copy": 770.0, "peak_nominal": 900.0,
                      "frac_of_measured": bus / 770.0, "frac_of_nominal": bus / 900.0}
        traffic = None
        tf = ROOT / "profiles" / "k2_traffic.json"
        if tf.exists() and world == 1:            # the ncu capture is of the single-GPU K2
            traffic = json.loads(tf.read_text()).get("dram_bytes_per_launch")
        line = {
            "metric": METRIC, "value": tokens_per_step * a.steps / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOADS[a.preset], "params": n_params, "global_batch": world, "seq_len": a.seq,
                       "parallelism": f"dp{world}", "grad_sync": (a.grad_sync + (" + NVSwitch multicast" if multicast else " (P2P)")) if world > 1 else "none (N=1)",
                       "l2": f"inputs >> L2: four {opt.arenas[0].layout.nbytes / 1e9:.1f} GB arenas streamed once per step",
                       "checkpoint": "none" if ckpt_at is None else f"1 {fmt} snapshot inside each timed region (staged pack on a side stream, D2H + file write in the writer thread)"},
            "e2e": {"value": tokens_per_step * a.steps / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms / a.steps, "last_loss": loss_val},
            "gpu_launches": launches_timed,
            "host_issue_ms_per_step": host_issue_ms,     # close to ms_per_step = the loop is launch-bound on the host
            "clocks": clocks.summary(),
            "roofline": {"kernel": (("allreduce_adamw_mc_kernel (K1+K2 fused, NVSwitch multicast)" if multicast else
                                     "allreduce_adamw_kernel<W> (K1+K2 fused, P2P)") if fused_multi else
                                    "adamw_bf16_scattered_kernel<0> (K2, gradients read where autograd left them)"),
                         "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": (achieved / hbm_peak) if achieved else None, "traffic": traffic,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": k2_bytes, "nvlink": nvlink,
                         "launch_ms_mean": None if overlapped else k2_mean, "launches_timed": len(k2_ms),
                         "exposed_ms_after_backward": k2_mean if overlapped else None,
                         "share_of_step": k2_mean / (ms / a.steps)},
            "cpu_baseline": cpu,
            "ckpt": ck,
        }
        print(json.dumps(line), flush=True)
    trainer.close()
    ckpt.release_checkpoint_engines()
    if rank == 0:
        shutil.rmtree(ckpt_dir, ignore_errors=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    args = parse()
    sys.exit(run_reference(args) if args.impl == "reference" else run_b200(args))
