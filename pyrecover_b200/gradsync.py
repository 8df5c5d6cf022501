"""GradSync -- the data-parallel gradient exchange that stands in for DistributedDataParallel.

Replaces `DDP(model, device_ids=[local_rank], output_device=local_rank,
find_unused_parameters=False)` of the reference (train.py:107-115) and the bucketed
allreduce its Reducer fires inside `loss.backward()` (train.py:269):

  * construction broadcasts rank 0's parameters to every rank (DDP's _sync_module_states,
    torch/nn/parallel/distributed.py:879-889) -- over torch.distributed (NCCL/gloo =
    bootstrap only), as is the exchange of the CUDA-IPC handles of the symmetric arenas;
  * gradients land in the flat g arena, which IS the NVLink communication buffer (no separate
    bucket copy as with DDP's gradient_as_bucket_view=False), by the library's own gather kernel
    (pr_gather_grads: one launch per bucket, reading every gradient where autograd left it) -- per
    bucket on the communication stream when the exchange overlaps backward, else once in
    optimizer.step().  The per-parameter hooks of the overlapping modes only count down (about a
    microsecond each); the autograd allocations are released after the update and come back from
    the caching allocator at the same addresses next step (no cudaMalloc in steady state);
  * mode "replicated": buckets of the arena are all-reduced by K1 on a side stream as soon as
    all their gradients have landed (in bucket order, like the Reducer), overlapping the rest
    of backward; optimizer.step() then runs K2 over the whole arena on every rank;
  * mode "fused": optimizer.step() runs ONE kernel: reduce-scatter -> AdamW on the rank's
    slice -> all-gather of the updated parameters (moments are sharded across ranks);
  * mode "overlap": gather + the same fused kernel, per bucket, launched on a side stream from the
    hook of the bucket's last gradient WHILE backward is still running: by the time a bucket's last gradient has landed, the
    autograd nodes that read those parameters have already produced their input gradients
    (weight and input gradients come out of the same node), and a peer can only push new
    parameter values into this rank's arena after this rank's own kernel for that bucket has
    reached the begin barrier.  Only the last bucket's exchange is exposed.  Requires that no
    parameter is used twice in the graph (no weight tying, no recomputation) -- true for the
    reference model -- and one backward per step.

The mean's 1/W is applied after the fp32 sum (one rounding), not before in bf16 as the
Reducer does (reducer.hpp:500) -- strictly more accurate, bounded in tests/.
"""
from __future__ import annotations

import os
import time
from typing import List, Optional

import torch
from torch.cuda import nvtx

from . import kernels as K
from .flat import FlatArena

REGION_G, REGION_P, REGION_M, REGION_V = 0, 1, 2, 3
_REGION = {"g": REGION_G, "p": REGION_P, "m": REGION_M, "v": REGION_V}


def slice_of(lo: int, hi: int, esize: int, rank: int, world: int):
    """Rank `rank`'s slice [a, b) of the element range [lo, hi) cut into `world` runs of whole 16-byte
    vectors -- the partition K1's kernels use (csrc/allreduce.cu slice_bounds: per = ceil(nvec / W))."""
    nvec = (hi - lo) * esize // 16
    per = (nvec + world - 1) // world
    a = lo + min(per * rank, nvec) * 16 // esize
    b = lo + min(per * (rank + 1), nvec) * 16 // esize
    return a, b


class _Bucket:
    """A contiguous run of arena slots [seg_lo, seg_hi) = elements [lo, hi) exchanged by one launch."""
    __slots__ = ("lo", "hi", "seg_lo", "seg_hi", "members", "pending")

    def __init__(self, lo, hi, seg_lo, seg_hi):
        self.lo, self.hi, self.seg_lo, self.seg_hi = lo, hi, seg_lo, seg_hi
        self.members = seg_hi - seg_lo
        self.pending = self.members


class GradSync:
    def __init__(self, process_group=None, mode: str = "replicated", bucket_bytes: int = 256 << 20,
                 overlap: bool = True, max_blocks: int = 0, overlap_blocks: int = 32,
                 comm: Optional[K.Comm] = None):
        import torch.distributed as dist
        if mode not in ("replicated", "fused", "overlap"):
            raise ValueError("mode must be 'replicated', 'fused' or 'overlap'")
        self.mode = mode
        self.pg = process_group
        if comm is not None:                       # emulated / injected communicator (tests)
            self.comm, self.rank, self.world = comm, comm.rank, comm.world
        else:
            if not dist.is_initialized():
                raise RuntimeError("GradSync needs an initialised torch.distributed process group (bootstrap)")
            self.rank, self.world = dist.get_rank(process_group), dist.get_world_size(process_group)
            self.comm = K.Comm(self.rank, self.world)
        self.bucket_bytes = int(os.environ.get("PYRECOVER_B200_BUCKET_BYTES", bucket_bytes))
        self.overlap = (overlap and mode == "replicated") or mode == "overlap"
        # grid of the fused launch after backward (0 = the kernel's default, one CTA per SM)
        self.max_blocks = int(os.environ.get("PYRECOVER_B200_FUSED_BLOCKS", max_blocks))
        # CTAs of a bucket launch that runs UNDER backward: NVLink saturates with a few dozen
        # CTAs, and every CTA taken is an SM the backward GEMMs cannot use
        self.overlap_blocks = int(os.environ.get("PYRECOVER_B200_OVERLAP_BLOCKS", overlap_blocks))
        self.arena: Optional[FlatArena] = None
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self._hooks = []
        self._buckets: List[_Bucket] = []
        self._next_bucket = 0
        self._plan: Optional[K.AdamWPlan] = None
        self._ev_pool: List[torch.cuda.Event] = []
        self._moments_replicated = True
        self._learning = False
        self._n = 0                   # exchanges completed (one per optimizer step)
        self._seq_after = [0] * 16    # communicator launch sequence number after exchange n, ring
        self._hyper = None            # callable -> dict(lr, beta1, beta2, eps, weight_decay, step) of the NEXT update

    def bind_hyperparams(self, fn, optimizer=None) -> None:
        """B200AdamW hands over a callable giving the hyper-parameters of the update that the
        running backward feeds (mode "overlap" launches the update from the hooks)."""
        self._hyper = fn
        self._optimizer = optimizer

    # ---- arena allocation: the symmetric tensors come from the communicator ----------------
    def alloc(self, name: str, numel: int, dtype: torch.dtype) -> torch.Tensor:
        symmetric = name == "g" or self.mode in ("fused", "overlap")
        if not symmetric:
            return torch.zeros(numel, dtype=dtype, device=self.device)
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        self.comm.alloc_region(_REGION[name], nbytes)
        return self.comm.region_tensor(_REGION[name], dtype, self.rank if self.comm.emulated else 0)

    def attach(self, arena: FlatArena) -> None:
        import torch.distributed as dist
        self.arena = arena
        self.comm.connect(self.pg)
        if self.world > 1 and not self.comm.emulated:
            dist.broadcast(arena.p, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0,
                           group=self.pg)            # DDP._sync_module_states
        # buckets: contiguous arena ranges, formed from the END of the arena because backward
        # produces gradients in reverse parameter order
        esize = arena.layout.esize
        slots = arena.layout.slots
        cur_hi, seg_hi = arena.numel, len(slots)
        for s in reversed(slots):
            if (cur_hi - s.offset) * esize >= self.bucket_bytes or s.index == 0:
                self._buckets.append(_Bucket(s.offset if s.index else 0, cur_hi, s.index, seg_hi))
                cur_hi, seg_hi = s.offset, s.index
        if arena.dtype == torch.bfloat16:
            self._plan = K.AdamWPlan([s.offset for s in slots], [s.numel for s in slots], arena.numel)
        if self.overlap:
            # Calling into Python from the autograd engine costs ~25 us per hook whatever the hook does
            # (measured: 147 hooks = 3.7 ms of a 25 ms step that is launch-bound on the host).  So only the
            # FIRST backward runs one hook per parameter, to learn in which order gradients arrive; from
            # then on there is ONE hook per bucket, on the parameter whose gradient arrived last in it.
            self._arrival: List[int] = []
            self._learning = True
            for b in self._buckets:
                for i in range(b.seg_lo, b.seg_hi):
                    self._hooks.append(arena.params[i].register_post_accumulate_grad_hook(self._make_learning_hook(b, i)))
        # without overlap nothing happens DURING backward: optimizer.step() gathers all gradients
        # into the arena with one launch

    def _make_learning_hook(self, bucket: _Bucket, index: int):
        def hook(_param):
            self._arrival.append(index)
            bucket.pending -= 1
            if bucket.pending == 0 and self.world > 1:
                self._launch_ready_buckets()
        return hook

    def _install_trigger_hooks(self) -> None:
        """After the learning backward: one hook per bucket, on its last-arriving parameter."""
        order = {idx: k for k, idx in enumerate(self._arrival)}
        if len(order) != len(self.arena.params):
            return                                   # incomplete backward: keep learning
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._learning = False

        def trigger(_param):
            if self.world > 1:
                self._launch_ready_buckets()
        for b in self._buckets:
            last = max(range(b.seg_lo, b.seg_hi), key=lambda i: order[i])
            self._hooks.append(self.arena.params[last].register_post_accumulate_grad_hook(trigger))

    def _bucket_complete(self, b: _Bucket) -> bool:
        if self._learning:
            return b.pending == 0
        # the trigger fired: normally every gradient of the bucket is there; if the engine ever delivers them
        # in another order the bucket simply waits for a later trigger (or for optimizer.step())
        params = self.arena.params
        for i in range(b.seg_lo, b.seg_hi):
            if params[i].grad is None:
                return False
        return True

    def _event(self) -> torch.cuda.Event:
        return self._ev_pool.pop() if self._ev_pool else torch.cuda.Event()

    def _gather(self, seg_lo: int, seg_hi: int, stream) -> None:
        """Gradients of tensors [seg_lo, seg_hi) -> their slots of the g arena, on `stream`."""
        arena = self.arena
        grads = [q.grad for q in arena.params[seg_lo:seg_hi]]
        if self._plan is not None:
            self._plan.gather([None] * seg_lo + grads, arena.g, seg_lo, seg_hi, stream=stream)
        else:                                      # fp32 arenas (not the reference configuration): library copy
            with torch.cuda.stream(stream):
                torch._foreach_copy_(arena.g_views[seg_lo:seg_hi], grads)

    def _launch_ready_buckets(self) -> None:
        # strictly in bucket order so every rank issues the same collectives in the same order
        while self._next_bucket < len(self._buckets) and self._bucket_complete(self._buckets[self._next_bucket]):
            b = self._buckets[self._next_bucket]
            ev = self._event()
            ev.record(torch.cuda.current_stream())
            self.comm_stream.wait_event(ev)
            self._ev_pool.append(ev)
            esize = self.arena.layout.esize
            torch.cuda.nvtx.range_push("grad_sync")
            # the gradients stay attached to their parameters until optimizer.step() has made the
            # training stream wait for this stream, so the allocator cannot hand their memory out
            # while the gather below is still reading it
            self._gather(b.seg_lo, b.seg_hi, self.comm_stream)
            if self.mode == "overlap":
                fence = getattr(getattr(self, "_optimizer", None), "_snapshot_fence", None)
                if fence is not None:      # a checkpoint snapshot may still be reading p / m / v
                    self.comm_stream.wait_event(fence)
                h = self._hyper()
                args = K.adamw_args(h["lr"], h["beta1"], h["beta2"], h["eps"], h["weight_decay"], h["step"],
                                    grad_scale=1.0 / self.world)
                self.comm.allreduce_adamw(REGION_G, REGION_P, REGION_M, REGION_V, b.lo, b.hi - b.lo, args,
                                          max_blocks=self.overlap_blocks, stream=self.comm_stream)
                self._moments_replicated = False
            else:
                self.comm.allreduce(REGION_G, b.lo * esize, (b.hi - b.lo) * esize, self.arena.dtype,
                                    1.0 / self.world, max_blocks=self.overlap_blocks, stream=self.comm_stream)
            torch.cuda.nvtx.range_pop()
            self._next_bucket += 1

    # ---- called by B200AdamW.step() ---------------------------------------------------------
    def grads_ready(self) -> bool:
        n = len(self.arena.params)
        if not self.overlap or self.world == 1:
            have = sum(q.grad is not None for q in self.arena.params)
            if have == 0:
                return False
            if have != n:
                raise RuntimeError(f"GradSync: {have} of {n} gradients arrived "
                                   "(reference semantics: find_unused_parameters=False)")
            self._gather(0, n, torch.cuda.current_stream())
            self._release_grads()          # same stream as the gather: the memory may be reused at once
            return True
        seen = sum(q.grad is not None for q in self.arena.params)
        if seen == 0:
            return False
        if seen != n:
            raise RuntimeError(f"GradSync: {seen} of {n} gradients arrived "
                               "(reference semantics: find_unused_parameters=False)")
        return True

    def _release_grads(self) -> None:
        for q in self.arena.params:
            q.grad = None

    def _rearm(self) -> None:
        self._next_bucket = 0
        for b in self._buckets:
            b.pending = b.members
        if self.overlap and getattr(self, "_learning", False):
            self._install_trigger_hooks()
            self._arrival = []

    def reduce_and_step(self, *, lr, beta1, beta2, eps, weight_decay, step, numerics, max_grad_norm=None,
                        norm=None) -> None:
        arena = self.arena
        cur = torch.cuda.current_stream()
        if max_grad_norm and self.world > 1 and self.mode != "replicated":
            self._reduce_clip_step(lr=lr, beta1=beta1, beta2=beta2, eps=eps, weight_decay=weight_decay, step=step,
                                   numerics=numerics, max_grad_norm=max_grad_norm, norm=norm)
            return
        events = getattr(self, "kernel_events", None)          # bench.py: CUDA events around the update kernel
        if events is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if self.mode == "overlap" and self.world > 1:
            if numerics != 0 or arena.dtype != torch.bfloat16:
                raise RuntimeError("overlap GradSync implements bf16 arenas with fp32 op-math (fused=True)")
            if events is not None:
                e0.record(cur)
            self._launch_ready_buckets()          # normally a no-op: every bucket was launched by the hooks
            ev = self._event()
            ev.record(self.comm_stream)
            cur.wait_event(ev)
            self._ev_pool.append(ev)
            self._release_grads()                 # the training stream is now ordered after the gathers
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
        elif self.mode == "fused" and self.world > 1:
            if numerics != 0 or arena.dtype != torch.bfloat16:
                raise RuntimeError("fused GradSync implements bf16 arenas with fp32 op-math (fused=True)")
            args = K.adamw_args(lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0 / self.world)
            if events is not None:
                e0.record(cur)
            nvtx.range_push("grad_sync")
            nvtx.range_push("adamw")             # one launch does both
            self.comm.allreduce_adamw(REGION_G, REGION_P, REGION_M, REGION_V, 0, arena.numel, args,
                                      max_blocks=self.max_blocks, stream=cur)
            nvtx.range_pop()
            nvtx.range_pop()
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
            self._moments_replicated = False
        else:
            if self.world > 1:
                if self.overlap:
                    self._launch_ready_buckets()
                    ev = self._event()
                    ev.record(self.comm_stream)
                    cur.wait_event(ev)
                    self._ev_pool.append(ev)
                    self._release_grads()
                else:
                    nvtx.range_push("grad_sync")
                    self.comm.allreduce(REGION_G, 0, arena.layout.nbytes, arena.dtype, 1.0 / self.world,
                                        max_blocks=self.max_blocks, stream=cur)
                    nvtx.range_pop()
            if events is not None:
                e0.record(cur)
            nvtx.range_push("adamw")
            coef = None
            if max_grad_norm:
                # after the exchange every rank holds the same mean gradient, bit for bit: each computes
                # the same norm from it, deterministically -- no second collective for the scalar
                norm.compute(arena.g, max_grad_norm)
                coef = norm.coef
            K.adamw_(arena.p, arena.g, arena.m, arena.v, lr=lr, beta1=beta1, beta2=beta2, eps=eps,
                     weight_decay=weight_decay, step=step, numerics=numerics, grad_scale_dev=coef)
            nvtx.range_pop()
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
        self._rearm()                             # for the next backward
        self._note_exchange()
        if self.comm.timed_out():
            raise RuntimeError("GradSync: a peer did not reach the gradient exchange (barrier timed out)")

    def _reduce_clip_step(self, *, lr, beta1, beta2, eps, weight_decay, step, numerics, max_grad_norm, norm) -> None:
        """Clipping with sharded optimizer state ("fused" / "overlap" requested): the global norm is
        needed before any element is updated, so the single fused launch becomes allreduce (K1) ->
        norm of the mean gradient (identical on every rank) -> K2 on the rank's own slice with the
        coefficient read on the device -> all-gather of the updated parameters."""
        if self.mode == "overlap":
            raise NotImplementedError("gradient clipping needs every gradient before the first update: "
                                      "use --grad-sync fused or replicated with --clip-grad-norm")
        if numerics != 0 or self.arena.dtype != torch.bfloat16:
            raise RuntimeError("fused GradSync implements bf16 arenas with fp32 op-math (fused=True)")
        arena = self.arena
        cur = torch.cuda.current_stream()
        events = getattr(self, "kernel_events", None)
        if events is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
        self.comm.allreduce(REGION_G, 0, arena.layout.nbytes, arena.dtype, 1.0 / self.world,
                            max_blocks=self.max_blocks, stream=cur)
        norm.compute(arena.g, max_grad_norm)
        lo, hi = self.slice_range()
        K.adamw_(arena.p[lo:hi], arena.g[lo:hi], arena.m[lo:hi], arena.v[lo:hi], lr=lr, beta1=beta1, beta2=beta2,
                 eps=eps, weight_decay=weight_decay, step=step, numerics=0, grad_scale_dev=norm.coef)
        self.comm.allgather_slices(REGION_P, 0, arena.layout.nbytes, max_blocks=self.max_blocks, stream=cur)
        if events is not None:
            e1.record(cur)
            events.append((e0, e1))
        self._moments_replicated = False
        self._rearm()
        self._note_exchange()
        if self.comm.timed_out():
            raise RuntimeError("GradSync: a peer did not reach the gradient exchange (barrier timed out)")

    # ---- stop signal: rank 0's decision reaches every rank inside the exchange itself --------------
    # The reference broadcasts `should_stop` and reads it back with .item() after EVERY step
    # (train.py:342-346): one more NCCL launch and a full host<->device synchronisation per step.  Here
    # rank 0 drops the number of the exchange after which everybody stops into its signal word; the begin
    # barrier of every K1 launch forwards that word to each rank's host (pr_comm_signal_*), and the host
    # looks at it without synchronising anything.  The decision takes effect STOP_LAG steps later, on
    # every rank at the same step: by then each rank has provably seen it (it only waits, if at all, for
    # an exchange that was enqueued STOP_LAG steps ago).
    STOP_LAG = 2

    def _note_exchange(self) -> None:
        self._n += 1
        self._seq_after[self._n % len(self._seq_after)] = self.comm.seq()

    def request_stop(self) -> int:
        """Rank 0, before the next step is enqueued: everybody stops after exchange number
        (completed + 1 + STOP_LAG).  Returns that number.  Other ranks' calls are ignored by the kernels."""
        at = self._n + 1 + self.STOP_LAG
        if self.rank == 0:
            self.comm.signal_post(at, stream=torch.cuda.current_stream())
        return at

    def stop_due(self) -> bool:
        """Every rank, after a step: has rank 0 asked to stop at or before this exchange?  No device
        synchronisation: at most a wait for the exchange enqueued STOP_LAG steps ago to have started."""
        if self._n <= self.STOP_LAG:
            return False
        need = self._seq_after[(self._n - self.STOP_LAG) % len(self._seq_after)]
        while True:
            seq, val = self.comm.signal_read()
            if ((seq - need) & 0xffffffff) < 0x80000000 or self.comm.timed_out():
                break
            time.sleep(0)
        return val != 0 and self._n >= val

    # ---- sharded-moment bookkeeping (fused mode) ------------------------------------------------
    def replicate_moments(self) -> None:
        """All-gather the per-rank moment slices so every rank holds full m, v (needed before a
        rank-0 vanilla checkpoint: reference save_ckpt_vanilla, pyrecover/checkpoint.py:58-74)."""
        if self.mode in ("fused", "overlap") and self.world > 1 and not self._moments_replicated:
            esize = self.arena.layout.esize
            # ownership follows the launches: one slice set over the whole arena ("fused") or one
            # per bucket ("overlap")
            ranges = [(0, self.arena.numel)] if self.mode == "fused" else [(b.lo, b.hi) for b in self._buckets]
            for lo, hi in ranges:
                self.comm.allgather_slices(REGION_M, lo * esize, (hi - lo) * esize, max_blocks=self.max_blocks)
                self.comm.allgather_slices(REGION_V, lo * esize, (hi - lo) * esize, max_blocks=self.max_blocks)
            self._moments_replicated = True

    def _ownership_ranges(self):
        """Element ranges over which slices are cut: the whole arena ("replicated", "fused") or every
        bucket ("overlap": one fused launch, hence one set of W slices, per bucket)."""
        if self.mode == "overlap":
            return [(b.lo, b.hi) for b in self._buckets]
        return [(0, self.arena.numel)]

    def owned_ranges(self, rank: Optional[int] = None, world: Optional[int] = None):
        """Element ranges [lo, hi) of the flat arenas whose optimizer state rank `rank` owns -- its
        reduce-scatter slices, which are also its checkpoint shard (no re-replication before a
        sharded save)."""
        r = self.rank if rank is None else rank
        w = self.world if world is None else world
        if w != self.world:          # a layout for another world size: plain arena slices
            ranges = [(0, self.arena.numel)]
        else:
            ranges = self._ownership_ranges()
        esize = self.arena.layout.esize
        out = []
        for lo, hi in ranges:
            a, b = slice_of(lo, hi, esize, r, w)
            if b > a:
                out.append((a, b))
        return out

    def allgather_arena(self, name: str, t: torch.Tensor) -> None:
        """Every rank holds its `owned_ranges()` of the flat arena tensor `t` (p, m or v); make every
        rank hold all of it.  Symmetric arenas are gathered in place by K1's all-gather kernel;
        a plain tensor (replicated mode keeps p / m / v outside the communicator) bounces through
        the gradient arena, whose contents are dead outside backward."""
        if self.world == 1:
            return
        arena = self.arena
        esize = arena.layout.esize
        symmetric = name == "g" or self.mode in ("fused", "overlap")
        region = _REGION[name] if symmetric else REGION_G
        if not symmetric:
            for lo, hi in self.owned_ranges():
                arena.g[lo:hi].copy_(t[lo:hi])
        for lo, hi in self._ownership_ranges():
            self.comm.allgather_slices(region, lo * esize, (hi - lo) * esize, max_blocks=self.max_blocks)
        if not symmetric:
            t.copy_(arena.g)

    def moments_loaded(self) -> None:
        self._moments_replicated = True

    def slice_range(self, rank: Optional[int] = None):
        """Element range [lo, hi) of the arena a rank owns (reduce-scatter slice; checkpoint shard)."""
        r = self.rank if rank is None else rank
        return slice_of(0, self.arena.numel, self.arena.layout.esize, r, self.world)

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []
        torch.cuda.synchronize()
        self.comm.close()
