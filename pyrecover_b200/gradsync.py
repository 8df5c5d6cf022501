"""GradSync -- the data-parallel gradient exchange that stands in for DistributedDataParallel.

Replaces `DDP(model, device_ids=[local_rank], output_device=local_rank,
find_unused_parameters=False)` of the reference (train.py:107-115) and the bucketed
allreduce its Reducer fires inside `loss.backward()` (train.py:269):

  * construction broadcasts rank 0's parameters to every rank (DDP's _sync_module_states,
    torch/nn/parallel/distributed.py:879-889) -- over torch.distributed (NCCL/gloo =
    bootstrap only), as is the exchange of the CUDA-IPC handles of the symmetric arenas;
  * gradients land in the flat g arena, which IS the NVLink communication buffer (no separate
    bucket copy as with DDP's gradient_as_bucket_view=False): from a post-accumulate-grad hook per
    parameter when the exchange overlaps backward, else by one multi-tensor copy in
    optimizer.step() (no Python runs during backward); the autograd allocations are freed;
  * mode "replicated": buckets of the arena are all-reduced by K1 on a side stream as soon as
    all their gradients have landed (in bucket order, like the Reducer), overlapping the rest
    of backward; optimizer.step() then runs K2 over the whole arena on every rank;
  * mode "fused": optimizer.step() runs ONE kernel: reduce-scatter -> AdamW on the rank's
    slice -> all-gather of the updated parameters (moments are sharded across ranks);
  * mode "overlap": the same fused kernel, but per bucket and launched from the hooks on a side
    stream WHILE backward is still running: by the time a bucket's last gradient has landed, the
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

from typing import List, Optional

import torch

from . import _lib
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
        self.bucket_bytes = int(bucket_bytes)
        self.overlap = (overlap and mode == "replicated") or mode == "overlap"
        self.max_blocks = max_blocks
        # CTAs of a bucket allreduce that runs UNDER backward: NVLink saturates with a few dozen
        # CTAs, and every CTA taken is an SM the backward GEMMs cannot use
        self.overlap_blocks = overlap_blocks
        self.arena: Optional[FlatArena] = None
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self._hooks = []
        self._buckets: List[dict] = []
        self._next_bucket = 0
        self._seen = 0
        self._moments_replicated = True
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
        cur_hi = arena.numel
        members: List[int] = []
        for s in reversed(slots):
            members.append(s.index)
            if (cur_hi - s.offset) * esize >= self.bucket_bytes or s.index == 0:
                self._buckets.append({"lo": s.offset if s.index else 0, "hi": cur_hi,
                                      "members": members, "pending": len(members)})
                cur_hi = s.offset
                members = []
        self._bucket_of = {}
        for bi, b in enumerate(self._buckets):
            for idx in b["members"]:
                self._bucket_of[idx] = bi
        if self.overlap:
            for q, gv, slot in zip(arena.params, arena.g_views, slots):
                self._hooks.append(q.register_post_accumulate_grad_hook(self._make_hook(slot.index, gv)))
        # without overlap nothing has to happen DURING backward: optimizer.step() gathers all
        # gradients into the arena with one multi-tensor copy (147 python hook calls per step cost
        # ~4 ms of host time, which a launch-bound loop cannot hide)

    def _make_hook(self, index: int, gview: torch.Tensor):
        def hook(param: torch.nn.Parameter):
            gview.copy_(param.grad)
            param.grad = None
            self._seen += 1
            b = self._buckets[self._bucket_of[index]]
            b["pending"] -= 1
            if self.overlap and self.world > 1:
                self._launch_ready_buckets()
        return hook

    def _launch_ready_buckets(self) -> None:
        # strictly in bucket order so every rank issues the same collectives in the same order
        while self._next_bucket < len(self._buckets) and self._buckets[self._next_bucket]["pending"] == 0:
            b = self._buckets[self._next_bucket]
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.comm_stream.wait_event(ev)
            esize = self.arena.layout.esize
            if self.mode == "overlap":
                fence = getattr(getattr(self, "_optimizer", None), "_snapshot_fence", None)
                if fence is not None:      # a checkpoint snapshot may still be reading p / m / v
                    self.comm_stream.wait_event(fence)
                h = self._hyper()
                args = K.adamw_args(h["lr"], h["beta1"], h["beta2"], h["eps"], h["weight_decay"], h["step"],
                                    grad_scale=1.0 / self.world)
                self.comm.allreduce_adamw(REGION_G, REGION_P, REGION_M, REGION_V, b["lo"], b["hi"] - b["lo"], args,
                                          max_blocks=self.overlap_blocks, stream=self.comm_stream)
                self._moments_replicated = False
            else:
                self.comm.allreduce(REGION_G, b["lo"] * esize, (b["hi"] - b["lo"]) * esize, self.arena.dtype,
                                    1.0 / self.world, max_blocks=self.overlap_blocks, stream=self.comm_stream)
            self._next_bucket += 1

    # ---- called by B200AdamW.step() ---------------------------------------------------------
    def grads_ready(self) -> bool:
        if not self.overlap:
            grads = [q.grad for q in self.arena.params]
            have = sum(g is not None for g in grads)
            if have == 0:
                return False
            if have != len(grads):
                raise RuntimeError(f"GradSync: {have} of {len(grads)} gradients arrived "
                                   "(reference semantics: find_unused_parameters=False)")
            torch._foreach_copy_(self.arena.g_views, grads)
            for q in self.arena.params:
                q.grad = None
            return True
        if self._seen == 0:
            return False
        if self._seen != len(self.arena.params):
            raise RuntimeError(f"GradSync: {self._seen} of {len(self.arena.params)} gradients arrived "
                               "(reference semantics: find_unused_parameters=False)")
        return True

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
            ev = torch.cuda.Event()
            ev.record(self.comm_stream)
            cur.wait_event(ev)
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
        elif self.mode == "fused" and self.world > 1:
            if numerics != 0 or arena.dtype != torch.bfloat16:
                raise RuntimeError("fused GradSync implements bf16 arenas with fp32 op-math (fused=True)")
            args = K.adamw_args(lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0 / self.world)
            if events is not None:
                e0.record(cur)
            self.comm.allreduce_adamw(REGION_G, REGION_P, REGION_M, REGION_V, 0, arena.numel, args,
                                      max_blocks=self.max_blocks, stream=cur)
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
            self._moments_replicated = False
        else:
            if self.world > 1:
                if self.overlap:
                    self._launch_ready_buckets()
                    ev = torch.cuda.Event()
                    ev.record(self.comm_stream)
                    cur.wait_event(ev)
                else:
                    self.comm.allreduce(REGION_G, 0, arena.layout.nbytes, arena.dtype, 1.0 / self.world,
                                        max_blocks=self.max_blocks, stream=cur)
            if events is not None:
                e0.record(cur)
            coef = None
            if max_grad_norm:
                # after the exchange every rank holds the same mean gradient, bit for bit: each computes
                # the same norm from it, deterministically -- no second collective for the scalar
                norm.compute(arena.g, max_grad_norm)
                coef = norm.coef
            K.adamw_(arena.p, arena.g, arena.m, arena.v, lr=lr, beta1=beta1, beta2=beta2, eps=eps,
                     weight_decay=weight_decay, step=step, numerics=numerics, grad_scale_dev=coef)
            if events is not None:
                e1.record(cur)
                events.append((e0, e1))
        # re-arm for the next backward
        self._seen = 0
        self._next_bucket = 0
        for b in self._buckets:
            b["pending"] = len(b["members"])
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
        self._seen = 0
        self._next_bucket = 0
        for b in self._buckets:
            b["pending"] = len(b["members"])
        if self.comm.timed_out():
            raise RuntimeError("GradSync: a peer did not reach the gradient exchange (barrier timed out)")

    # ---- sharded-moment bookkeeping (fused mode) ------------------------------------------------
    def replicate_moments(self) -> None:
        """All-gather the per-rank moment slices so every rank holds full m, v (needed before a
        rank-0 vanilla checkpoint: reference save_ckpt_vanilla, pyrecover/checkpoint.py:58-74)."""
        if self.mode in ("fused", "overlap") and self.world > 1 and not self._moments_replicated:
            esize = self.arena.layout.esize
            # ownership follows the launches: one slice set over the whole arena ("fused") or one
            # per bucket ("overlap")
            ranges = [(0, self.arena.numel)] if self.mode == "fused" else [(b["lo"], b["hi"]) for b in self._buckets]
            for lo, hi in ranges:
                self.comm.allgather_slices(REGION_M, lo * esize, (hi - lo) * esize, max_blocks=self.max_blocks)
                self.comm.allgather_slices(REGION_V, lo * esize, (hi - lo) * esize, max_blocks=self.max_blocks)
            self._moments_replicated = True

    def _ownership_ranges(self):
        """Element ranges over which slices are cut: the whole arena ("replicated", "fused") or every
        bucket ("overlap": one fused launch, hence one set of W slices, per bucket)."""
        if self.mode == "overlap":
            return [(b["lo"], b["hi"]) for b in self._buckets]
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
