"""B200AdamW -- drop-in for `torch.optim.AdamW(model.parameters(), lr=..., fused=...)`.

Replaces the optimizer the reference builds at train.py:120-122 and steps at train.py:274.
It honours the torch.optim.Optimizer protocol that the rest of the reference consumes:
`param_groups` (mutated by LambdaLR, utils.py:59-81 / train.py:275), `zero_grad()`
(train.py:260), `state_dict()/load_state_dict()` (pyrecover/checkpoint.py:69,188), with a
state dict that is key-for-key what torch.optim.AdamW produces, so checkpoints interchange
with the reference in both directions.

What changes underneath: parameters, gradients and both moments live in flat HBM arenas
(flat.FlatArena) and one K2 launch (csrc/adamw.cu) updates a whole param group.  With a
GradSync attached (world > 1) the gradient arena is the NVLink communication buffer of K1
and `step()` can run the fused reduce-scatter -> AdamW -> all-gather kernel instead.

`fused` keeps the reference flag's meaning as a NUMERICS selector: fused=True -> fp32
op-math rounded once (class of torch._fused_adamw_), fused=False/None -> every torch op's
result rounded to the storage dtype (class of _single_tensor_adam, the reference default).
Both are a single launch here.
"""
from __future__ import annotations

import time
from typing import Any, Dict, Iterable, List, Optional

import torch

from . import kernels as K
from .flat import FlatArena


class B200AdamW(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, amsgrad: bool = False, *, maximize: bool = False,
                 foreach: Optional[bool] = None, capturable: bool = False, differentiable: bool = False,
                 fused: Optional[bool] = None, numerics: Optional[int] = None, grad_sync=None,
                 max_grad_norm: Optional[float] = None):
        if amsgrad or maximize or capturable or differentiable:
            raise NotImplementedError("B200AdamW implements the reference configuration only "
                                      "(amsgrad/maximize/capturable/differentiable = False)")
        if isinstance(lr, torch.Tensor):
            lr = float(lr)
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("invalid AdamW hyper-parameter")
        self._numerics_override = numerics
        self._grad_sync = grad_sync
        # global-norm clipping folded into the update (reference utils.py:84-89, call commented out at
        # train.py:271-272): the norm is two small launches over the gradients, the coefficient stays
        # on the device and K2 multiplies it in -- no host sync, no extra pass, no re-rounded gradients
        if max_grad_norm is not None and not max_grad_norm > 0:
            raise ValueError("max_grad_norm must be positive (None disables clipping)")
        self.max_grad_norm = max_grad_norm
        self._grad_norm: Optional[K.GradNorm] = None
        self._arenas: List[FlatArena] = []
        self._steps: List[int] = []
        self._step_tensors: List[torch.Tensor] = []
        self._plans: List[Optional[K.AdamWPlan]] = []
        self.in_place_grads = True       # False: always gather gradients into the flat g arena first
        self.step_period_s: Optional[float] = None     # smoothed host time between step() calls (the snapshot sizes its grid by it)
        self._last_step_t: Optional[float] = None
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad,
                        maximize=maximize, foreach=foreach, capturable=capturable,
                        differentiable=differentiable, fused=fused, decoupled_weight_decay=True)
        super().__init__(params, defaults)

    # ------------------------------------------------------------------ construction ----
    def add_param_group(self, param_group: Dict[str, Any]) -> None:
        super().add_param_group(param_group)
        group = self.param_groups[-1]
        params = group["params"]
        if self._grad_sync is not None:
            if self._arenas:
                raise NotImplementedError("GradSync supports a single param group (as the reference uses)")
            if self.max_grad_norm and self._grad_sync.mode == "overlap":
                raise NotImplementedError("gradient clipping needs every gradient before the first update, but "
                                          "'overlap' updates bucket by bucket under backward: use --grad-sync "
                                          "fused or replicated with --clip-grad-norm")
            arena = FlatArena(params, world=self._grad_sync.world, alloc=self._grad_sync.alloc)
            self._grad_sync.attach(arena)

            def next_update_hyper(group=group):
                b1, b2 = group["betas"]
                return dict(lr=float(group["lr"]), beta1=b1, beta2=b2, eps=group["eps"],
                            weight_decay=group["weight_decay"], step=self._steps[0] + 1)
            self._grad_sync.bind_hyperparams(next_update_hyper, self)
        else:
            arena = FlatArena(params)
        self._arenas.append(arena)
        self._plans.append(None)
        self._steps.append(0)
        # one shared scalar per group: every parameter of the reference steps in lock-step
        step_t = torch.zeros((), dtype=torch.float32)
        self._step_tensors.append(step_t)
        for q, mv, vv in zip(params, arena.m_views, arena.v_views):
            # eager state (torch creates it lazily at the first step -- adam.py:168-184 -- which is
            # why a DCP resume into a fresh reference optimizer silently drops the moments)
            self.state[q] = {"step": step_t, "exp_avg": mv, "exp_avg_sq": vv}

    @property
    def arenas(self) -> List[FlatArena]:
        return self._arenas

    def grad_norm_state(self) -> Optional[K.GradNorm]:
        """Device-side [norm, coefficient, non-finite flag] of the last clipped step (None when
        clipping is off).  Reading it synchronises; do it at logging frequency."""
        return self._grad_norm

    def _norm_buffers(self) -> K.GradNorm:
        if self._grad_norm is None:
            self._grad_norm = K.GradNorm(self._arenas[0].device)
            self._nonfinite_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        return self._grad_norm

    def _note_norm(self) -> None:
        """After a clipped step has been enqueued: bring the sticky non-finite counter to the host with
        an asynchronous 4-byte copy, so check_finite() never has to synchronise."""
        if self._grad_norm is not None:
            self._nonfinite_host.copy_(self._grad_norm.out[3:4], non_blocking=True)

    def check_finite(self) -> None:
        """The reference raises as soon as the total gradient norm is not finite (utils.py:87,
        get_total_norm(error_if_nonfinite=True)).  Here the step is enqueued without waiting for the
        norm, so the device withholds the update by itself (K2 returns when the coefficient is the -1
        that pr_grad_norm writes for a non-finite norm: p, m, v and the next checkpoint stay clean) and
        this check raises at the caller's convenience -- train.py calls it at logging frequency and
        before every save, i.e. up to a few steps late: the one deviation from the reference."""
        if self._grad_norm is not None and float(self._nonfinite_host[0]) > 0:
            raise RuntimeError("The total norm of the gradients is non-finite, so it cannot be clipped "
                               f"({int(self._nonfinite_host[0])} step(s) so far had their update withheld)")

    def _numerics(self, group) -> int:
        if self._numerics_override is not None:
            return int(self._numerics_override)
        if self._arenas[0].dtype != torch.bfloat16:
            return 0
        return 0 if group.get("fused") else 1

    # ------------------------------------------------------------------ step -------------
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        now = time.perf_counter()
        if self._last_step_t is not None:
            dt = now - self._last_step_t
            if self.step_period_s is None:
                self.step_period_s = dt
            elif dt < 2.5 * self.step_period_s:         # a pause (logging sync, a blocking save, evaluation) is not a step
                self.step_period_s = 0.8 * self.step_period_s + 0.2 * dt
            else:
                self._period_outliers = getattr(self, "_period_outliers", 0) + 1
                if self._period_outliers >= 8:          # the cadence really changed
                    self.step_period_s, self._period_outliers = dt, 0
        self._last_step_t = now
        fence = getattr(self, "_snapshot_fence", None)
        if fence is not None:      # a checkpoint snapshot may still be reading the live arenas
            torch.cuda.current_stream().wait_event(fence)
            self._snapshot_fence = None
        for gi, group in enumerate(self.param_groups):
            arena = self._arenas[gi]
            beta1, beta2 = group["betas"]
            lr = float(group["lr"])
            if self._grad_sync is not None:
                if not self._grad_sync.grads_ready():
                    continue
                self._steps[gi] += 1
                self._step_tensors[gi].fill_(self._steps[gi])
                self._grad_sync.reduce_and_step(
                    lr=lr, beta1=beta1, beta2=beta2, eps=group["eps"], weight_decay=group["weight_decay"],
                    step=self._steps[gi], numerics=self._numerics(group),
                    max_grad_norm=self.max_grad_norm, norm=self._norm_buffers() if self.max_grad_norm else None)
                if self.max_grad_norm:
                    self._note_norm()
                continue
            have = [q.grad is not None for q in arena.params]
            if not any(have):
                continue
            if not all(have):
                raise RuntimeError("B200AdamW: some parameters have no gradient; the flat arena updates "
                                   "all parameters of a group together (reference: find_unused_parameters=False)")
            self._steps[gi] += 1
            self._step_tensors[gi].fill_(self._steps[gi])
            events = getattr(self, "kernel_events", None)      # bench.py: CUDA events around K2
            scattered = arena.dtype == torch.bfloat16 and self.in_place_grads
            grads = arena.grads_in_place() if scattered else None
            if not scattered:
                arena.gather_grads()
            if events is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            torch.cuda.nvtx.range_push("adamw")
            coef = None
            if self.max_grad_norm:
                if len(self._arenas) != 1:
                    raise NotImplementedError("gradient clipping supports a single param group (as the reference uses)")
                coef = self._norm_buffers().coef
            if scattered:
                # world size 1: read every gradient where autograd left it -- no copy into the g arena
                if self._plans[gi] is None:
                    self._plans[gi] = K.AdamWPlan([s.offset for s in arena.layout.slots],
                                                  [s.numel for s in arena.layout.slots], arena.numel)
                if coef is not None:
                    self._plans[gi].grad_norm(grads, self._grad_norm, self.max_grad_norm)
                self._plans[gi].step(arena.p, grads, arena.m, arena.v, lr=lr, beta1=beta1, beta2=beta2,
                                     eps=group["eps"], weight_decay=group["weight_decay"], step=self._steps[gi],
                                     numerics=self._numerics(group), grad_scale_dev=coef)
            else:
                if coef is not None:
                    self._grad_norm.compute(arena.g, self.max_grad_norm)
                K.adamw_(arena.p, arena.g, arena.m, arena.v, lr=lr, beta1=beta1, beta2=beta2, eps=group["eps"],
                         weight_decay=group["weight_decay"], step=self._steps[gi], numerics=self._numerics(group),
                         grad_scale_dev=coef)
            torch.cuda.nvtx.range_pop()
            if coef is not None:
                self._note_norm()
            if events is not None:
                e1.record()
                events.append((e0, e1))
        return loss

    # ------------------------------------------------------------------ state dict -------
    def state_dict(self) -> Dict[str, Any]:
        if self._grad_sync is not None:
            self._grad_sync.replicate_moments()
        sd = super().state_dict()
        # torch.optim.AdamW hands out one scalar per parameter; give each entry its own tensor
        sd["state"] = {k: {**st, "step": st["step"].clone()} for k, st in sd["state"].items()}
        return sd

    @torch.no_grad()
    def load_state_dict(self, state_dict: Dict[str, Any], tensors_restored: bool = False) -> None:
        """torch.optim.Optimizer.load_state_dict semantics.  tensors_restored=True (used by the
        fast checkpoint loader) means the unpack kernel already put exp_avg / exp_avg_sq into the
        arenas: only hyper-parameters and the step count are taken from the dict."""
        groups = state_dict["param_groups"]
        if len(groups) != len(self.param_groups):
            raise ValueError("loaded state dict has a different number of parameter groups")
        state = state_dict.get("state", {})
        for gi, (group, saved) in enumerate(zip(self.param_groups, groups)):
            if len(saved["params"]) != len(group["params"]):
                raise ValueError("loaded state dict contains a parameter group that doesn't match the size of "
                                 "optimizer's group")
            arena = self._arenas[gi]
            step_val = None
            for slot, idx in enumerate(saved["params"]):
                st = state.get(idx, state.get(str(idx)))
                if st is None:
                    continue     # torch semantics: parameter never stepped
                if not tensors_restored:
                    arena.m_views[slot].copy_(st["exp_avg"].to(arena.device, arena.dtype).view_as(arena.m_views[slot]))
                    arena.v_views[slot].copy_(st["exp_avg_sq"].to(arena.device, arena.dtype).view_as(arena.v_views[slot]))
                s = float(st["step"]) if not torch.is_tensor(st["step"]) else float(st["step"].item())
                if step_val is not None and s != step_val:
                    raise ValueError("B200AdamW: parameters of one group carry different step counts")
                step_val = s
            if step_val is None:      # empty state: fresh optimizer
                step_val = 0.0
                if not tensors_restored:
                    arena.m.zero_()
                    arena.v.zero_()
            self._steps[gi] = int(step_val)
            self._step_tensors[gi].fill_(step_val)
            for k, val in saved.items():
                if k != "params":
                    group[k] = val
        if self._grad_sync is not None:
            self._grad_sync.moments_loaded()

    def set_step(self, group_index: int, step: int) -> None:
        """Used by the fast checkpoint loader, which restores the arenas with the unpack kernel."""
        self._steps[group_index] = int(step)
        self._step_tensors[group_index].fill_(float(step))
