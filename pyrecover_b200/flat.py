"""Flat HBM arenas for parameters, gradients and AdamW moments.

Data layout ("HBM layout"): all trainable tensors of one dtype live back to back
in ONE contiguous arena, each tensor starting on a 256-byte boundary, in the order of
`model.parameters()` (which is also state_dict order, so a checkpoint slice is a byte range).
Four arenas share the layout: p (parameters; nn.Parameter.data become views), g (gradients;
the communication buffer of K1), m and v (AdamW moments).  Padding elements are zero and stay
zero under AdamW (p=g=m=v=0 is a fixed point), so kernels sweep the arena without a mask.

This replaces the reference's per-tensor storage (model built at train.py:100-101, one
autograd-allocated .grad per parameter, `zeros_like` moments created lazily inside
torch.optim.AdamW.step -- torch/optim/adam.py:168-184).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence

import torch

ALIGN_BYTES = 256


@dataclass
class Slot:
    index: int          # position in the optimizer's flattened parameter list
    offset: int         # element offset in the arena
    numel: int
    shape: torch.Size


class FlatLayout:
    """Element offsets of every tensor in an arena of one dtype."""

    def __init__(self, shapes: Sequence[torch.Size], dtype: torch.dtype, multiple_of: int = 1):
        esize = torch.empty((), dtype=dtype).element_size()
        align = ALIGN_BYTES // esize
        self.dtype = dtype
        self.esize = esize
        self.slots: List[Slot] = []
        cur = 0
        for i, shp in enumerate(shapes):
            n = 1
            for d in shp:
                n *= int(d)
            self.slots.append(Slot(i, cur, n, torch.Size(shp)))
            cur += (n + align - 1) // align * align
        # total is a multiple of (align * multiple_of) so it splits into equal 16-B aligned
        # slices for `multiple_of` ranks
        unit = align * max(1, multiple_of)
        self.numel = max(unit, (cur + unit - 1) // unit * unit)
        self.payload = sum(s.numel for s in self.slots)

    @property
    def nbytes(self) -> int:
        return self.numel * self.esize

    def views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        assert flat.numel() == self.numel and flat.dtype == self.dtype
        return [flat[s.offset:s.offset + s.numel].view(s.shape) for s in self.slots]


class FlatArena:
    """p / g / m / v arenas for a list of parameters of ONE dtype on ONE CUDA device."""

    def __init__(self, params: Sequence[torch.nn.Parameter], world: int = 1,
                 alloc: Optional[Callable[[str, int, torch.dtype], torch.Tensor]] = None):
        params = list(params)
        if not params:
            raise ValueError("FlatArena needs at least one parameter")
        dtype, device = params[0].dtype, params[0].device
        for p in params:
            if p.dtype != dtype or p.device != device:
                raise ValueError("all parameters of an arena must share dtype and device")
        if device.type != "cuda":
            raise RuntimeError("FlatArena lives in HBM: parameters must be CUDA tensors (no CPU fallback)")
        self.params = params
        self.dtype, self.device = dtype, device
        self.layout = FlatLayout([p.shape for p in params], dtype, multiple_of=world)
        n = self.layout.numel
        self._alloc = alloc
        self._g = None if alloc is None else alloc("g", n, dtype)   # communication buffer: eager
        self._g_views = None
        if alloc is None:
            def alloc(name, numel, dt):
                return torch.zeros(numel, dtype=dt, device=device)
            self._alloc = alloc
        self.p = alloc("p", n, dtype)
        self.m = alloc("m", n, dtype)
        self.v = alloc("v", n, dtype)
        self.p_views = self.layout.views(self.p)
        self.m_views = self.layout.views(self.m)
        self.v_views = self.layout.views(self.v)
        with torch.no_grad():
            torch._foreach_copy_(self.p_views, [q.data for q in params])
            for q, view in zip(params, self.p_views):
                q.data = view            # the module now reads/writes the arena directly

    @property
    def numel(self) -> int:
        return self.layout.numel

    @property
    def g(self) -> torch.Tensor:
        """Flat gradient arena.  At world size 1 K2 reads gradients in place, so it is only
        allocated when something asks for it (a non-contiguous gradient, the flat K2 path)."""
        if self._g is None:
            self._g = self._alloc("g", self.layout.numel, self.dtype)
        return self._g

    @property
    def g_views(self) -> List[torch.Tensor]:
        if self._g_views is None:
            self._g_views = self.layout.views(self.g)
        return self._g_views

    def gather_grads(self) -> bool:
        """Copy every p.grad that is not already the arena view into the g arena (one
        multi-tensor copy).  Returns False when some parameter has no gradient."""
        src, dst = [], []
        complete = True
        for q, gv in zip(self.params, self.g_views):
            if q.grad is None:
                complete = False
                continue
            if q.grad.data_ptr() != gv.data_ptr():
                src.append(q.grad)
                dst.append(gv)
        if src:
            torch._foreach_copy_(dst, src)
        return complete

    def grads_in_place(self):
        """Per-tensor gradient tensors for the scattered-gradient K2: p.grad itself when it is a
        contiguous, 16-byte aligned tensor of the arena dtype, else a copy in the g arena slot.
        Returns None when some parameter has no gradient."""
        out = []
        for i, q in enumerate(self.params):
            g = q.grad
            if g is None:
                return None
            if g.dtype != self.dtype or not g.is_contiguous() or (g.data_ptr() & 15):
                gv = self.g_views[i]
                gv.copy_(g)
                g = gv
            out.append(g)
        return out

    def named_slots(self, names: Sequence[str]) -> Dict[str, Slot]:
        return {n: s for n, s in zip(names, self.layout.slots)}
