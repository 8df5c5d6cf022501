"""GPU parity tests of global-norm gradient clipping folded into the optimizer kernel
(pr_grad_norm / pr_grad_norm_scattered + pr_adamw_args.grad_scale_dev) against the oracle's
restatement of the reference's clip_grad_norm_ (utils.py:84-89) and against torch itself.

Bars: the norm is a floating-point reduction whose order differs from the oracle's (fp32 per
thread, double above) -> relative tolerance 2e-6, stated below; everything downstream of a given
norm (coefficient, update) is bit-exact; two runs give identical bits (resume depends on it)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
NORM_RTOL = 2e-6


def _bits(t):
    from oracle import oracle as O
    return O.to_bits(t)


def _coef32(norm: float, max_norm: float) -> float:
    c = np.float32(max_norm) / (np.float32(norm) + np.float32(1e-6))
    return float(min(c, np.float32(1.0)))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("n", [1, 7, 8, 9, 1000, 4096 * 37 + 5, (1 << 22) + 3])
def test_grad_norm_vs_oracle(cuda_lib, oracle, dtype, n):
    from pyrecover_b200 import kernels as K
    g = (torch.randn(n, generator=torch.Generator().manual_seed(n)) * 3e-3).to(dtype)
    gn = K.GradNorm("cuda")
    gc = g.cuda()
    for max_norm in (1.0, 1e-3, 1e3):
        gn.compute(gc, max_norm)
        norm, coef, bad = gn.out[:3].tolist()
        o_norm, o_coef, o_bad = oracle.grad_norm(_bits(g) if dtype == torch.bfloat16 else g.numpy(), max_norm)
        assert abs(norm - o_norm) <= NORM_RTOL * o_norm, (norm, o_norm)
        assert bad == 0.0 and not o_bad
        assert coef == _coef32(norm, max_norm)              # bit-exact given the kernel's own norm
        assert abs(coef - o_coef) <= 4 * NORM_RTOL * o_coef
    first = gn.out.clone()
    gn.compute(gc, 1e3)
    assert torch.equal(first.view(torch.int32), gn.out.view(torch.int32)), "norm is not deterministic"
    # prescale (the 1/W of a mean taken later) scales the norm
    gn.compute(gc, 1.0, prescale=0.125)
    assert abs(gn.out[0].item() - 0.125 * first[0].item()) <= 1e-6 * first[0].item()


def test_grad_norm_nonfinite_zero_and_disabled(cuda_lib, oracle):
    from pyrecover_b200 import kernels as K
    gn = K.GradNorm("cuda")
    g = torch.zeros(4096, dtype=torch.bfloat16, device="cuda")
    gn.compute(g, 1.0)
    assert gn.out[:3].tolist() == [0.0, 1.0, 0.0]             # zero gradient: coefficient clamps to 1
    g[17] = float("inf")
    gn.compute(g, 1.0)
    norm, coef, bad, sticky = gn.out.tolist()
    # the reference raises here (error_if_nonfinite); the device writes the "withhold the update" coefficient
    assert bad == 1.0 and coef == -1.0 and not np.isfinite(norm) and sticky == 1.0
    assert oracle.grad_norm(_bits(g.cpu()), 1.0)[2] is True
    # K2 given that coefficient leaves p, m, v untouched -- flat, scattered, both numerics
    n = 4096
    p = (torch.randn(n) * 0.02).bfloat16().cuda()
    m = (torch.randn(n) * 1e-3).bfloat16().cuda()
    v = (torch.rand(n) * 1e-6).bfloat16().cuda()
    p0, m0, v0 = p.clone(), m.clone(), v.clone()
    for numerics in (0, 1):
        K.adamw_(p, g, m, v, lr=1e-3, step=2, numerics=numerics, grad_scale_dev=gn.coef)
        plan = K.AdamWPlan([0, 2048], [2048, 2048], n)
        plan.step(p, [g[:2048], g[2048:]], m, v, lr=1e-3, step=2, numerics=numerics, grad_scale_dev=gn.coef)
        torch.cuda.synchronize()
        assert torch.equal(p, p0) and torch.equal(m, m0) and torch.equal(v, v0)
    gn.compute(g, 1.0)
    assert gn.out[3].item() == 2.0                            # sticky count of non-finite norms
    g[17] = 2.0
    gn.compute(g, 0.0)                                        # max_norm <= 0: report the norm, never scale
    assert gn.out[:3].tolist() == [2.0, 1.0, 0.0]
    with pytest.raises(RuntimeError, match="16-B aligned"):
        gn.compute(g[1:], 1.0)


def test_grad_norm_scattered_equals_flat(cuda_lib, oracle):
    """Gradients left where autograd put them (ragged sizes, tensors larger than a tile)."""
    from pyrecover_b200 import kernels as K
    from pyrecover_b200.flat import FlatLayout
    shapes = [torch.Size(s) for s in [(3, 5), (16384 * 2 + 24,), (128, 64), (7,), (8,), (40000,)]]
    lay = FlatLayout(shapes, torch.bfloat16)
    gen = torch.Generator().manual_seed(5)
    grads = [(torch.randn(s.shape, generator=gen) * 2e-2).bfloat16().cuda() for s in lay.slots]
    flat = torch.cat([g.reshape(-1) for g in grads])
    plan = K.AdamWPlan([s.offset for s in lay.slots], [s.numel for s in lay.slots], lay.numel)
    a, b = K.GradNorm("cuda"), K.GradNorm("cuda")
    plan.grad_norm(grads, a, 0.5)
    pad = (-flat.numel()) % 8
    b.compute(torch.cat([flat, flat.new_zeros(pad)]), 0.5)
    o_norm, o_coef, _ = oracle.grad_norm(_bits(flat.cpu()), 0.5)
    for gn in (a, b):
        assert abs(gn.out[0].item() - o_norm) <= NORM_RTOL * o_norm
        assert gn.out[1].item() == _coef32(gn.out[0].item(), 0.5) and gn.out[1].item() < 1.0
    first = a.out.clone()
    plan.grad_norm(grads, a, 0.5)
    assert torch.equal(first.view(torch.int32), a.out.view(torch.int32))


@pytest.mark.parametrize("numerics", [0, 1])
def test_adamw_with_device_coefficient_is_bit_exact(cuda_lib, oracle, numerics):
    """K2 reading the coefficient from the device == K2 given the same coefficient by the host ==
    the oracle with the coefficient multiplied into the gradient scale in binary32."""
    from pyrecover_b200 import kernels as K
    n = 8 * 1000 + 3
    gen = torch.Generator().manual_seed(11)
    p = (torch.randn(n, generator=gen) * 0.02).bfloat16()
    g = (torch.randn(n, generator=gen) * 5e-2).bfloat16()
    m = (torch.randn(n, generator=gen) * 1e-3).bfloat16()
    v = (torch.rand(n, generator=gen) * 1e-6).bfloat16()
    gn = K.GradNorm("cuda")
    pad = (-n) % 8
    gn.compute(torch.cat([g, g.new_zeros(pad)]).cuda(), 1.0)
    coef = gn.out[1].item()
    assert 0.0 < coef < 1.0
    for gscale in (1.0, 0.5):
        pd, md, vd = p.cuda(), m.cuda(), v.cuda()
        K.adamw_(pd, g.cuda(), md, vd, lr=3e-4, step=7, numerics=numerics, grad_scale=gscale, grad_scale_dev=gn.coef)
        ph, mh, vh = p.cuda(), m.cuda(), v.cuda()
        K.adamw_(ph, g.cuda(), mh, vh, lr=3e-4, step=7, numerics=numerics,
                 grad_scale=float(np.float32(gscale) * np.float32(coef)))
        assert torch.equal(pd, ph) and torch.equal(md, mh) and torch.equal(vd, vh)
        pb, mb, vb = _bits(p), _bits(m), _bits(v)
        oracle.adamw_bf16(pb, _bits(g), mb, vb, lr=3e-4, step=7, mode=numerics, gscale=gscale, coef=coef)
        assert np.array_equal(_bits(pd), pb) and np.array_equal(_bits(md), mb) and np.array_equal(_bits(vd), vb)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 2.0 ** -6)])
def test_optimizer_clipping_vs_reference_clip_grad_norm(cuda_lib, dtype, tol):
    """B200AdamW(max_grad_norm=...) against the reference's own sequence on torch: clip_grad_norm_
    (utils.py:84-89) then torch.optim.AdamW(fused=True).step(), 4 steps with gradients well above
    the threshold.  fp32: same arithmetic up to reduction and operation order, compounding over the
    steps (2e-5 on the weights, 2e-6 on the norm).  bf16: torch rounds the norm
    and the clipped gradients to bf16, ours does neither -> a few bf16 ulps (2^-6 relative)."""
    from pyrecover_b200.optim import B200AdamW
    from utils import clip_grad_norm_
    torch.manual_seed(3)
    shapes = [(64, 48), (48,), (300, 17), (5,)]
    init = [torch.randn(s) * 0.05 for s in shapes]
    ours = [torch.nn.Parameter(t.clone().to(dtype).cuda()) for t in init]
    ref = [torch.nn.Parameter(t.clone().to(dtype).cuda()) for t in init]
    opt = B200AdamW(ours, lr=1e-2, fused=True, max_grad_norm=0.25)
    ropt = torch.optim.AdamW(ref, lr=1e-2, fused=True)
    for step in range(4):
        grads = [torch.randn(s, generator=torch.Generator().manual_seed(100 * step + i)).to(dtype).cuda() * 0.3
                 for i, s in enumerate(shapes)]
        for q, r, g in zip(ours, ref, grads):
            q.grad, r.grad = g.clone(), g.clone()
        total = clip_grad_norm_(ref, 0.25)
        ropt.step()
        opt.step()
        st = opt.grad_norm_state().out[:3].tolist()
        assert st[2] == 0.0 and st[1] < 1.0
        assert abs(st[0] - float(total)) <= (2.0 ** -8 if dtype == torch.bfloat16 else NORM_RTOL) * float(total)
        for q, r in zip(ours, ref):
            assert torch.allclose(q.detach().float(), r.detach().float(), rtol=tol, atol=tol * 1e-2), step
    # clipping off: the coefficient path is not taken at all
    assert B200AdamW([torch.nn.Parameter(torch.zeros(8, device="cuda"))], lr=1e-3).grad_norm_state() is None
    with pytest.raises(ValueError):
        B200AdamW([torch.nn.Parameter(torch.zeros(8, device="cuda"))], lr=1e-3, max_grad_norm=0.0)
