"""CPU: pin the oracle (oracle/pr_oracle.c) against vectors produced by the reference stack
(scripts/make_golden.py).  Bit-exact: integer/bit-pattern comparisons throughout."""
import hashlib

import numpy as np
import pytest
import torch


def _load(golden_dir, name):
    return np.load(golden_dir / name)


@pytest.mark.parametrize("tag,mode", [("single", 1), ("fused", 0)])
def test_adamw_bf16_bit_exact_vs_torch(oracle, golden_dir, tag, mode):
    """orc_adamw_bf16 == torch.optim.AdamW(fused=False|True) (reference train.py:120-122,274)
    for 12 warm-up steps of the reference LR schedule, every step, every element."""
    G = _load(golden_dir, "adamw_bf16.npz")
    b1, b2, eps, wd = G["hyper"]
    p = G["p0"].copy()
    m = np.zeros_like(p)
    v = np.zeros_like(p)
    for s, lr in enumerate(G["lrs"]):
        oracle.adamw_bf16(p, G["grads"][s], m, v, lr=lr, beta1=b1, beta2=b2, eps=eps, wd=wd,
                          step=s + 1, mode=mode)
        assert np.array_equal(m, G[f"{tag}_m_{s + 1}"]), f"exp_avg differs at step {s + 1}"
        assert np.array_equal(v, G[f"{tag}_v_{s + 1}"]), f"exp_avg_sq differs at step {s + 1}"
        assert np.array_equal(p, G[f"{tag}_p_{s + 1}"]), f"param differs at step {s + 1}"


def test_adamw_f32_vs_torch(oracle, golden_dir):
    """fp32 arenas: bit-exact against torch fused=True (CPU kernel); fused=False evaluates
    the same update with a different association, tolerance rtol 2e-6 (stated)."""
    G = _load(golden_dir, "adamw_f32.npz")
    b1, b2, eps, wd = G["hyper"]
    p = G["p0"].copy()
    m = np.zeros_like(p)
    v = np.zeros_like(p)
    for s, lr in enumerate(G["lrs"]):
        oracle.adamw_f32(p, G["grads"][s], m, v, lr=lr, beta1=b1, beta2=b2, eps=eps, wd=wd, step=s + 1)
        assert np.array_equal(m.view(np.uint32), G[f"fused_m_{s + 1}"].view(np.uint32))
        assert np.array_equal(v.view(np.uint32), G[f"fused_v_{s + 1}"].view(np.uint32))
        assert np.array_equal(p.view(np.uint32), G[f"fused_p_{s + 1}"].view(np.uint32))
        np.testing.assert_allclose(p, G[f"single_p_{s + 1}"], rtol=2e-6, atol=1e-9)
        np.testing.assert_allclose(m, G[f"single_m_{s + 1}"], rtol=2e-6, atol=1e-12)


@pytest.mark.parametrize("world", [2, 4])
def test_allreduce_vs_ddp_gloo(oracle, golden_dir, world):
    """Reference: DistributedDataParallel on gloo (train.py:110-115,269).  Its mean is
    sum_r bf16(g_r / W) accumulated in bf16 (W-1 roundings, each <= 2^-9 relative to a partial
    sum that never exceeds S = sum_r |g_r| / W); ours is ONE rounding of the fp32 sum.  For W a
    power of two the pre-scale is exact, hence |ours - ref| <= W * 2^-9 * S.  (SURVEY.md
    section 7 quotes a tighter constant that gloo's own output violates.)"""
    G = _load(golden_dir, f"ddp_mean_w{world}.npz")
    bufs = [np.ascontiguousarray(G["local"][r]) for r in range(world)]
    ours = oracle.from_bits(oracle.allreduce_bf16(bufs, 1.0 / world)).float()
    ref = oracle.from_bits(G["mean"]).float()
    S = torch.stack([oracle.from_bits(b).float().abs() for b in bufs]).sum(0) / world
    bound = world * 2.0 ** -9 * S + 1e-30
    assert bool(((ours - ref).abs() <= bound).all())
    ddp_order = oracle.allreduce_bf16_ddp(bufs)
    if world == 2:      # two addends commute: the restatement is bit-equal to gloo
        assert np.array_equal(ddp_order, G["mean"])
    else:               # gloo's ring rotates the accumulation order per chunk: bound only
        d = (oracle.from_bits(ddp_order).float() - ref).abs()
        assert bool((d <= 2 * bound).all())


def test_allreduce_f32_is_plain_sum(oracle):
    rng = np.random.default_rng(0)
    bufs = [rng.standard_normal(1000).astype(np.float32) for _ in range(8)]
    out = oracle.allreduce_f32(bufs, 0.125)
    acc = np.zeros(1000, np.float32)
    for b in bufs:
        acc = acc + b
    assert np.array_equal(out, acc * np.float32(0.125))


def test_checksum_known_answers(oracle, golden_dir):
    K = _load(golden_dir, "checksum_kat.npz")
    for i in range(int(K["count"][0])):
        data = K[f"data_{i}"]
        base, A, B = (int(x) for x in K[f"meta_{i}"])
        assert oracle.checksum64(data, base) == (A, B)
        assert oracle.checksum64_numpy(data, base) == (A, B)


def test_checksum_chunks_combine(oracle):
    """Checksum of checksums: partial sums taken with their own word_base add up."""
    rng = np.random.default_rng(5)
    data = rng.integers(0, 256, size=4 * 10007, dtype=np.uint8)
    whole = oracle.checksum64(data)
    cuts = [0, 4 * 13, 4 * 4096, 4 * 9000, data.size]
    parts = [oracle.checksum64(data[a:b], a // 4) for a, b in zip(cuts[:-1], cuts[1:])]
    assert oracle.combine(parts) == whole
    flipped = data.copy()
    flipped[12345] ^= 0x10
    assert oracle.checksum64(flipped) != whole
    swapped = data.copy()
    swapped[0:4], swapped[4:8] = data[4:8].copy(), data[0:4].copy()
    if not np.array_equal(swapped, data):
        assert oracle.checksum64(swapped) != whole


def test_pack_roundtrip_and_ragged(oracle):
    rng = np.random.default_rng(3)
    sizes = [0, 1, 15, 16, 17, 4096, 70001]
    segs = [rng.integers(0, 256, size=s, dtype=np.uint8) for s in sizes]
    offs, cur = [], 0
    for s in sizes:
        offs.append(cur)
        cur += (s + 63) // 64 * 64
    image, csum = oracle.pack(segs, offs, cur)
    assert csum == oracle.checksum64(image)
    back = oracle.unpack(image, sizes, offs)
    for a, b in zip(segs, back):
        assert np.array_equal(a, b)


def test_reference_written_checkpoint_fixture(golden_dir):
    """The fixture written by the reference's save_ckpt_vanilla is a torch.load-able file with
    the md5 side-car the reference produces (pyrecover/checkpoint.py:74-84)."""
    path = golden_dir / "ref_vanilla.pt"
    assert hashlib.md5(path.read_bytes()).hexdigest() == (golden_dir / "ref_vanilla.pt.md5").read_text()
    ck = torch.load(path, map_location="cpu")
    assert set(ck) == {"epoch", "step", "model", "optimizer", "lr_scheduler"}
    assert ck["step"] == 3 and ck["epoch"] == 1
    assert set(ck["optimizer"]["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}


@pytest.mark.parametrize("tag", ["big", "small", "tight"])
def test_grad_norm_and_clip_vs_reference_clip_grad_norm(oracle, golden_dir, tag):
    """oracle.grad_norm == the reference's clip_grad_norm_ (utils.py:84-89; vectors made by
    scripts/make_golden.py --clip-only from /root/reference/utils.py).
    fp32 gradients: norm within 1e-6 (torch reduces in fp32, the oracle in double) and the clipped
    gradients g * coef equal to torch's within 1 fp32 ulp.  bf16 gradients: torch rounds every
    per-tensor norm and the total to bf16 (2^-8 each), which the oracle does not -- 2^-7 stated."""
    G = _load(golden_dir, "clip_grad_norm.npz")
    max_norm = float(G[f"{tag}_f32_max_norm"][0])
    g = G[f"{tag}_f32_grads"]
    norm, coef, bad = oracle.grad_norm(g, max_norm)
    total = float(G[f"{tag}_f32_total"][0])
    assert not bad and abs(norm - total) <= 1e-6 * total
    assert coef == (1.0 if total + 1e-6 <= max_norm else coef) and 0.0 < coef <= 1.0
    mine = g * np.float32(coef)
    assert np.allclose(mine, G[f"{tag}_f32_clipped"], rtol=3e-7, atol=0)
    if tag == "small":
        assert coef == 1.0 and np.array_equal(G[f"{tag}_f32_clipped"], g)      # below the threshold: untouched
    else:
        assert coef < 1.0
    gb = G[f"{tag}_bf16_grads"]
    nb, cb, _ = oracle.grad_norm(gb, max_norm)
    tb = float(G[f"{tag}_bf16_total"][0])
    assert abs(nb - tb) <= 2.0 ** -7 * tb
    # prescale multiplies the norm
    assert abs(oracle.grad_norm(g, max_norm, prescale=0.25)[0] - 0.25 * norm) <= 1e-6 * norm
