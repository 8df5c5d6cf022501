"""GPU tests of the host-side mirror of the reference interface: B200AdamW (torch.optim protocol),
the pyrecover.checkpoint functions on the B200 snapshot path, and train.py end to end."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _tiny_model(seed=0, dtype=torch.bfloat16):
    from pyrecover_b200.workload import Transformer, preset_args
    torch.manual_seed(seed)
    torch.set_default_dtype(dtype)
    try:
        m = Transformer(preset_args("tiny", 64)).cuda()
    finally:
        torch.set_default_dtype(torch.float32)
    return m


def _run_steps(model, opt, sched, steps, seed=0, start=0):
    from pyrecover_b200.workload import SyntheticTokens
    data = SyntheticTokens(1024, 64, 2, seed=seed)
    for s in range(start, start + steps):
        x, y = data.batch(s + 1)
        opt.zero_grad()
        logits = model(x.cuda())
        loss = torch.nn.functional.cross_entropy(logits.flatten(0, 1).float(), y.cuda().flatten(), reduction="sum") / y.numel()
        loss.backward()
        opt.step()
        sched.step()
    return loss


def _build(seed=0, fused=True):
    from pyrecover_b200.optim import B200AdamW
    from utils import build_lr_scheduler
    model = _tiny_model(seed)
    opt = B200AdamW(model.parameters(), lr=1e-3, fused=fused)
    return model, opt, build_lr_scheduler(opt, 10)


@pytest.mark.parametrize("fused", [True, False])
def test_b200_adamw_matches_oracle_over_training(cuda_lib, oracle, fused):
    """Same gradients through B200AdamW (one launch over the flat arena) and through the oracle,
    tensor by tensor, with the reference's warm-up schedule: bit-exact after 4 steps."""
    model, opt, sched = _build(fused=fused)
    ref = {n: [oracle.to_bits(p.data), np.zeros(p.numel(), np.uint16), np.zeros(p.numel(), np.uint16)]
           for n, p in model.named_parameters()}
    from pyrecover_b200.workload import SyntheticTokens
    data = SyntheticTokens(1024, 64, 2)
    for s in range(4):
        x, y = data.batch(s + 1)
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(x.cuda()).flatten(0, 1).float(), y.cuda().flatten())
        loss.backward()
        lr = opt.param_groups[0]["lr"]
        grads = {n: oracle.to_bits(p.grad) for n, p in model.named_parameters()}
        opt.step()
        sched.step()
        for n, p in model.named_parameters():
            pb, mb, vb = ref[n]
            pb = pb.reshape(-1)
            oracle.adamw_bf16(pb, grads[n].reshape(-1), mb, vb, lr=lr, step=s + 1, mode=0 if fused else 1)
            assert np.array_equal(oracle.to_bits(p.data).reshape(-1), pb), f"{n} step {s + 1}"
            st = opt.state[p]
            assert np.array_equal(oracle.to_bits(st["exp_avg"]).reshape(-1), mb)
            assert np.array_equal(oracle.to_bits(st["exp_avg_sq"]).reshape(-1), vb)
            assert float(st["step"]) == s + 1


def test_state_dict_interchanges_with_torch_adamw(cuda_lib):
    """Key-for-key torch.optim.AdamW compatibility in both directions."""
    model, opt, sched = _build()
    _run_steps(model, opt, sched, 3)
    sd = opt.state_dict()
    twin = _tiny_model()
    topt = torch.optim.AdamW(twin.parameters(), lr=1e-3, fused=True)
    tsd_keys = None
    twin.load_state_dict(model.state_dict())
    topt.load_state_dict(sd)                       # ours -> torch
    for (n, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        assert torch.equal(opt.state[p]["exp_avg"], topt.state[q]["exp_avg"])
        assert float(topt.state[q]["step"]) == 3.0
    tsd = topt.state_dict()
    assert set(tsd["param_groups"][0]) == set(sd["param_groups"][0])
    assert set(tsd["state"][0]) == set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}
    model2, opt2, _ = _build(seed=1)
    opt2.load_state_dict(tsd)                      # torch -> ours
    for p, q in zip(model2.parameters(), twin.parameters()):
        assert torch.equal(opt2.state[p]["exp_avg_sq"], topt.state[q]["exp_avg_sq"])
    assert opt2._steps[0] == 3


def _equal_models(a, b):
    return all(torch.equal(p, q) for p, q in zip(a.parameters(), b.parameters()))


@pytest.mark.parametrize("mode", ["staged", "direct"])
def test_vanilla_save_is_torch_loadable_and_resume_bit_exact(cuda_lib, oracle, tmp_path, mode, monkeypatch):
    import pyrecover
    from pyrecover import checkpoint as ck
    monkeypatch.setenv("PYRECOVER_B200_SNAPSHOT", mode)
    model, opt, sched = _build()
    _run_steps(model, opt, sched, 3)
    path = tmp_path / "exp" / "ckpt_3.pt"
    path.parent.mkdir()
    out = pyrecover.save_ckpt_vanilla(model, opt, sched, None, 3, 1, path, max_keep=3, verify=True)
    assert out == path
    _run_steps(model, opt, sched, 2, start=3)      # training continues while the writer works
    pyrecover.wait_for_checkpoints()
    # (1) the file is what the reference's tools expect: bare torch.load, top-level "model" dict
    ref_model, ref_opt, ref_sched = _build()
    _run_steps(ref_model, ref_opt, ref_sched, 3)
    blob = torch.load(path, map_location="cuda")
    assert set(blob) >= {"epoch", "step", "model", "optimizer", "lr_scheduler"}
    for (n, p) in ref_model.state_dict().items():
        assert torch.equal(blob["model"][n], p), n
    assert blob["optimizer"]["param_groups"][0]["lr"] == ref_opt.param_groups[0]["lr"]
    assert float(blob["optimizer"]["state"][5]["step"]) == 3.0
    import hashlib
    assert hashlib.md5(path.read_bytes()).hexdigest() == Path(str(path) + ".md5").read_text()
    # (2) the stored PR-SUM64 equals the oracle's checksum of the image bytes in the file
    from pyrecover_b200 import ptfile
    side = json.loads(Path(str(path) + ".b200sum").read_text())
    offs = ptfile.record_offsets(str(path))
    raw = path.read_bytes()
    image = b"".join(raw[o:o + n] for k, (o, n) in offs.items() if k.startswith("data/") and not k.startswith("data/s"))
    assert list(oracle.checksum64(np.frombuffer(image, np.uint8))) == side["sum64"]
    # (3) resume into a differently-initialised model: bit-exact state, then identical training
    m2, o2, s2 = _build(seed=123)
    epoch, step = pyrecover.load_ckpt_vanilla(m2, o2, s2, None, "latest", experiment_dir=path.parent, verify=True)
    assert (epoch, step) == (1, 3)
    assert _equal_models(m2, ref_model)
    assert torch.equal(o2.arenas[0].m, ref_opt.arenas[0].m) and torch.equal(o2.arenas[0].v, ref_opt.arenas[0].v)
    assert o2._steps[0] == 3 and s2.last_epoch == ref_sched.last_epoch
    _run_steps(m2, o2, s2, 2, start=3)
    assert _equal_models(m2, model)
    # (4) a corrupted file is caught by the on-GPU checksum
    bad = bytearray(raw)
    first = min(o for k, (o, n) in offs.items() if k == "data/0")
    bad[first + 1000] ^= 0x40
    path.write_bytes(bytes(bad))
    with pytest.raises(RuntimeError, match="Checksum mismatch"):
        pyrecover.load_ckpt_vanilla(m2, o2, s2, None, str(path), verify=True)
    ck.release_checkpoint_engines()


def test_reference_written_checkpoint_loads(cuda_lib, golden_dir, tmp_path):
    """A file written by the reference's own save_ckpt_vanilla (fixture) restores through our
    load_ckpt_vanilla into a B200AdamW, md5 side-car verified; and our loader's result equals
    torch.load of the same file."""
    import shutil

    import pyrecover
    from pyrecover_b200.optim import B200AdamW
    from utils import build_lr_scheduler
    for f in ("ref_vanilla.pt", "ref_vanilla.pt.md5"):
        shutil.copy(golden_dir / f, tmp_path / f)
    torch.manual_seed(99)
    model = torch.nn.Sequential(torch.nn.Linear(24, 16, bias=False), torch.nn.Linear(16, 8, bias=True)).to(torch.bfloat16).cuda()
    opt = B200AdamW(model.parameters(), lr=1e-5)
    sched = build_lr_scheduler(opt, 10)
    epoch, step = pyrecover.load_ckpt_vanilla(model, opt, sched, None, str(tmp_path / "ref_vanilla.pt"), verify=True)
    assert (epoch, step) == (1, 3)
    blob = torch.load(tmp_path / "ref_vanilla.pt", map_location="cuda")
    for n, p in model.state_dict().items():
        assert torch.equal(p, blob["model"][n])
    for i, p in enumerate(model.parameters()):
        assert torch.equal(opt.state[p]["exp_avg"], blob["optimizer"]["state"][i]["exp_avg"])
    assert opt._steps[0] == 3 and sched.last_epoch == blob["lr_scheduler"]["last_epoch"]
    (tmp_path / "ref_vanilla.pt.md5").write_text("0" * 32)
    with pytest.raises(RuntimeError, match="Checksum mismatch"):
        pyrecover.load_ckpt_vanilla(model, opt, sched, None, str(tmp_path / "ref_vanilla.pt"), verify=True)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_sharded_save_load_roundtrip(cuda_lib, tmp_path, world):
    """Every rank's shard written (here: one process plays all ranks in turn, rank 0 last because it waits
    for the others before it writes .metadata), then restored into a fresh model: bit-exact; per-shard
    PR-SUM64 verified on the GPU; a flipped byte is caught.  The directory is a torch.distributed.checkpoint
    checkpoint: stock dist_cp.load fills a CPU state dict with the same bits (the reference's
    load_ckpt_distributed, pyrecover/checkpoint.py:352-356, is that call)."""
    import pyrecover
    from pyrecover import checkpoint as ck
    model, opt, sched = _build()
    _run_steps(model, opt, sched, 2)
    d = tmp_path / "exp" / "ckpt_2"
    snap = ck._snapshotter(model, opt)
    small = {"epoch": 1, "step": 2, "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in opt.param_groups],
             "opt_steps": list(opt._steps), "rng_state": ck._rng_state(), "lr_scheduler": sched.state_dict()}
    for r in reversed(range(world)):
        snap.save_sharded(str(d), small, r, world, max_keep=3, blocking=True)
        assert (pyrecover.get_latest_checkpoint(d.parent, distributed=True) == str(d)) == (r == 0)
    total = sum((d / f"__{r}_0.distcp").stat().st_size for r in range(world))
    assert total >= 3 * opt.arenas[0].layout.payload * 2
    # stock torch.distributed.checkpoint reads it, with the reference's key names
    import torch.distributed.checkpoint as dcp
    prefix = "module." if world > 1 else ""
    tmpl = {"model": {prefix + k: torch.zeros_like(v, device="cpu") for k, v in model.state_dict().items()},
            "optimizer": {"state": {i: {"step": torch.zeros(()), "exp_avg": torch.zeros_like(p, device="cpu"),
                                        "exp_avg_sq": torch.zeros_like(p, device="cpu")}
                                    for i, p in enumerate(model.parameters())},
                          "param_groups": [{"lr": 0.0, "betas": (0.0, 0.0), "params": []}]},
            "metadata": {"epoch": 0, "step": 0}, "lr_scheduler": {"last_epoch": -1}}
    dcp.load(state_dict=tmpl, checkpoint_id=str(d), no_dist=True)
    for k, v in model.state_dict().items():
        assert torch.equal(tmpl["model"][prefix + k], v.cpu()), k
    for i, p in enumerate(model.parameters()):
        assert torch.equal(tmpl["optimizer"]["state"][i]["exp_avg"], opt.state[p]["exp_avg"].cpu())
        assert torch.equal(tmpl["optimizer"]["state"][i]["exp_avg_sq"], opt.state[p]["exp_avg_sq"].cpu())
        assert float(tmpl["optimizer"]["state"][i]["step"]) == 2.0
    assert tmpl["metadata"] == {"epoch": 1, "step": 2} and tmpl["lr_scheduler"]["last_epoch"] == sched.last_epoch
    assert tmpl["optimizer"]["param_groups"][0]["lr"] == opt.param_groups[0]["lr"]
    assert tmpl["optimizer"]["param_groups"][0]["params"] == list(range(len(list(model.parameters()))))
    m2, o2, s2 = _build(seed=5)
    epoch, step = pyrecover.load_ckpt_distributed(m2, o2, s2, None, "latest", experiment_dir=d.parent)
    assert (epoch, step) == (1, 2) and _equal_models(m2, model)
    assert torch.equal(o2.arenas[0].v, opt.arenas[0].v) and o2._steps[0] == 2
    assert s2.last_epoch == sched.last_epoch
    import json
    f = d / f"__{world - 1}_0.distcp"
    raw = bytearray(f.read_bytes())
    payload = json.loads((d / f"b200_{world - 1:05d}.json").read_text())["pieces"][0][2]
    raw[payload + 40] ^= 1
    f.write_bytes(bytes(raw))
    with pytest.raises(RuntimeError, match="Checksum mismatch"):
        pyrecover.load_ckpt_distributed(m2, o2, s2, None, str(d))
    (d / ".metadata").unlink()
    assert pyrecover.get_latest_checkpoint(d.parent, distributed=True) is None
    ck.release_checkpoint_engines()


def test_sharded_directory_converts_offline_to_vanilla_and_dcp(cuda_lib, tmp_path):
    """pyrecover.convert reads a sharded directory with nothing but torch on the CPU (the layout table
    travels in b200_meta.pt): the vanilla file it writes equals save_ckpt_vanilla's tensors, loads back
    through load_ckpt_vanilla, and the DCP directory it writes loads with torch.distributed.checkpoint."""
    import pyrecover
    from pyrecover import checkpoint as ck
    from pyrecover import convert
    model, opt, sched = _build()
    _run_steps(model, opt, sched, 3)
    d = tmp_path / "exp" / "ckpt_3"
    pyrecover.save_ckpt_distributed(model, opt, sched, None, 3, 1, d, max_keep=3, verify=True)
    pyrecover.save_ckpt_vanilla(model, opt, sched, None, 3, 1, tmp_path / "direct.pt", max_keep=0, verify=False)
    ck.wait_for_checkpoints()
    state = convert.load_state_cpu(str(d))
    direct = torch.load(tmp_path / "direct.pt", map_location="cpu")
    assert list(state["model"]) == list(direct["model"]) and state["step"] == 3 and state["epoch"] == 1
    for k in direct["model"]:
        assert torch.equal(state["model"][k], direct["model"][k]), k
    for i, st in direct["optimizer"]["state"].items():
        for key in ("exp_avg", "exp_avg_sq"):
            assert torch.equal(state["optimizer"]["state"][i][key], st[key]), (i, key)
        assert float(state["optimizer"]["state"][i]["step"]) == float(st["step"])
    assert state["optimizer"]["param_groups"][0]["params"] == direct["optimizer"]["param_groups"][0]["params"]
    assert convert.main([str(d), str(tmp_path / "conv.pt"), "--to", "vanilla"]) == 0
    m2, o2, s2 = _build(seed=9)
    assert pyrecover.load_ckpt_vanilla(m2, o2, s2, None, str(tmp_path / "conv.pt"), verify=False) == (1, 3)
    assert _equal_models(m2, model) and torch.equal(o2.arenas[0].m, opt.arenas[0].m) and o2._steps[0] == 3
    assert convert.main([str(d), str(tmp_path / "dcp" / "ckpt_3"), "--to", "dcp"]) == 0
    import torch.distributed.checkpoint as dcp
    tmpl = {"model": {k: torch.zeros_like(v) for k, v in direct["model"].items()}}
    dcp.load(state_dict=tmpl, checkpoint_id=str(tmp_path / "dcp" / "ckpt_3"), no_dist=True)
    for k in direct["model"]:
        assert torch.equal(tmpl["model"][k], direct["model"][k]), k
    ck.release_checkpoint_engines()


def _train(tmp, name, extra, steps):
    cmd = [sys.executable, str(ROOT / "train.py"), "--synthetic-data", "--model-preset", "tiny", "--sequence-length", "256",
           "--training-steps", str(steps), "--checkpoint-frequency", "10", "--checkpoint-dir", str(tmp), "--experiment_name", name,
           "--fused-optimizer", "--verify-checkpoints", "--logging-frequency", "10", "--learning-rate", "1e-3"] + extra
    r = subprocess.run(cmd, cwd=str(ROOT), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


@pytest.mark.parametrize("sharded", [False, True])
def test_train_straight_vs_interrupted_resumed_weight_equality(cuda_lib, tmp_path, sharded):
    """BASELINE config 1's experiment on the GPU path: 30 steps straight vs 20 steps, kill, resume
    to 30 -- final checkpoints equal with tolerance 0 (the reference's check_weights_equality.py
    criterion, tests/check_weights_equality.py:166-178, which its own runs cannot meet)."""
    extra = ["--use-torch-distributed-ckpt"] if sharded else []
    _train(tmp_path, "straight", extra, 30)
    _train(tmp_path, "resumed", extra, 20)
    _train(tmp_path, "resumed", extra + ["--resume-from-checkpoint", "latest"], 30)
    if sharded:
        import pyrecover
        a, oa, sa = _build_tiny256()
        b, ob, sb = _build_tiny256(seed=3)
        pyrecover.load_ckpt_distributed(a, oa, sa, None, str(tmp_path / "straight" / "ckpt_30"))
        pyrecover.load_ckpt_distributed(b, ob, sb, None, str(tmp_path / "resumed" / "ckpt_30"))
        assert _equal_models(a, b) and torch.equal(oa.arenas[0].m, ob.arenas[0].m)
        from pyrecover import checkpoint as ck
        ck.release_checkpoint_engines()
    else:
        A = torch.load(tmp_path / "straight" / "ckpt_30.pt", map_location="cuda")
        B = torch.load(tmp_path / "resumed" / "ckpt_30.pt", map_location="cuda")
        assert A["model"].keys() == B["model"].keys() and len(A["model"]) == 39
        for k in A["model"]:
            assert float((A["model"][k].float() - B["model"][k].float()).abs().max()) == 0.0, k
        for i in A["optimizer"]["state"]:
            assert torch.equal(A["optimizer"]["state"][i]["exp_avg_sq"], B["optimizer"]["state"][i]["exp_avg_sq"])
        assert A["lr_scheduler"] == B["lr_scheduler"]


def _build_tiny256(seed=0):
    from pyrecover_b200.optim import B200AdamW
    from pyrecover_b200.workload import Transformer, preset_args
    from utils import build_lr_scheduler
    torch.manual_seed(seed)
    torch.set_default_dtype(torch.bfloat16)
    try:
        m = Transformer(preset_args("tiny", 256)).cuda()
    finally:
        torch.set_default_dtype(torch.float32)
    o = B200AdamW(m.parameters(), lr=1e-3, fused=True)
    return m, o, build_lr_scheduler(o, 10)
