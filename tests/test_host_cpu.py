"""CPU tests (no GPU): the C-ABI library loads and exports every declared symbol, the boundary
(flags, pyrecover API), host logic (time-aware policy, rotation, archive writer, layouts) and
BASELINE config 1 (tiny model, DDP on gloo, world 2, checkpoint every 10 steps, resume equality)."""
import hashlib
import inspect
import json
import os
import re
import subprocess
import sys
import time
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]


# ------------------------------------------------------------------------- C ABI ----------
def _declared_symbols():
    text = (ROOT / "include" / "pyrecover_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pr_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    from pyrecover_b200 import _lib
    _lib.build()
    L = _lib.lib()
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/pyrecover_b200.h but not exported"
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared, "ctypes table and header disagree"
    assert L.pr_abi_version() == _lib.PR_ABI_VERSION
    assert L.pr_last_error() is not None


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under pyrecover_b200/ or pyrecover/ (nor the
    train.py boundary) may import or open it."""
    for base in ("pyrecover_b200", "pyrecover"):
        for f in (ROOT / base).rglob("*.py"):
            assert "oracle" not in f.read_text(), f
    for f in ("train.py", "utils.py", "dist_utils.py"):
        assert "oracle" not in (ROOT / f).read_text(), f


def test_kernels_refuse_cpu_tensors():
    from pyrecover_b200 import kernels as K
    t = torch.zeros(16, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        K.adamw_(t, t, t, t, lr=1e-3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        K.checksum(t)
    from pyrecover_b200.optim import B200AdamW
    with pytest.raises(RuntimeError, match="CUDA"):
        B200AdamW([torch.nn.Parameter(torch.zeros(4))], lr=1e-3)


# ------------------------------------------------------------------------- boundary -------
def test_reference_flags_names_types_defaults():
    """tests/golden/reference_flags.json was dumped from the reference's own get_args()
    (utils.py:105-261) by scripts/make_golden.py."""
    import utils
    ref = json.loads((ROOT / "tests" / "golden" / "reference_flags.json").read_text())
    ours = vars(utils.get_args([]))
    assert len(ref) == 28
    for k, v in ref.items():
        assert k in ours, k
        assert ours[k] == v, (k, ours[k], v)
    a = utils.get_args(["--use-torch-distributed-ckpt", "--timeaware-checkpointing", "--experiment_name=x",
                        "--sequence-length=128", "--resume-from-checkpoint", "latest", "--fused-optimizer"])
    assert a.use_torch_distributed_ckpt and a.timeaware_checkpointing and a.sequence_length == 128


def test_pyrecover_api_signatures_match_reference():
    """Positional order matters: the reference's train.py passes 4 (load) / 7 (save) positionally."""
    import pyrecover
    from pyrecover import checkpoint as ck
    ref = json.loads((ROOT / "tests" / "golden" / "reference_signatures.json").read_text())
    for name, params in ref.items():
        sig = inspect.signature(getattr(ck, name))
        got = [[p.name, None if p.default is inspect._empty else p.default] for p in sig.parameters.values()]
        assert got == params, name
    assert callable(pyrecover.setup_resubmission) and callable(pyrecover.get_remaining_time)


def test_lr_schedule_matches_reference_formula():
    import utils
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=1e-5)
    sched = utils.build_lr_scheduler(opt, 10)
    lrs = []
    for _ in range(13):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
    ref = np.load(ROOT / "tests" / "golden" / "adamw_bf16.npz")["lrs"]
    assert np.array_equal(np.array(lrs[:12]), ref)
    assert lrs[12] == 1e-5
    assert sched.state_dict()["lr_lambdas"] == [{"warmup_steps": 10}]


# ------------------------------------------------------------------------- time-aware -----
def test_time_aware_policy_thresholds():
    from pyrecover.timelimit import TimeAwarePolicy, get_remaining_time
    now = [1000.0]
    pol = TimeAwarePolicy(1.0, 10.0, end_time=1100.0, clock=lambda: now[0])
    assert pol.buffer_time == 10 * 1.0 + 2 * 10.0            # train.py:166-176
    assert pol.threshold == 1.0 + 10.0 + 30.0
    assert not pol.should_stop()
    now[0] = 1059.5
    assert pol.should_stop()                                 # 40.5 left < 41 (train.py:226-228)
    assert pol.observe_iteration(2.0) and pol.max_iter_time == 2.0
    assert pol.buffer_time == 5 * 2.0 + 1 * 10.0             # train.py:303
    assert not pol.observe_iteration(1.5)
    assert pol.observe_checkpoint(12.0) and pol.max_ckpt_time == 12.0
    assert TimeAwarePolicy(1.0, 10.0, end_time=None).should_stop() is False
    os.environ["SLURM_JOB_END_TIME"] = str(time.time() + 50)
    try:
        assert 49 < get_remaining_time() <= 50
        os.environ["SLURM_JOB_END_TIME"] = "garbage"
        assert get_remaining_time() is None
    finally:
        del os.environ["SLURM_JOB_END_TIME"]


def test_resubmission_handler_sets_flag():
    import signal

    from pyrecover.resubmit import setup_resubmission
    old = signal.getsignal(signal.SIGUSR1)
    try:
        seen = []
        st = setup_resubmission(None, signals=(signal.SIGUSR1,), on_signal=seen.append)
        os.kill(os.getpid(), signal.SIGUSR1)
        time.sleep(0.05)
        assert st.stop_requested and seen == [signal.SIGUSR1] and not st.resubmitted
    finally:
        signal.signal(signal.SIGUSR1, old)


# ------------------------------------------------------------------------- rotation -------
def test_rotation_is_by_step_number_and_scoped(tmp_path):
    from pyrecover_b200.snapshot import rotate_dirs, rotate_files
    for s in (90, 100, 200, 1000):
        (tmp_path / f"ckpt_{s}.pt").write_bytes(b"x")
        (tmp_path / f"ckpt_{s}.pt.md5").write_text("m")
    rotate_files(tmp_path, 2)
    assert sorted(p.name for p in tmp_path.glob("*.pt")) == ["ckpt_1000.pt", "ckpt_200.pt"]   # lexicographic would keep 90 & 200
    assert not (tmp_path / "ckpt_90.pt.md5").exists()
    rotate_files(tmp_path, 0)
    assert len(list(tmp_path.glob("*.pt"))) == 2
    for s in (5, 10, 15):
        d = tmp_path / f"ckpt_{s}"
        d.mkdir()
        (d / "shard_00000.bin").write_bytes(b"x")
        (d / "shard_00000.json").write_text("{}")
        (d / "COMPLETE_00000").write_text("1")
    (tmp_path / "logs").mkdir()
    # a directory a slower rank is still writing (2 ranks, one marker) is never pulled from under it
    busy = tmp_path / "ckpt_7"
    busy.mkdir()
    (busy / "shard_00001.bin.tmp").write_bytes(b"x")
    (busy / "COMPLETE_00000").write_text("2")
    rotate_dirs(tmp_path, 1)
    assert (tmp_path / "logs").exists() and (tmp_path / "ckpt_15").exists() and not (tmp_path / "ckpt_5").exists()
    assert not (tmp_path / "ckpt_10").exists() and busy.exists()
    old = time.time() - 3600                          # ... unless it has been dead for a long time
    for f in busy.iterdir():
        os.utime(f, (old, old))
    rotate_dirs(tmp_path, 1)
    assert not busy.exists()


# ------------------------------------------------------------------------- archive --------
def test_archive_writer_is_torch_loadable_and_aligned(tmp_path):
    from collections import OrderedDict

    from pyrecover_b200 import ptfile as PT
    n = 1000
    p, m = torch.randn(n).bfloat16(), torch.randn(n).bfloat16()
    image = torch.cat([p, m]).view(torch.uint8).numpy().tobytes()
    rec = PT.RecordRef("0", torch.bfloat16, 2 * n)
    srec = PT.RecordRef("s0", torch.float32, 1)
    state = {"epoch": 1, "step": 7,
             "model": OrderedDict(w=PT.TensorStub(rec, 0, (10, 50)), b=PT.TensorStub(rec, 500, (500,))),
             "optimizer": {"state": {0: {"step": PT.TensorStub(srec, 0, ()), "exp_avg": PT.TensorStub(rec, n, (10, 50))}},
                           "param_groups": [{"lr": 1e-5, "betas": (0.9, 0.999), "params": [0]}]}}
    path = tmp_path / "ckpt_7.pt"
    info = PT.write_archive(str(path), PT.dumps_state(state), [("0", image), ("s0", np.float32([7.0]).tobytes())], md5=True)
    assert info["md5"] == hashlib.md5(path.read_bytes()).hexdigest()
    assert all(o % 64 == 0 for o in info["offsets"].values())
    assert not Path(str(path) + ".tmp").exists()
    for kw in ({}, {"mmap": True}, {"weights_only": True}):
        ck = torch.load(path, map_location="cpu", **kw)
        assert torch.equal(ck["model"]["w"], p[:500].view(10, 50)) and torch.equal(ck["model"]["b"], p[500:])
        assert torch.equal(ck["optimizer"]["state"][0]["exp_avg"], m[:500].view(10, 50))
        assert float(ck["optimizer"]["state"][0]["step"]) == 7.0 and ck["step"] == 7
    offs = PT.record_offsets(str(path))
    assert offs["data/0"] == (info["offsets"]["0"], len(image))
    import zipfile
    assert zipfile.ZipFile(path).testzip() is None


@pytest.mark.parametrize("zip64", [False, True])
def test_archive_writer_streams_big_records_direct_and_zip64(tmp_path, monkeypatch, zip64):
    """The streamed-record path: page-aligned buffers go out with O_DIRECT where the filesystem
    takes it (buffered otherwise -- same bytes either way), the CRC comes from the helper thread,
    the md5 is that of the finished file, and the zip64 framing (forced here with a small limit;
    real checkpoints cross 2 GiB) is accepted by torch.load and by the zipfile module."""
    import mmap
    import zipfile
    from collections import OrderedDict

    from pyrecover_b200 import ptfile as PT
    if zip64:
        monkeypatch.setattr(PT, "_ZIP64_LIMIT", (1 << 20) - 1)
    chunk = 1 << 20
    bufs = []
    for i in range(3):
        m = mmap.mmap(-1, chunk)                      # page aligned, like the pinned arena
        m.write(np.random.default_rng(i).integers(0, 256, chunk, dtype=np.uint8).tobytes())
        bufs.append(m)
    tail = 4096 + 320                                 # last chunk: one aligned block + an unaligned rest
    total = 2 * chunk + tail
    parts = [memoryview(bufs[0]), memoryview(bufs[1]), memoryview(bufs[2])[:tail]]
    want = b"".join(bytes(x) for x in parts)
    rec = PT.RecordRef("0", torch.uint8, total)
    state = {"step": 3, "model": OrderedDict(head=PT.TensorStub(rec, 7, (100,)), tail=PT.TensorStub(rec, total - 50, (50,)))}
    out = {}
    for direct in (None, False):
        (tmp_path / str(direct)).mkdir()
        path = tmp_path / str(direct) / "big.pt"
        info = PT.write_archive(str(path), PT.dumps_state(state), [("0", (total, iter(parts)))], md5=True, direct=direct)
        raw = path.read_bytes()
        out[direct] = raw
        assert info["md5"] == hashlib.md5(raw).hexdigest() and info["bytes"] == len(raw)
        assert info["offsets"]["0"] % 4096 == 0
        assert raw[info["offsets"]["0"]:info["offsets"]["0"] + total] == want
        assert zipfile.ZipFile(path).testzip() is None
        for kw in ({"weights_only": True}, {"mmap": True}):
            ck = torch.load(path, map_location="cpu", **kw)
            assert bytes(ck["model"]["head"].numpy().tobytes()) == want[7:107]
            assert bytes(ck["model"]["tail"].numpy().tobytes()) == want[-50:]
        assert PT.record_offsets(str(path))["data/0"] == (info["offsets"]["0"], total)
    assert out[None] == out[False]
    # a short stream is an error and leaves nothing behind
    with pytest.raises(IOError):
        PT.write_archive(str(tmp_path / "short.pt"), PT.dumps_state(state), [("0", (total, iter(parts[:2])))])
    assert not (tmp_path / "short.pt").exists() and not (tmp_path / "short.pt.tmp").exists()
    # raw shard writer
    n = PT.write_raw(str(tmp_path / "shard.bin"), iter(parts))
    assert n == total and (tmp_path / "shard.bin").read_bytes() == want


def test_flat_layout_alignment_and_slices():
    from pyrecover_b200.flat import FlatLayout
    shapes = [torch.Size(s) for s in [(1024, 256), (128, 256), (256,), (7, 3)]]
    lay = FlatLayout(shapes, torch.bfloat16, multiple_of=8)
    assert all(s.offset * 2 % 256 == 0 for s in lay.slots)
    assert lay.numel % (128 * 8) == 0 and lay.payload == 1024 * 256 + 128 * 256 + 256 + 21
    flat = torch.arange(lay.numel, dtype=torch.float32).bfloat16()
    views = lay.views(flat)
    assert [tuple(v.shape) for v in views] == [tuple(s) for s in shapes]
    assert views[2].data_ptr() == flat.data_ptr() + lay.slots[2].offset * 2


# ------------------------------------------------------------------------- host path ------
def _small_model(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(24, 16, bias=False), torch.nn.Linear(16, 8)).to(torch.bfloat16)


def test_host_path_checkpoint_roundtrip_and_reference_fixture(tmp_path, golden_dir):
    """CPU models take the reference's own route (torch.save / torch.load): round trip is
    bit-exact, the .md5 side-car verifies, a reference-written file loads, latest-lookup works."""
    import shutil

    import pyrecover
    import utils
    model = _small_model(0)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=False)
    sched = utils.build_lr_scheduler(opt, 10)
    for _ in range(3):
        opt.zero_grad()
        model(torch.randn(4, 24).bfloat16()).float().pow(2).sum().backward()
        opt.step()
        sched.step()
    exp = tmp_path / "exp"
    exp.mkdir()
    for step in (1, 2, 3, 4):
        pyrecover.save_ckpt_vanilla(model, opt, sched, None, step, 1, exp / f"ckpt_{step}.pt", max_keep=3, verify=True)
    assert sorted(p.name for p in exp.glob("*.pt")) == ["ckpt_2.pt", "ckpt_3.pt", "ckpt_4.pt"]
    assert pyrecover.get_latest_checkpoint(exp) == str(exp / "ckpt_4.pt")
    m2 = _small_model(1)
    o2 = torch.optim.AdamW(m2.parameters(), lr=1e-5, fused=False)
    s2 = utils.build_lr_scheduler(o2, 10)
    assert pyrecover.load_ckpt_vanilla(m2, o2, s2, None, "latest", experiment_dir=exp, verify=True) == (1, 4)
    for a, b in zip(model.parameters(), m2.parameters()):
        assert torch.equal(a, b)
        assert torch.equal(opt.state[a]["exp_avg_sq"], o2.state[b]["exp_avg_sq"])
    assert s2.state_dict() == sched.state_dict()
    # reference-written fixture
    for f in ("ref_vanilla.pt", "ref_vanilla.pt.md5"):
        shutil.copy(golden_dir / f, tmp_path / f)
    assert pyrecover.load_ckpt_vanilla(m2, o2, s2, None, str(tmp_path / "ref_vanilla.pt"), verify=True) == (1, 3)
    with pytest.raises(RuntimeError, match="No checkpoint found"):
        pyrecover.load_ckpt_vanilla(m2, o2, s2, None, "latest", experiment_dir=tmp_path / "nothing")
    (exp / "ckpt_4.pt.md5").write_text("deadbeef")
    with pytest.raises(RuntimeError, match="Checksum mismatch"):
        pyrecover.load_ckpt_vanilla(m2, o2, s2, None, str(exp / "ckpt_4.pt"), verify=True)


def test_check_weights_equality_criterion_on_our_files(tmp_path):
    """What the reference's tests/check_weights_equality.py does in vanilla mode (:112-115,
    :166-178): bare torch.load(path)["model"], max |a-b| <= tolerance per tensor."""
    import pyrecover
    model = _small_model(3)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5)
    a, b = tmp_path / "a" / "ckpt_1.pt", tmp_path / "b" / "ckpt_1.pt"
    a.parent.mkdir()
    b.parent.mkdir()
    pyrecover.save_ckpt_vanilla(model, opt, None, None, 1, 1, a, verify=False)
    pyrecover.save_ckpt_vanilla(model, opt, None, None, 1, 1, b, verify=False)
    A, B = torch.load(a)["model"], torch.load(b)["model"]
    assert A.keys() == B.keys()
    assert all(float((A[k].float() - B[k].float()).abs().max()) <= 1e-7 for k in A)


# ------------------------------------------------------------------------- config 1 -------
def _launch(tmp, name, steps, extra=(), nproc=2):
    env = dict(os.environ, OMP_NUM_THREADS="2", MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + (os.getpid() % 200)),
           str(ROOT / "train.py"), "--distributed", "--device", "cpu", "--synthetic-data", "--model-preset", "tiny",
           "--sequence-length", "256", "--batch-size", str(nproc), "--training-steps", str(steps), "--checkpoint-frequency", "10",
           "--checkpoint-dir", str(tmp), "--experiment_name", name, "--verify-checkpoints", "--logging-frequency", "10",
           "--learning-rate", "1e-3", *extra]
    r = subprocess.run(cmd, cwd=str(ROOT), env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r


@pytest.mark.parametrize("sharded", [False, True])
def test_config1_gloo_world2_resume_weight_equality(tmp_path, sharded):
    """BASELINE.json config 1 as stated (dim 256, 4 layers, seq 256): tiny transformer, DDP on CPU/gloo world_size 2, synthetic tokens,
    checkpoint every 10 steps + resume: the interrupted+resumed run ends with exactly the weights
    of the straight run (tolerance 0)."""
    extra = ("--use-torch-distributed-ckpt",) if sharded else ()
    _launch(tmp_path, "straight", 30, extra)
    _launch(tmp_path, "resumed", 20, extra)
    out = _launch(tmp_path, "resumed", 30, extra + ("--resume-from-checkpoint", "latest"))
    assert "step 20" in out.stdout
    if sharded:
        import torch.distributed.checkpoint as dcp
        from torch.distributed.checkpoint.format_utils import dcp_to_torch_save
        for name in ("straight", "resumed"):
            dcp_to_torch_save(str(tmp_path / name / "ckpt_30"), str(tmp_path / f"{name}.pt"))
        A = torch.load(tmp_path / "straight.pt", weights_only=False)
        B = torch.load(tmp_path / "resumed.pt", weights_only=False)
        A, B = A["model"], B["model"]
    else:
        A = torch.load(tmp_path / "straight" / "ckpt_30.pt")["model"]
        B = torch.load(tmp_path / "resumed" / "ckpt_30.pt")["model"]
    assert len(A) >= 39 and A.keys() == B.keys()
    for k in A:
        assert float((A[k].float() - B[k].float()).abs().max()) == 0.0, k


def test_slice_partition_covers_every_range_once():
    """The reduce-scatter / shard partition (gradsync.slice_of == csrc/allreduce.cu slice_bounds): for
    every world size the W slices tile the range exactly, on 16-byte boundaries."""
    from pyrecover_b200.gradsync import slice_of
    for esize in (2, 4):
        for lo, hi in ((0, 1024 * 8), (4096, 4096 + 8 * 1000), (0, 8 * 7), (128, 128)):
            for world in range(1, 9):
                cur = lo
                for r in range(world):
                    a, b = slice_of(lo, hi, esize, r, world)
                    assert a == cur and b >= a and (a - lo) * esize % 16 == 0
                    cur = b
                assert cur == lo + (hi - lo) * esize // 16 * 16 // esize


def test_clip_grad_norm_flag_on_the_host_path(tmp_path):
    """--clip-grad-norm on a CPU model runs the reference's own clip_grad_norm_ (utils.py:84-89) at the
    call site the reference left commented out (train.py:271-272); a tight threshold changes the run."""
    def run(name, extra):
        cmd = [sys.executable, str(ROOT / "train.py"), "--device", "cpu", "--synthetic-data", "--model-preset", "tiny",
               "--sequence-length", "32", "--batch-size", "1", "--training-steps", "4", "--checkpoint-frequency", "4",
               "--checkpoint-dir", str(tmp_path), "--experiment_name", name, "--logging-frequency", "2",
               "--learning-rate", "1e-3", *extra]
        r = subprocess.run(cmd, cwd=str(ROOT), env=dict(os.environ, OMP_NUM_THREADS="2"), capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        return torch.load(tmp_path / name / "ckpt_4.pt", map_location="cpu")["model"]
    A = run("plain", ())
    B = run("clip", ("--clip-grad-norm", "--grad-max-norm", "0.01"))
    C = run("loose", ("--clip-grad-norm", "--grad-max-norm", "1e6"))
    assert any(not torch.equal(A[k], B[k]) for k in A)
    assert all(torch.equal(A[k], C[k]) for k in A)          # coefficient clamps to 1: nothing changes


# ------------------------------------------------------------------------- converter ------
def test_convert_reference_vanilla_to_dcp_and_back(tmp_path, golden_dir):
    """pyrecover.convert on the CPU: a checkpoint written by the REFERENCE's save_ckpt_vanilla becomes a
    torch.distributed.checkpoint directory with the keys the reference's save_ckpt_distributed
    produces, which then loads through load_ckpt_distributed (the reference's DCP call) into a fresh
    model / optimizer / scheduler with every tensor equal; and back to a vanilla file."""
    import pyrecover
    from pyrecover import convert
    from utils import build_lr_scheduler
    src = golden_dir / "ref_vanilla.pt"
    state = convert.load_state_cpu(str(src))
    dcp_dir = tmp_path / "ckpt_3"
    assert convert.main([str(src), str(dcp_dir), "--to", "dcp"]) == 0
    assert (dcp_dir / ".metadata").exists()
    import pickle
    keys = set(pickle.load(open(dcp_dir / ".metadata", "rb")).state_dict_metadata)
    assert {"model.0.weight", "model.1.bias", "optimizer.state.0.exp_avg", "optimizer.state.2.exp_avg_sq",
            "optimizer.param_groups.0.lr", "metadata.epoch", "metadata.step", "lr_scheduler.last_epoch"} <= keys
    model = torch.nn.Sequential(torch.nn.Linear(24, 16, bias=False), torch.nn.Linear(16, 8, bias=True)).to(torch.bfloat16)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=False)
    sched = build_lr_scheduler(opt, 10)
    # (no scheduler in the load template: DCP wants every template key in the checkpoint, and this repo's
    # warm-up lambda is an object with state where the reference's functools.partial has none)
    assert pyrecover.load_ckpt_distributed(model, opt, None, None, str(dcp_dir)) == (1, 3)
    for k, t in model.state_dict().items():
        assert torch.equal(t, state["model"][k]), k
    sd = opt.state_dict()["state"]
    for i, st in state["optimizer"]["state"].items():
        assert torch.equal(sd[i]["exp_avg"], st["exp_avg"]) and torch.equal(sd[i]["exp_avg_sq"], st["exp_avg_sq"])
        assert float(sd[i]["step"]) == float(st["step"])
    back = tmp_path / "back.pt"
    assert convert.main([str(src), str(back), "--to", "vanilla"]) == 0
    again = torch.load(back, map_location="cpu", weights_only=False)
    assert again["step"] == 3 and all(torch.equal(again["model"][k], state["model"][k]) for k in state["model"])
    with pytest.raises(RuntimeError, match="neither"):
        convert.load_state_cpu(str(tmp_path))


def test_convert_reads_a_sharded_directory_without_gpu(tmp_path):
    """The sharded format is self-describing (shard_<r>.json pieces + the layout table in meta.pt):
    a directory laid out by hand for world size 2, with the two ranks' runs interleaved, comes back
    as the reference's state dict; a missing COMPLETE marker or a truncated shard is an error."""
    from pyrecover import convert
    w = torch.arange(96, dtype=torch.float32).reshape(8, 12).bfloat16()
    b = torch.arange(8, dtype=torch.float32).bfloat16()
    m = (torch.arange(96, dtype=torch.float32) * 0.5).reshape(8, 12).bfloat16()
    v = (torch.arange(96, dtype=torch.float32) * 0.25).reshape(8, 12).bfloat16()
    # image: w @0 (192 B) | b @256 (16 B) | m @512 | v @768 ; total 1024
    image = np.zeros(1024, dtype=np.uint8)
    for off, t in ((0, w), (256, b), (512, m), (768, v)):
        raw = t.contiguous().view(torch.uint8).numpy().reshape(-1)
        image[off:off + raw.size] = raw
    d = tmp_path / "ckpt_9"
    d.mkdir()
    runs = {0: [(0, 96), (256, 16), (512, 96), (768, 96)], 1: [(96, 96), (608, 96), (864, 96)]}
    for r, pieces in runs.items():
        blob, table, loff = bytearray(), [], 0
        for goff, n in pieces:
            table.append([goff, n, loff])
            blob += image[goff:goff + n].tobytes() + b"\0" * ((-n) % 256)
            loff += n + ((-n) % 256)
        (d / f"shard_{r:05d}.bin").write_bytes(bytes(blob))
        (d / f"shard_{r:05d}.json").write_text(json.dumps({"format": 1, "rank": r, "world": 2, "sum64": [0, 0],
                                                            "image_bytes": loff, "full_image_bytes": 1024, "pieces": table}))
        (d / f"COMPLETE_{r:05d}").write_text("2")
    layout = {"model": {"lin.weight": ["bfloat16", 0, [8, 12]], "lin.bias": ["bfloat16", 256, [8]]},
              "optimizer": {0: {"exp_avg": ["bfloat16", 512, [8, 12]], "exp_avg_sq": ["bfloat16", 768, [8, 12]]}, 1: {}},
              "image_bytes": 1024}
    torch.save({"epoch": 2, "step": 9, "opt_steps": [9], "param_groups": [{"lr": 1e-3, "betas": (0.9, 0.999)}],
                "lr_scheduler": {"last_epoch": 9}, "layout": layout}, d / "meta.pt")
    st = convert.load_state_cpu(str(d))
    assert st["epoch"] == 2 and st["step"] == 9 and list(st["model"]) == ["lin.weight", "lin.bias"]
    assert torch.equal(st["model"]["lin.weight"], w) and torch.equal(st["model"]["lin.bias"], b)
    assert torch.equal(st["optimizer"]["state"][0]["exp_avg"], m) and torch.equal(st["optimizer"]["state"][0]["exp_avg_sq"], v)
    assert float(st["optimizer"]["state"][0]["step"]) == 9.0 and 1 not in st["optimizer"]["state"]
    assert st["optimizer"]["param_groups"][0]["params"] == [0, 1] and st["lr_scheduler"] == {"last_epoch": 9}
    out = tmp_path / "v.pt"
    assert convert.main([str(d), str(out), "--to", "vanilla"]) == 0
    assert torch.equal(torch.load(out, map_location="cpu", weights_only=False)["model"]["lin.weight"], w)
    (d / "shard_00001.bin").write_bytes((d / "shard_00001.bin").read_bytes()[:300])
    with pytest.raises(RuntimeError, match="truncated"):
        convert.load_state_cpu(str(d))
    (d / "COMPLETE_00001").unlink()
    with pytest.raises(RuntimeError, match="incomplete"):
        convert.load_state_cpu(str(d))


# ------------------------------------------------------------------------- round 2 --------
REFERENCE = Path("/root/reference")


def test_reference_check_weights_equality_runs_unmodified_on_our_files(tmp_path):
    """north_star's correctness criterion: the reference's OWN tests/check_weights_equality.py
    (:112-115 loads with a bare torch.load(path)["model"], :166-178 compares), executed unmodified as a
    subprocess on two checkpoints written by this package: exit 0 when equal, 1 when a weight differs."""
    script = REFERENCE / "tests" / "check_weights_equality.py"
    if not script.exists():
        pytest.skip("the reference checkout is not present on this machine")
    import pyrecover
    model = _small_model(3)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5)
    a, b, c = (tmp_path / n / "ckpt_1.pt" for n in "abc")
    for f in (a, b, c):
        f.parent.mkdir()
    pyrecover.save_ckpt_vanilla(model, opt, None, None, 1, 1, a, verify=False)
    pyrecover.save_ckpt_vanilla(model, opt, None, None, 1, 1, b, verify=False)
    with torch.no_grad():
        next(model.parameters()).add_(1.0)
    pyrecover.save_ckpt_vanilla(model, opt, None, None, 1, 1, c, verify=False)
    env = dict(os.environ, PYTHONPATH=str(ROOT), CUDA_VISIBLE_DEVICES="")     # `from pyrecover.checkpoint import ...` -> ours
    help_text = subprocess.run([sys.executable, str(script), "--help"], cwd=str(tmp_path), env=env, capture_output=True, text=True).stdout
    dev = ["--device", "cpu"] if "--device" in help_text else []
    same = subprocess.run([sys.executable, str(script), str(a), str(b), *dev], cwd=str(tmp_path), env=env, capture_output=True, text=True)
    assert same.returncode == 0, same.stdout[-2000:] + same.stderr[-2000:]
    diff = subprocess.run([sys.executable, str(script), str(a), str(c), *dev], cwd=str(tmp_path), env=env, capture_output=True, text=True)
    assert diff.returncode == 1, diff.stdout[-2000:] + diff.stderr[-2000:]


def _lay_native_dcp_dir(d, world, module_prefix=""):
    """What Snapshotter.save_sharded does, on the CPU: a (p | m | v) image cut into `world` flat
    16-byte-granular slices, every rank's boxes framed into __<r>_0.distcp, rank 0 writing .metadata."""
    from pyrecover_b200 import dcpfile as D
    from pyrecover_b200.gradsync import slice_of
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(37, 53), torch.nn.Linear(53, 11)).bfloat16()
    params = list(model.named_parameters())
    offs, cur = [], 0
    for _, p in params:
        offs.append(cur)
        cur += (p.numel() * 2 + 255) // 256 * 256
    arena = (cur + 256 * world - 1) // (256 * world) * (256 * world)
    image = np.zeros(3 * arena, np.uint8)
    m = {n: torch.randn_like(p) for n, p in params}
    v = {n: torch.rand_like(p) for n, p in params}
    layout = {"model": {}, "optimizer": {}, "image_bytes": 3 * arena}
    for i, ((n, p), o) in enumerate(zip(params, offs)):
        nb = p.numel() * 2
        for t, base in ((p.data, 0), (m[n], arena), (v[n], 2 * arena)):
            image[base + o: base + o + nb] = np.frombuffer(t.contiguous().view(torch.int16).numpy().tobytes(), np.uint8)
        layout["model"][n] = ["bfloat16", o, list(p.shape)]
        layout["optimizer"][i] = {"exp_avg": ["bfloat16", arena + o, list(p.shape)],
                                  "exp_avg_sq": ["bfloat16", 2 * arena + o, list(p.shape)]}
    table = D.tensor_table(layout, module_prefix)
    group = {"lr": 2e-3, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0.01, "amsgrad": False, "maximize": False,
             "foreach": None, "capturable": False, "differentiable": False, "fused": None, "decoupled_weight_decay": True,
             "initial_lr": 1e-3, "params": list(range(len(params)))}
    small_state = {"epoch": 1, "step": 7, "opt_steps": [7], "param_groups": [group],
                   "lr_scheduler": {"base_lrs": [1e-3], "last_epoch": 7, "_step_count": 8, "_is_initial": False,
                                    "_get_lr_called_within_step": False, "_last_lr": [2e-3], "lr_lambdas": [None]}}
    plans, placed0, flat, paths = {}, None, None, None
    for r in reversed(range(world)):
        runs = []
        for base in (0, arena, 2 * arena):
            a, b = slice_of(0, arena // 2, 2, r, world)
            runs.append((base + a * 2, (b - a) * 2))
        items, nbytes = D.plan_rank(table, runs)
        plans[r] = items
        img = np.zeros(nbytes, np.uint8)                       # the pack kernel's job
        for it in items:
            img[it.payload_off: it.payload_off + it.nbytes] = image[it.global_off: it.global_off + it.nbytes]
        small = None
        if r == 0:
            flat, paths = D.small_state_items(small_state, len(params))
            small = flat
        mv = memoryview(img)
        placed = D.write_rank_file(str(d / D.rank_file(r)), items, nbytes, [mv[:1000], mv[1000:]], small)
        if r == 0:
            placed0 = placed
    D.write_metadata(str(d), D.build_metadata(table, plans, flat, paths, placed0))
    return model, m, v, plans


@pytest.mark.parametrize("world", [1, 2, 3])
def test_native_dcp_directory_is_read_by_stock_dcp_and_by_the_reference_loader(tmp_path, world):
    """the sharded directory the B200 path writes IS a torch.distributed.checkpoint directory
    with the reference's keys (pyrecover/checkpoint.py:254-274).  Flat arena slices become at most
    2 * ndim - 1 boxes per tensor and rank; here: stock dist_cp.load restores model, moments, step scalars,
    param_groups, scheduler and metadata from it bit for bit, every item is a valid zip (CRCs included),
    and the reference's own load_ckpt_distributed (:300-368, loaded from /root/reference when present)
    returns (epoch, step) and fills a model + fresh AdamW."""
    import io
    import zipfile

    import torch.distributed.checkpoint as dist_cp
    from pyrecover_b200 import dcpfile as D
    d = tmp_path / "ckpt_7"
    d.mkdir()
    model, m, v, plans = _lay_native_dcp_dir(d, world)
    params = list(model.named_parameters())

    def fresh():
        m2 = torch.nn.Sequential(torch.nn.Linear(37, 53), torch.nn.Linear(53, 11)).bfloat16()
        o2 = torch.optim.AdamW(m2.parameters(), lr=5.0)
        s2 = torch.optim.lr_scheduler.LambdaLR(o2, lambda t: 1.0)
        return m2, o2, s2

    def check(m2, o2, s2):
        for (n, p), (_, q) in zip(params, m2.named_parameters()):
            assert torch.equal(p, q), n
            assert torch.equal(o2.state[q]["exp_avg"], m[n]) and torch.equal(o2.state[q]["exp_avg_sq"], v[n]), n
            assert float(o2.state[q]["step"]) == 7.0
        assert o2.param_groups[0]["lr"] == 2e-3 and s2.last_epoch == 7
    m2, o2, s2 = fresh()
    for p in m2.parameters():
        o2.state[p] = {"step": torch.zeros(()), "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p)}
    objs = {"model": m2, "optimizer": o2, "metadata": {"epoch": 0, "step": 0}, "lr_scheduler": s2}
    dist_cp.load(state_dict=objs, checkpoint_id=str(d), storage_reader=dist_cp.FileSystemReader(str(d)))
    check(m2, o2, s2)
    assert objs["metadata"] == {"epoch": 1, "step": 7}
    for r, items in plans.items():
        raw = (d / D.rank_file(r)).read_bytes()
        for it in items:
            assert zipfile.ZipFile(io.BytesIO(raw[it.file_off: it.file_off + it.archive.length])).testzip() is None
    if (REFERENCE / "pyrecover" / "checkpoint.py").exists():
        import importlib.util
        spec = importlib.util.spec_from_file_location("_ref_ckpt_for_dcp_test", REFERENCE / "pyrecover" / "checkpoint.py")
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        import pyrecover
        m3, o3, s3 = fresh()
        pyrecover.checkpoint._materialize_adam_state(o3)      # the reference drops moments into a fresh optimizer
        assert ref.load_ckpt_distributed(m3, o3, s3, None, str(d), verify=False) == (1, 7)
        check(m3, o3, s3)


def test_bench_parity_oracle_accepts_exact_results_and_rejects_a_flipped_bit():
    """bench.py's N > 1 parity verdict (the CPU half, the part that decides `parity.ok`): fed with results
    computed by the oracle itself it passes in all four kernel configurations; one flipped bit fails it."""
    import bench
    from oracle import oracle as O
    O.build()
    for world, mc, fused in ((2, False, True), (2, False, False), (8, True, True), (4, True, False)):
        n = 128 * world * 20
        gb = [O.to_bits((torch.randn(n, generator=torch.Generator().manual_seed(1234 + r)) * 1e-3).bfloat16()) for r in range(world)]
        g0 = torch.Generator().manual_seed(99)
        p0, m0 = (O.to_bits((torch.randn(n, generator=g0) * s).bfloat16()) for s in (0.02, 1e-3))
        v0 = O.to_bits((torch.randn(n, generator=g0) * 1e-3).square().bfloat16())
        scale = 1.0 / world
        got, sums = O.allreduce_bf16(gb, scale), O.allreduce_bf16(gb, 1.0)
        pb, mb, vb = p0.copy(), m0.copy(), v0.copy()
        if mc and fused:
            O.adamw_bf16(pb, sums, mb, vb, lr=3e-4, step=3, mode=0, gscale=scale)
        elif fused:
            O.adamw_bf16_gf32(pb, O.allreduce_bf16_to_f32(gb, 1.0), mb, vb, lr=3e-4, step=3, gscale=scale)
        else:
            O.adamw_bf16(pb, got, mb, vb, lr=3e-4, step=3, mode=0)
        r = bench.parity_oracle(world, n, mc, fused, gb, p0, m0, v0, 3e-4, 3, got, sums, pb, mb, vb)
        assert r["_ok"] and json.dumps(r)
        bad = pb.copy()
        bad[17] ^= 1
        assert not bench.parity_oracle(world, n, mc, fused, gb, p0, m0, v0, 3e-4, 3, got, sums, bad, mb, vb)["_ok"]
        far = got.copy()
        far[5] ^= 0x0040                                   # 64 ulp away: outside any bound
        assert not bench.parity_oracle(world, n, mc, fused, gb, p0, m0, v0, 3e-4, 3, far, sums, pb, mb, vb)["_ok"]


def test_policy_learns_from_durable_checkpoint_time_and_signal_stops(tmp_path):
    """Fake clock + real signal.  (1) The policy is fed the DURABLE time of an asynchronous write (the
    reference times its synchronous save, train.py:334-337), so a write slower than --default-ckpt-time raises
    the stop threshold; the milliseconds a save takes to enqueue never lower it.  (2) A decision that takes
    effect a fixed number of steps later widens the threshold by that many iterations.  (3) SIGUSR1 sets the
    resubmission state and runs the queue command once."""
    import signal

    from pyrecover.resubmit import setup_resubmission
    from pyrecover.timelimit import TimeAwarePolicy
    now = [0.0]
    pol = TimeAwarePolicy(1.0, 10.0, end_time=100.0, clock=lambda: now[0])
    pol.observe_iteration(1.0)
    assert pol.threshold == 1.0 + 10.0 + (5 * 1.0 + 10.0)
    assert not pol.observe_checkpoint(0.004)                 # enqueue time: the running maximum does not move
    assert pol.observe_checkpoint(12.5) and pol.max_ckpt_time == 12.5      # durable time of the previous save
    pol.observe_iteration(1.0)
    assert pol.threshold == 1.0 + 12.5 + (5 * 1.0 + 12.5)
    pol.extra_iterations = 2
    assert pol.threshold == 3 * 1.0 + 12.5 + (5 * 1.0 + 12.5)
    now[0] = 100.0 - pol.threshold + 0.1
    assert pol.should_stop()
    marker = tmp_path / "requeued"
    old = signal.getsignal(signal.SIGUSR1)
    try:
        st = setup_resubmission([sys.executable, "-c", f"open({str(marker)!r}, 'w').write('x')"], signals=(signal.SIGUSR1,))
        os.kill(os.getpid(), signal.SIGUSR1)
        os.kill(os.getpid(), signal.SIGUSR1)
        for _ in range(100):
            if marker.exists():
                break
            time.sleep(0.05)
        assert st.stop_requested and st.resubmitted and marker.exists()
    finally:
        signal.signal(signal.SIGUSR1, old)


def test_flat_ranges_become_contiguous_boxes():
    """dcpfile.flat_range_boxes: any flat element range of a row-major tensor is covered exactly once, in order,
    by at most 2 * ndim - 1 boxes, each contiguous in memory (what lets a reduce-scatter slice be written as
    torch.distributed.checkpoint chunks without moving a byte)."""
    import random

    from pyrecover_b200.dcpfile import flat_range_boxes
    rng = random.Random(7)
    for shape in [(7,), (5, 6), (3, 4, 5), (2, 3, 4, 5), (1, 9), (6, 1), ()]:
        n = int(np.prod(shape)) if shape else 1
        t = torch.arange(n).reshape(shape)
        for _ in range(150):
            lo = rng.randrange(0, n)
            hi = rng.randrange(lo, n + 1)
            boxes = flat_range_boxes(shape, lo, hi)
            assert len(boxes) <= max(1, 2 * len(shape) - 1)
            got = []
            for off, sz, first in boxes:
                sub = t[tuple(slice(o, o + s) for o, s in zip(off, sz))] if shape else t
                flat = sub.reshape(-1)
                assert torch.equal(flat, torch.arange(first, first + flat.numel()))      # contiguous run of the flat order
                got.append(flat)
            got = torch.cat(got) if got else torch.empty(0, dtype=torch.long)
            if shape:
                assert torch.equal(got, torch.arange(lo, hi))
    assert flat_range_boxes((4, 4), 5, 5) == []


def test_pack_grid_follows_image_size_and_step_period():
    """snapshot.pack_blocks_for: the pack that runs beside the next forward is spread over ~0.7 of a step
    (one CTA moves ~85 GB/s): 8 CTAs for the 6 GB image of the 1 B model at a 25 ms step, the 4-CTA floor for a
    0.75 GB shard, all SMs at most; an explicit PYRECOVER_B200_PACK_BLOCKS wins."""
    from pyrecover_b200 import snapshot as S
    assert S.pack_blocks_for(6_040_203_264, 0.0254) == 8
    assert S.pack_blocks_for(755_000_000, 0.028) == 4
    assert S.pack_blocks_for(48_000_000_000, 0.15) == 11
    assert S.pack_blocks_for(48_000_000_000, 0.001) == 148
    assert S.pack_blocks_for(6_040_203_264, None) == 10
    old = S.PACK_BLOCKS
    try:
        S.PACK_BLOCKS = 32
        assert S.pack_blocks_for(6_040_203_264, 0.0254) == 32
    finally:
        S.PACK_BLOCKS = old
