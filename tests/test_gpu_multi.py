"""Multi-GPU tests (skipped on a single-GPU box): the data-parallel path through train.py with one
process per GPU -- GradSync (K1, replicated and fused K1+K2), checkpoint formats, resume."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _need(n):
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _launch(tmp, name, steps, world, extra=(), env_extra=None):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", PR_COMM_TIMEOUT_S="30", **(env_extra or {}))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29800 + (os.getpid() % 100)),
           str(ROOT / "train.py"), "--distributed", "--synthetic-data", "--model-preset", "tiny", "--sequence-length", "256",
           "--batch-size", str(world), "--training-steps", str(steps), "--checkpoint-frequency", "10",
           "--checkpoint-dir", str(tmp), "--experiment_name", name, "--fused-optimizer", "--verify-checkpoints",
           "--logging-frequency", "10", "--learning-rate", "1e-3", *extra]
    r = subprocess.run(cmd, cwd=str(ROOT), env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    return r


@pytest.mark.parametrize("mode", ["replicated", "fused", "overlap"])
@pytest.mark.parametrize("sharded", [False, True])
def test_two_gpu_train_resume_weight_equality(tmp_path, mode, sharded):
    """W = 2 over NVLink: 20 steps straight vs 10 + resume to 20 -> final weights equal with
    tolerance 0, for both gradient-exchange modes and both checkpoint formats; and the
    interrupted run's step-10 checkpoint equals the straight run's."""
    _need(2)
    extra = ("--grad-sync", mode) + (("--use-torch-distributed-ckpt",) if sharded else ())
    _launch(tmp_path, "straight", 20, 2, extra)
    _launch(tmp_path, "resumed", 10, 2, extra)
    _launch(tmp_path, "resumed", 20, 2, extra + ("--resume-from-checkpoint", "latest"))
    if not sharded:
        A = torch.load(tmp_path / "straight" / "ckpt_20.pt", map_location="cpu")
        B = torch.load(tmp_path / "resumed" / "ckpt_20.pt", map_location="cpu")
        for k in A["model"]:
            assert float((A["model"][k].float() - B["model"][k].float()).abs().max()) == 0.0, k
        for i in A["optimizer"]["state"]:
            assert torch.equal(A["optimizer"]["state"][i]["exp_avg"], B["optimizer"]["state"][i]["exp_avg"])
            assert torch.equal(A["optimizer"]["state"][i]["exp_avg_sq"], B["optimizer"]["state"][i]["exp_avg_sq"])
    else:
        for r in range(2):
            a = (tmp_path / "straight" / "ckpt_20" / f"__{r}_0.distcp").read_bytes()
            b = (tmp_path / "resumed" / "ckpt_20" / f"__{r}_0.distcp").read_bytes()
            assert a == b, f"shard {r} differs"


@pytest.mark.parametrize("mode", ["replicated", "fused"])
def test_two_gpu_multicast_train_resume_weight_equality(tmp_path, mode):
    """The same experiment with the NVSwitch multicast kernels forced on at W = 2
    (PR_COMM_MULTICAST=1; a box without multicast support runs the P2P kernels instead and the
    test still holds): the in-switch sum is deterministic, so straight and resumed runs end with
    identical weights; and the multicast run stays within bf16 rounding of the P2P run."""
    _need(2)
    extra = ("--grad-sync", mode)
    mc = {"PR_COMM_MULTICAST": "1"}
    _launch(tmp_path, "straight", 20, 2, extra, mc)
    _launch(tmp_path, "resumed", 10, 2, extra, mc)
    _launch(tmp_path, "resumed", 20, 2, extra + ("--resume-from-checkpoint", "latest"), mc)
    _launch(tmp_path, "p2p", 20, 2, extra, {"PR_COMM_MULTICAST": "0"})
    A = torch.load(tmp_path / "straight" / "ckpt_20.pt", map_location="cpu")
    B = torch.load(tmp_path / "resumed" / "ckpt_20.pt", map_location="cpu")
    C = torch.load(tmp_path / "p2p" / "ckpt_20.pt", map_location="cpu")
    for k in A["model"]:
        assert torch.equal(A["model"][k], B["model"][k]), k
        # 20 AdamW steps at lr 1e-3 move a weight by at most ~2e-2; the two exchanges differ by one
        # bf16 ulp of a gradient per step, far below that
        assert float((A["model"][k].float() - C["model"][k].float()).abs().max()) < 5e-2, k
    for i in A["optimizer"]["state"]:
        assert torch.equal(A["optimizer"]["state"][i]["exp_avg"], B["optimizer"]["state"][i]["exp_avg"])
        assert torch.equal(A["optimizer"]["state"][i]["exp_avg_sq"], B["optimizer"]["state"][i]["exp_avg_sq"])


@pytest.mark.parametrize("multicast", ["0", "1"])
def test_two_gpu_allreduce_kernels_parity(tmp_path, multicast):
    """K1 over real NVLink at W = 2 (scripts/allreduce_bench.py, small sizes): P2P two-shot / one-shot
    bit-exact against the fp32 rank-order sum on every rank, the fused kernel bit-equal to
    allreduce -> K2; with the NVSwitch multicast kernels forced on: within 1 bf16 ulp + the
    summation-order term, identical across ranks and from run to run, fp32 payload within its bound,
    fused multicast bit-equal to multicast allreduce -> K2."""
    _need(2)
    import json
    out = tmp_path / "ar.json"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", PR_COMM_TIMEOUT_S="30", PR_COMM_MULTICAST=multicast)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29900 + (os.getpid() % 90)), str(ROOT / "scripts" / "allreduce_bench.py"),
           "--min-log2", "21", "--max-log2", "23", "--mc-blocks", "16", "--json", str(out)]
    r = subprocess.run(cmd, cwd=str(ROOT), env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    res = json.loads(out.read_text())
    assert res["parity_two_shot"] and res["parity_one_shot"] and res["parity_fused"]
    if multicast == "1" and res["multicast_available"]:
        assert res["parity_multimem_within_bound"] and res["multimem_identical_across_ranks"]
        assert res["multimem_deterministic"] and res["parity_multimem_f32"] and res["parity_fused_multimem"]
        assert max(abs(int(k)) for k, v in res["multimem_ulp_diff_hist"].items() if v) <= 1
    if multicast == "0":
        assert not res["multicast_available"]


def test_two_gpu_gradient_clipping_modes_agree(tmp_path):
    """--clip-grad-norm at W = 2: the global norm is taken from the exchanged mean gradient, which
    is bit-identical on both ranks, so "replicated" (allreduce -> norm -> K2 everywhere) and
    "fused" (allreduce -> norm -> K2 on the own slice -> parameter all-gather) end with the same
    weights bit for bit; and a tight threshold changes the result (the coefficient is applied)."""
    _need(2)
    clip = ("--clip-grad-norm", "--grad-max-norm", "0.05")
    p2p = {"PR_COMM_MULTICAST": "0"}
    _launch(tmp_path, "rep", 10, 2, ("--grad-sync", "replicated") + clip, p2p)
    _launch(tmp_path, "fus", 10, 2, ("--grad-sync", "fused") + clip, p2p)
    _launch(tmp_path, "noclip", 10, 2, ("--grad-sync", "replicated"), p2p)
    A = torch.load(tmp_path / "rep" / "ckpt_10.pt", map_location="cpu")
    B = torch.load(tmp_path / "fus" / "ckpt_10.pt", map_location="cpu")
    C = torch.load(tmp_path / "noclip" / "ckpt_10.pt", map_location="cpu")
    differs = False
    for k in A["model"]:
        assert torch.equal(A["model"][k], B["model"][k]), k
        differs = differs or not torch.equal(A["model"][k], C["model"][k])
    assert differs, "clipping at 0.05 left the weights unchanged"
    for i in A["optimizer"]["state"]:
        assert torch.equal(A["optimizer"]["state"][i]["exp_avg"], B["optimizer"]["state"][i]["exp_avg"])


def test_two_gpu_matches_vanilla_and_sharded(tmp_path):
    """The same W = 2 run checkpointed in both formats holds the same bytes: the sharded
    directory, loaded on ONE GPU (different world size), equals the vanilla file."""
    _need(2)
    _launch(tmp_path, "v", 10, 2)
    _launch(tmp_path, "s", 10, 2, ("--use-torch-distributed-ckpt",))
    sys.path.insert(0, str(ROOT))
    import pyrecover
    from pyrecover_b200.optim import B200AdamW
    from pyrecover_b200.workload import Transformer, preset_args
    from utils import build_lr_scheduler
    torch.set_default_dtype(torch.bfloat16)
    m = Transformer(preset_args("tiny", 256)).cuda()
    torch.set_default_dtype(torch.float32)
    o = B200AdamW(m.parameters(), lr=1e-3, fused=True)
    s = build_lr_scheduler(o, 10)
    assert pyrecover.load_ckpt_distributed(m, o, s, None, str(tmp_path / "s" / "ckpt_10")) == (1, 10)
    V = torch.load(tmp_path / "v" / "ckpt_10.pt", map_location="cuda")
    for k, t in m.state_dict().items():
        assert torch.equal(t, V["model"][k]), k
    for i, p in enumerate(m.parameters()):
        assert torch.equal(o.state[p]["exp_avg_sq"], V["optimizer"]["state"][i]["exp_avg_sq"])


def test_two_gpu_overlap_many_buckets_resume_weight_equality(tmp_path):
    """"overlap" with buckets small enough that the tiny model is exchanged in many launches under backward
    (one learned trigger hook per bucket): straight vs resumed runs still end bit-identical, and equal to the
    single-launch "fused" run (same slices of the same sums, only cut per bucket)."""
    _need(2)
    env = {"PYRECOVER_B200_BUCKET_BYTES": str(1 << 20), "PR_COMM_MULTICAST": "0"}
    extra = ("--grad-sync", "overlap")
    _launch(tmp_path, "straight", 20, 2, extra, env)
    _launch(tmp_path, "resumed", 10, 2, extra, env)
    _launch(tmp_path, "resumed", 20, 2, extra + ("--resume-from-checkpoint", "latest"), env)
    _launch(tmp_path, "fused", 20, 2, ("--grad-sync", "fused"), {"PR_COMM_MULTICAST": "0"})
    A = torch.load(tmp_path / "straight" / "ckpt_20.pt", map_location="cpu")
    B = torch.load(tmp_path / "resumed" / "ckpt_20.pt", map_location="cpu")
    C = torch.load(tmp_path / "fused" / "ckpt_20.pt", map_location="cpu")
    for k in A["model"]:
        assert torch.equal(A["model"][k], B["model"][k]), k
        assert torch.equal(A["model"][k], C["model"][k]), k
    for i in A["optimizer"]["state"]:
        assert torch.equal(A["optimizer"]["state"][i]["exp_avg"], B["optimizer"]["state"][i]["exp_avg"])
        assert torch.equal(A["optimizer"]["state"][i]["exp_avg_sq"], C["optimizer"]["state"][i]["exp_avg_sq"])


@pytest.mark.parametrize("mode", ["fused", "overlap"])
def test_two_gpu_timeaware_stop_rides_in_the_exchange(tmp_path, mode):
    """Time-aware stop at W = 2 (reference train.py:163-190, 223-232, 342-375) with the decision carried by
    K1's begin barrier instead of a per-step broadcast + .item(): with a simulated SLURM_JOB_END_TIME both
    ranks stop after the SAME step, write ckpt_<step>_final as a complete sharded directory, and a resumed run
    continues from it.  PYRECOVER_B200_SYNC_DEBUG=1 makes train.py run the steady-state steps under
    torch.cuda.set_sync_debug_mode("error"): any host<->device synchronisation in the loop fails the run."""
    _need(2)
    import re
    import time
    env = {"SLURM_JOB_END_TIME": str(int(time.time()) + 45), "PYRECOVER_B200_SYNC_DEBUG": "1"}
    extra = ("--grad-sync", mode, "--timeaware-checkpointing", "--use-torch-distributed-ckpt", "--default-iter-time", "0.5",
             "--default-ckpt-time", "2", "--checkpoint-frequency", "-1", "--logging-frequency", "1000000")
    r = _launch(tmp_path, "ta", 100000, 2, extra, env)
    out = r.stdout + r.stderr
    finals = sorted((tmp_path / "ta").glob("ckpt_*_final"))
    assert len(finals) == 1, out[-3000:]
    step = int(re.search(r"ckpt_(\d+)_final", finals[0].name).group(1))
    assert step > 10 and "[STOP] every rank stops after optimizer step" in out
    assert (finals[0] / ".metadata").exists() and (finals[0] / "__0_0.distcp").exists() and (finals[0] / "__1_0.distcp").exists()
    r2 = _launch(tmp_path, "ta", step + 5, 2, ("--grad-sync", mode, "--use-torch-distributed-ckpt", "--resume-from-checkpoint",
                                               "latest", "--checkpoint-frequency", "-1"))
    assert f"step {step})" in (r2.stdout + r2.stderr)
