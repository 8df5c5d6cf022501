"""Training driver: `python train.py <flags>` -- the command submit-training-simple.sh runs.

Keeps the reference loop's observable behaviour (train.py:37-400): flags, log line, checkpoint
naming (`ckpt_<step>[.pt]`, `ckpt_<step>_final[.pt]`), time-aware stop, resume semantics, end-of-
run timing summary.  Underneath, the data-parallel hot path (gradient exchange, AdamW, checkpoint
snapshot) runs on the sm_100a kernels of pyrecover_b200 when the model lives on a GPU.
"""
import csv
import time
from pathlib import Path

import torch

from dist_utils import (get_rank, is_distributed_activated, is_rank0, log_rank0, maybe_cleanup_distributed,
                        maybe_init_distributed)
from pyrecover.checkpoint import (load_ckpt_distributed, load_ckpt_vanilla, save_ckpt_distributed,
                                  save_ckpt_vanilla, take_durable_checkpoint_time, wait_for_checkpoints)
from pyrecover.resubmit import setup_resubmission
from pyrecover.timelimit import TimeAwarePolicy
from pyrecover_b200.trainer import Trainer, model_args_from_cli
from pyrecover_b200.workload import SyntheticTokens
from utils import (PRECISION_STR_TO_DTYPE, get_args, get_num_flop_per_token, get_num_params, init_logger,
                   set_default_dtype)

H100_PEAK_FLOPS = 989e12      # the reference's MFU denominator (train.py:287), kept for log parity


def build_data(args, local_batch_size, rank):
    """(fetch(step) -> (input_ids, labels) on the host, vocab_size, sampler)."""
    if args.synthetic_data:
        vocab = args.vocab_size or 131072
        if args.model_preset and not args.vocab_size:
            from pyrecover_b200.workload import MODEL_PRESETS
            vocab = MODEL_PRESETS[args.model_preset][5]
        data = SyntheticTokens(vocab, args.sequence_length, local_batch_size, seed=args.seed, rank=rank,
                               pin=torch.cuda.is_available() and args.device != "cpu")
        return data.batch, vocab, None
    # real data: the reference's dataset.py / tokenizer stack, kept as-is by the user (out of scope here)
    try:
        from torch.utils.data import DataLoader
        from transformers import AutoTokenizer

        from dataset import CollatorForCLM, ParquetDataset  # type: ignore
    except ImportError as e:
        raise SystemExit(f"real-data training needs the reference's dataset.py next to train.py ({e}); "
                         "use --synthetic-data otherwise")
    tokenizer = AutoTokenizer.from_pretrained(args.tokenizer_name_or_path)
    ds = ParquetDataset(args.dataset, tokenizer, args.sequence_length, args.batch_size * args.training_steps)
    sampler = None
    world = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
    if world > 1:
        from torch.utils.data.distributed import DistributedSampler
        sampler = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True)
    dl = DataLoader(ds, batch_size=local_batch_size, collate_fn=CollatorForCLM(args.sequence_length, tokenizer.pad_token_id),
                    sampler=sampler, shuffle=sampler is None)
    state = {"it": iter(dl), "epoch": 1}

    def fetch(step):
        if sampler is not None:
            sampler.set_epoch(state["epoch"])
        try:
            return next(state["it"])
        except StopIteration:
            state["it"], state["epoch"] = iter(dl), state["epoch"] + 1
            return next(state["it"])
    return fetch, args.vocab_size or tokenizer.vocab_size, sampler


def train(args):
    run_start = time.perf_counter()
    ckpt_store_total = ckpt_load_total = 0.0
    use_cuda = torch.cuda.is_available() if args.device is None else args.device == "cuda"
    local_rank, world_size = maybe_init_distributed(args.distributed, backend=args.dist_backend, use_cuda=use_cuda)
    rank = get_rank()
    log_rank0(f"Experiment args: {args}")
    device = torch.device(f"cuda:{local_rank}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if args.snapshot_mode:
        import os
        os.environ["PYRECOVER_B200_SNAPSHOT"] = args.snapshot_mode

    global_batch_size = int(args.batch_size)
    local_batch_size = max(global_batch_size // world_size, 1)
    log_rank0(f"Global batch size: {global_batch_size}\nLocal batch size: {local_batch_size}")
    fetch, vocab_size, train_sampler = build_data(args, local_batch_size, rank)

    if args.deterministic:
        import os
        os.environ.setdefault("CUBLAS_WORKSPACE_CONFIG", ":4096:8")
        torch.use_deterministic_algorithms(True)
        torch.backends.cudnn.benchmark = False
        if use_cuda:      # cuDNN's attention backward accumulates with atomics; the flash backend has a deterministic backward
            torch.backends.cuda.enable_cudnn_sdp(False)
    log_rank0("Setting up Model...")
    torch.manual_seed(args.seed)             # identical initial weights on every rank and every run
    Transformer, model_config = model_args_from_cli(args, vocab_size)
    with set_default_dtype(PRECISION_STR_TO_DTYPE[args.model_dtype]), torch.device(device):
        model = Transformer(model_config)       # built in place on the device (the reference builds on the
                                                # host and copies, train.py:100-101: minutes at 8 B)
    if args.compile:
        log_rank0("Using `torch.compile`")
        model = torch.compile(model, fullgraph=True)
    trainer = Trainer(model, device, lr=args.learning_rate, fused=args.fused_optimizer,
                      warmup_steps=args.lr_warmup_steps, world_size=world_size, local_rank=local_rank,
                      grad_sync_mode=args.grad_sync,
                      max_grad_norm=args.grad_max_norm if args.clip_grad_norm else None)
    model, optimizer, lr_scheduler = trainer.model, trainer.optimizer, trainer.lr_scheduler
    num_flop_per_token = get_num_flop_per_token(get_num_params(model, exclude_embedding=True), model_config)

    ckpt_root = Path(args.checkpoint_dir)
    if ckpt_root.exists() and not ckpt_root.is_dir():
        exit(f"Checkpoint dir {ckpt_root} exists as file already! Abort!")
    exp_dir = ckpt_root / args.experiment_name
    exp_dir.mkdir(parents=True, exist_ok=True)
    csv_file = csv_writer = None
    if args.log_loss_to_csv and is_rank0():
        csv_file = open(exp_dir / f"{args.experiment_name}_loss_log.csv", "w", newline="")
        csv_writer = csv.writer(csv_file)
        csv_writer.writerow(["Step", "Loss"])
        csv_file.flush()

    sharded = args.use_torch_distributed_ckpt
    save_fn, load_fn = (save_ckpt_distributed, load_ckpt_distributed) if sharded else (save_ckpt_vanilla, load_ckpt_vanilla)
    log_rank0("Using sharded checkpoint directories" if sharded else "Using vanilla single-file checkpointing")

    policy = None
    if args.timeaware_checkpointing:
        policy = TimeAwarePolicy(args.default_iter_time, args.default_ckpt_time)
        log_rank0(f"Initial max_iter_time: {policy.max_iter_time}, max_ckpt_time: {policy.max_ckpt_time}, "
                  f"buffer_time: {policy.buffer_time}")
        if policy.end_time is None:
            log_rank0("Warning: SLURM_JOB_END_TIME is not set. Time-check logic will be skipped.")
        log_rank0(f"SLURM_JOB_END_TIME: {policy.end_time}")

    # SLURM's pre-termination signal (#SBATCH --signal=USR1@<s>) or a plain SIGTERM: stop like the time-aware
    # policy does (final checkpoint, clean exit) and, when a command is given, queue the continuation job
    resub = None
    if args.resubmit_command is not None:
        import shlex
        resub = setup_resubmission(shlex.split(args.resubmit_command) or None)
        log_rank0(f"Resubmission on SIGUSR1/SIGTERM armed: {args.resubmit_command or '(stop only)'}")
    stop_enabled = policy is not None or resub is not None
    grad_sync = trainer.grad_sync            # K1 communicator (CUDA, world > 1): carries the stop decision
    if policy is not None and grad_sync is not None:
        policy.extra_iterations = grad_sync.STOP_LAG      # the decision takes effect that many steps later

    def ckpt_path(step, final=False):
        name = f"ckpt_{step}" + ("_final" if final else "") + ("" if sharded else ".pt")
        return exp_dir / name

    def store(step, epoch, final=False):
        t0 = time.perf_counter()
        save_fn(model, optimizer, lr_scheduler, train_sampler, step, epoch, ckpt_path(step, final),
                max_keep=args.max_kept_checkpoints, verify=args.verify_checkpoints,
                is_distributed=(world_size > 1), rank=rank)
        if args.sync_checkpoints:
            wait_for_checkpoints()
        return time.perf_counter() - t0

    train_step, epoch = 0, 1
    if args.resume_from_checkpoint is not None:
        log_rank0(f"Try resume from checkpoint {args.resume_from_checkpoint}")
        t0 = time.perf_counter()
        epoch, train_step = load_fn(model, optimizer, lr_scheduler, train_sampler, args.resume_from_checkpoint,
                                    experiment_dir=exp_dir, verify=args.verify_checkpoints,
                                    is_distributed=(world_size > 1), rank=rank)
        ckpt_load_total += time.perf_counter() - t0
        log_rank0(f"Checkpoint loading completed in {ckpt_load_total:.2f} seconds")
    if is_distributed_activated() and world_size > 1:
        torch.distributed.barrier()

    # test aid: fail the run on ANY host<->device synchronisation inside a steady-state step
    sync_debug = use_cuda and __import__("os").environ.get("PYRECOVER_B200_SYNC_DEBUG") == "1"
    should_stop = stop_posted = False
    stop_flag = torch.zeros(1, dtype=torch.int32, device=device) if world_size > 1 and grad_sync is None else None
    ntokens = ntraining_tokens = 0
    time_last_log = time.perf_counter()
    log_rank0("Starting training!")
    while train_step < args.training_steps:
        train_step += 1
        if policy is not None:
            durable = take_durable_checkpoint_time()       # an asynchronous write finished since the last look
            if durable is not None and policy.observe_checkpoint(durable):
                log_rank0(f"Updated max_ckpt_time: {policy.max_ckpt_time} (durable time of the previous checkpoint)")
        if policy is not None and is_rank0() and not should_stop and policy.should_stop():
            should_stop = True
            log_rank0(f"[TIME CHECK] Remaining time ({policy.time_left():.2f}s) < threshold "
                      f"({policy.threshold:.2f}s). should_stop set to True.")
        if resub is not None and is_rank0() and resub.stop_requested and not should_stop:
            should_stop = True
            log_rank0("[SIGNAL] termination signal received. should_stop set to True.")
        if grad_sync is not None and stop_enabled and should_stop and not stop_posted and is_rank0():
            at = grad_sync.request_stop()          # rides in this step's gradient exchange: no collective, no sync
            stop_posted = True
            log_rank0(f"[STOP] every rank stops after optimizer step {at} of this run")
        iter_start = time.perf_counter()
        if args.profile and args.profile_step_start == train_step and use_cuda:
            torch.cuda.cudart().cudaProfilerStart()
            torch.autograd.profiler.emit_nvtx(record_shapes=True).__enter__()

        logging_step = train_step == 1 or train_step % args.logging_frequency == 0
        strict = sync_debug and train_step > 3 and not logging_step and csv_writer is None
        if strict:
            torch.cuda.set_sync_debug_mode("error")
        input_ids, labels = fetch(train_step)
        ntokens += global_batch_size * args.sequence_length
        ntraining_tokens += int(labels.ne(-100).sum()) * world_size
        loss = trainer.step_from_host(input_ids, labels)

        if csv_writer is not None:
            csv_writer.writerow([train_step, loss.item()])
            csv_file.flush()
        if logging_step:
            dt = time.perf_counter() - time_last_log
            tps = ntokens / dt
            if hasattr(optimizer, "check_finite"):
                optimizer.check_finite()           # non-finite gradient norm (clipping on): raise like the reference
            log_rank0(f"Epoch: {epoch} | Step: {train_step} | Loss: {loss.item():.2f} | Tokens per second: {tps:.2f} | "
                      f"Training tokens per second (%): {100 * (ntraining_tokens / dt) / tps:.2f} | "
                      f"MFU (%): {100 * num_flop_per_token * tps / H100_PEAK_FLOPS:.2f} | "
                      f"TFLOPs: {num_flop_per_token * tps / 1e12:.2f}")
            ntokens = ntraining_tokens = 0
            time_last_log = time.perf_counter()
        if policy is not None:
            if policy.observe_iteration(time.perf_counter() - iter_start):
                log_rank0(f"Updated max_iter_time: {policy.max_iter_time}")
            if train_step % args.logging_frequency == 0:
                log_rank0(f"Current buffer_time: {policy.buffer_time}")

        stop_now = should_stop
        if world_size > 1 and stop_enabled:
            if grad_sync is not None:
                stop_now = grad_sync.stop_due()        # rank 0's word, forwarded by K1's begin barrier
            else:
                # host path (CPU / gloo): rank 0 decides and broadcasts, as the reference (train.py:342-346)
                stop_flag.fill_(int(should_stop))
                torch.distributed.broadcast(stop_flag, src=0)
                stop_now = should_stop = bool(stop_flag.item())
        if strict:
            torch.cuda.set_sync_debug_mode("default")
        saved_this_step = False
        if args.checkpoint_frequency != -1 and train_step % args.checkpoint_frequency == 0:
            log_rank0(f"Saving checkpoint to {ckpt_path(train_step)}")
            dt = store(train_step, epoch)
            ckpt_store_total += dt
            saved_this_step = True
            if policy is not None and policy.observe_checkpoint(dt):
                log_rank0(f"Updated max_ckpt_time: {policy.max_ckpt_time}")
            log_rank0(f"Checkpoint store completed in {dt:.2f} seconds")

        if stop_enabled and stop_now:
            log_rank0(f"[TIME CHECK] Saving final checkpoint to {ckpt_path(train_step, final=True)} before exit.")
            t0 = time.perf_counter()
            store(train_step, epoch, final=True)
            wait_for_checkpoints()
            dt = time.perf_counter() - t0
            ckpt_store_total += dt
            log_rank0(f"[TIME CHECK] Final checkpoint store completed in {dt:.2f} seconds")
            break
        if args.profile and args.profile_step_end == train_step and use_cuda:
            torch.cuda.cudart().cudaProfilerStop()

    wait_for_checkpoints()
    if csv_file is not None:
        csv_file.close()
    log_rank0(f"Training completed in {time.perf_counter() - run_start:.2f} seconds")
    log_rank0(f"Total checkpoint loading time: {ckpt_load_total:.2f} seconds")
    log_rank0(f"Total checkpoint storing time: {ckpt_store_total:.2f} seconds")
    log_rank0(f"Total checkpointing time: {(ckpt_load_total + ckpt_store_total):.2f} seconds")
    trainer.close()
    maybe_cleanup_distributed()


if __name__ == "__main__":
    init_logger()
    train(get_args())
