"""Process-group bootstrap and rank-aware logging.

Same public helpers as the reference's dist_utils.py (:14-101).  Differences: the rank / world /
local rank come from SLURM_* (the reference's only source, :45-47) OR from the torchrun
variables RANK / WORLD_SIZE / LOCAL_RANK; the backend is NCCL on CUDA and gloo on CPU (the
reference hard-wires NCCL, :50-54).  torch.distributed is bootstrap only here: rendezvous,
barriers, the initial parameter broadcast and the exchange of CUDA-IPC handles.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from utils import logger


def _env_world() -> Tuple[Optional[int], int, int]:
    """(rank, world, local_rank) from SLURM or torchrun variables; rank None = not distributed."""
    if "SLURM_PROCID" in os.environ and int(os.environ.get("SLURM_NTASKS", "1")) > 1:
        return (int(os.environ["SLURM_PROCID"]), int(os.environ["SLURM_NTASKS"]),
                int(os.environ.get("SLURM_LOCALID", 0)))
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    return None, 1, 0


def is_distributed_slurm_env() -> bool:
    return _env_world()[0] is not None


def is_distributed_activated() -> bool:
    return "DISTRIBUTED_RUN" in os.environ


def get_rank() -> int:
    if is_distributed_slurm_env() and is_distributed_activated():
        import torch.distributed as dist
        if dist.is_initialized():
            return dist.get_rank()
    return 0


def is_rank_eq(rank: int) -> bool:
    return get_rank() == rank


def is_rank0() -> bool:
    return get_rank() == 0


def maybe_init_distributed(activate_distributed: bool, backend: Optional[str] = None,
                           use_cuda: Optional[bool] = None) -> Tuple[int, int]:
    """Returns (local_rank, world_size)."""
    if not activate_distributed:
        print("--NOT RUNNING DISTRIBUTED TRAINING!--")
        return 0, 1
    os.environ["DISTRIBUTED_RUN"] = "1"
    rank, world, local_rank = _env_world()
    if rank is None:
        exit("Try running distributed training but environment is not setup for this!")
    import torch.distributed as dist
    use_cuda = torch.cuda.is_available() if use_cuda is None else use_cuda
    backend = backend or ("nccl" if use_cuda else "gloo")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if use_cuda:
        torch.cuda.set_device(local_rank)
    dist.init_process_group(backend=backend, rank=rank, world_size=world,
                            **({"device_id": torch.device("cuda", local_rank)} if use_cuda and backend == "nccl" else {}))
    print(f"[Rank {rank}] world_size={world}, local_rank={local_rank}, backend={backend}, pid={os.getpid()}", flush=True)
    log_rank0(f"Process group initialised: world_size={world}, backend={backend}")
    return local_rank, world


def maybe_cleanup_distributed():
    try:
        from pyrecover.checkpoint import release_checkpoint_engines, wait_for_checkpoints
        wait_for_checkpoints()
        release_checkpoint_engines()
    except ImportError:
        pass
    if is_distributed_activated() and is_distributed_slurm_env():
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
            print("Process group cleaned up!")


def log_rank(msg, rank):
    if not is_distributed_activated() or is_rank_eq(rank):
        logger.info(msg)


def log_rank0(msg):
    log_rank(msg, 0)


def get_slurm_job_end_time_env() -> Optional[float]:
    """SLURM_JOB_END_TIME as a UNIX timestamp, or None (unset / not a number)."""
    from pyrecover.timelimit import job_end_time
    return job_end_time()
