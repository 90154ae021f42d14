"""Process bootstrap: launcher env contract, process-group init/teardown, device pick, seeding.

Behavioural parity with the reference (train_ddp.py:49-78, 318-329):
  * distributed iff ``WORLD_SIZE > 1`` (env contract of torchrun: WORLD_SIZE / RANK / LOCAL_RANK /
    MASTER_ADDR / MASTER_PORT);
  * ``setup_distributed()`` returns ``(rank, world_size, local_rank)`` and ``(0, 1, 0)`` when not
    distributed;
  * every rank seeds with ``seed + rank`` (train_ddp.py:76-78) — replicas become identical only
    through the wrapper's rank-0 broadcast, exactly as with torch DDP.

Differences: the backend is selectable.  ``native`` (default on CUDA) still
creates a torch process group — NCCL on GPUs, gloo on CPU — but only for bootstrap (exchange of
IPC handles through its store) and cold-path scalar collectives; the gradient data plane is the
peer-memory kernels in csrc/comm.
"""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist


def is_distributed() -> bool:
    """Reference train_ddp.py:49-50."""
    return int(os.environ.get("WORLD_SIZE", "1")) > 1


def pick_backend(backend: str | None = None) -> str:
    """Torch process-group backend used for bootstrap / cold-path collectives."""
    if backend in (None, "native", "auto"):
        return "nccl" if torch.cuda.is_available() else "gloo"
    if backend not in ("nccl", "gloo"):
        raise ValueError(f"unknown backend {backend!r} (native|nccl|gloo)")
    return backend


def setup_distributed(backend: str | None = None, timeout_s: float = 600.0):
    """Initialise the process group from the torchrun env contract (train_ddp.py:53-68)."""
    if not is_distributed():
        return 0, 1, 0
    world_size = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    pg_backend = pick_backend(backend)
    if torch.cuda.is_available():
        # Bind the device before the group exists so NCCL never touches cuda:0 from every rank.
        torch.cuda.set_device(local_rank)
    if not dist.is_initialized():
        kwargs = {}
        if pg_backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", local_rank)
        timeout = datetime.timedelta(seconds=timeout_s)
        store = _restart_safe_store(rank, world_size, timeout)
        if store is not None:
            dist.init_process_group(backend=pg_backend, store=store, rank=rank,
                                    world_size=world_size, timeout=timeout, **kwargs)
        else:
            dist.init_process_group(backend=pg_backend, init_method="env://", timeout=timeout,
                                    **kwargs)
    return rank, world_size, local_rank


def _restart_safe_store(rank: int, world_size: int, timeout):
    """Under ``torchrun --max-restarts`` with a static rendezvous (``--master-addr/--master-port``)
    every restarted worker group talks to the SAME agent-hosted TCPStore, and ``env://`` uses it
    without a per-attempt prefix: a restarted rank can read the dead incarnation's rendezvous keys
    (gloo: its peer's old listening address -> "connection refused" or a stalled connect; observed
    on this code base's restart test about every other run).  The reference inherits that
    behaviour (train_ddp.py:65); here the keys of every RESTARTED attempt live under their own
    prefix.  Returns None outside torchrun and for the first attempt (plain ``env://`` then)."""
    import os
    attempt = os.environ.get("TORCHELASTIC_RESTART_COUNT")
    addr, port = os.environ.get("MASTER_ADDR"), os.environ.get("MASTER_PORT")
    if attempt is None or addr is None or port is None or int(attempt) == 0:
        return None        # first incarnation: nothing stale in the store, plain env://
    agent_hosts_store = os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "False") == "True"
    base = dist.TCPStore(addr, int(port), world_size, is_master=(rank == 0 and not agent_hosts_store),
                         timeout=timeout, wait_for_workers=False)
    return dist.PrefixStore(f"dpt_b200/attempt_{attempt}", base)


def cleanup_distributed() -> None:
    """Reference train_ddp.py:71-73."""
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def set_seed(seed: int, rank: int) -> None:
    """Reference train_ddp.py:76-78 (note: seed + rank, replicas differ until the broadcast)."""
    torch.manual_seed(seed + rank)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed + rank)


def pick_device(local_rank: int) -> torch.device:
    """Reference train_ddp.py:321-324."""
    if torch.cuda.is_available():
        return torch.device(f"cuda:{local_rank}")
    return torch.device("cpu")


def world_info() -> tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1
