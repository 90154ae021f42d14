"""In-process launcher: ``spawn(fn, nprocs)`` starts one worker per rank with the same env
contract torchrun provides (WORLD_SIZE / RANK / LOCAL_RANK / MASTER_ADDR / MASTER_PORT), so the
reference-style entry point runs unchanged under either."""
from __future__ import annotations

import os
import socket

import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(local_rank: int, fn, nprocs: int, port: int, args: tuple) -> None:
    os.environ.update({"WORLD_SIZE": str(nprocs), "RANK": str(local_rank),
                       "LOCAL_RANK": str(local_rank), "LOCAL_WORLD_SIZE": str(nprocs),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    fn(*args)


def spawn(fn, nprocs: int, args: tuple = (), join: bool = True):
    """Run ``fn(*args)`` in ``nprocs`` processes; each sees the torchrun environment."""
    port = free_port()
    return mp.spawn(_worker, args=(fn, nprocs, port, args), nprocs=nprocs, join=join)
