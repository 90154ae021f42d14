"""Built-in communication hooks with torch's signature ``hook(state, bucket) -> Future``
(torch distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-135; C++ built-ins ALLREDUCE
and FP16_COMPRESS in c10d/default_comm_hooks.hpp:9-52).

They run on the torch.distributed reducer (``register_comm_hook`` switches to it): the native
fused peer-memory kernels are this framework's default "hook", these exist so code written
against torch DDP's hook API keeps working.  ``state`` is a process group or
None, exactly as in torch.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def _world(group) -> int:
    return dist.get_world_size(group) if group is not None else dist.get_world_size()


def _done(t: torch.Tensor) -> torch.futures.Future:
    fut: torch.futures.Future = torch.futures.Future()
    fut.set_result(t)
    return fut


def allreduce_hook(state, bucket) -> torch.futures.Future:
    """Average the bucket across ranks in its own dtype (fp32)."""
    buf = bucket.buffer()
    buf.div_(_world(state))
    dist.all_reduce(buf, group=state)
    return _done(buf)


def _compress_hook(state, bucket, dtype) -> torch.futures.Future:
    buf = bucket.buffer()
    low = buf.to(dtype).div_(_world(state))
    dist.all_reduce(low, group=state)
    buf.copy_(low)
    return _done(buf)


def fp16_compress_hook(state, bucket) -> torch.futures.Future:
    """Cast to fp16, average, cast back (halves the bytes on the wire)."""
    return _compress_hook(state, bucket, torch.float16)


def bf16_compress_hook(state, bucket) -> torch.futures.Future:
    """Cast to bf16, average, cast back."""
    return _compress_hook(state, bucket, torch.bfloat16)


def noop_hook(state, bucket) -> torch.futures.Future:
    """No communication at all — for measuring the compute-only step (torch debugging_hooks)."""
    return _done(bucket.buffer())
