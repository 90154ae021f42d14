"""Training / validation loops with the reference's observable behaviour (train_ddp.py:170-300):
per-GPU batches, rank-0 running loss/acc every ``print_freq`` steps, global throughput over the
window, epoch-end SUM all-reduce of (loss_sum, correct, total), validation on the full set on
every rank without sharding.

Differences that do not change the printed numbers: running metrics accumulate on the device
and are read only when a line is printed (the reference syncs the host twice per step,
train_ddp.py:217,220); the three epoch-end scalars travel in one packed all-reduce.
"""
from __future__ import annotations

import time

import torch
import torch.distributed as dist

from . import metrics as M
from .runtime import is_distributed
from .utils import nvtx


def reduce_tensor(tensor: torch.Tensor, op=None) -> torch.Tensor:
    """In-place all-reduce of a (scalar) tensor; identity when not distributed
    (reference train_ddp.py:159-167)."""
    if not (is_distributed() and dist.is_available() and dist.is_initialized()):
        return tensor
    dist.all_reduce(tensor, op=op if op is not None else dist.ReduceOp.SUM)
    return tensor


def _autocast(args, device):
    if getattr(args, "amp", False) and device.type == "cuda":
        dtype = torch.bfloat16 if getattr(args, "amp_dtype", "bf16") == "bf16" else torch.float16
        return torch.autocast("cuda", dtype=dtype)
    if getattr(args, "amp", False) and device.type == "cpu" and getattr(args, "amp_dtype", "") == "bf16":
        return torch.autocast("cpu", dtype=torch.bfloat16)
    import contextlib
    return contextlib.nullcontext()


def _device_batch(loader, images, device):
    """--gpu-augment: uint8 batches are cropped / flipped / normalised by one kernel on the GPU
    (reference: torchvision transforms per sample in 4 worker processes, train_ddp.py:91-101)."""
    tf = getattr(loader, "gpu_transform", None)
    if tf is None or images.dtype != torch.uint8:
        return images
    return tf(images.to(device, non_blocking=True))


def _epoch_reduce(loss_sum: float, correct: float, total: float, device, rank: int):
    packed = torch.tensor([loss_sum, correct, total], device=device, dtype=torch.float64)
    reduce_tensor(packed)
    if rank == 0:
        l, c, t = packed.tolist()
        return l / max(t, 1.0), 100.0 * c / max(t, 1.0)
    return None, None


def train_one_epoch(epoch, model, train_loader, optimizer, criterion, device, args, rank,
                    world_size, scaler=None, train_sampler=None, train_step=None, profiler=None):
    """Returns (global_loss, global_acc, epoch_time) — rank 0 gets numbers, others None."""
    model.train()
    if train_sampler is not None:
        train_sampler.set_epoch(epoch)
    use_scaler = getattr(args, "amp", False) and scaler is not None and scaler.is_enabled()
    acc = torch.zeros(3, device=device, dtype=torch.float64)  # loss*B, correct, total
    if train_step is not None:
        train_step.metrics.zero_()
    start_epoch = time.time()
    accum_time, accum_samples = 0.0, 0
    # On CUDA the steps are asynchronous (no per-step host sync), so the window's step time is
    # taken from CUDA events around each step: device time, loader wait excluded — the same
    # quantity the reference obtains from time.time() + .item() (train_ddp.py:196,217,224).
    cuda = device.type == "cuda"
    win_events: list = []
    max_steps = getattr(args, "steps", 0) or 0
    check_every = getattr(args, "check_sync_every", 0) or 0
    n_steps = len(train_loader) if not max_steps else min(max_steps, len(train_loader))
    for i, (images, targets) in enumerate(train_loader):
        if max_steps and i >= max_steps:
            break
        batch_start = time.time()
        bsz = images.size(0)
        images = _device_batch(train_loader, images, device)
        if cuda:
            ev_s = torch.cuda.Event(enable_timing=True)
            ev_e = torch.cuda.Event(enable_timing=True)
            ev_s.record()
        fast = train_step is not None and tuple(images.shape) == train_step.batch_shape
        if profiler is not None and profiler.should_sample():
            profiler.step_begin()
        else:
            if profiler is not None:
                profiler._step += 1
                profiler._ev0 = None
        if fast:
            with nvtx.range("train_step(graph)"):
                train_step(images, targets)
        else:
            with nvtx.range("h2d"):
                images = images.to(device, non_blocking=True)
                targets = targets.to(device, non_blocking=True)
            optimizer.zero_grad(set_to_none=True)
            with nvtx.range("forward"), _autocast(args, device):
                outputs = model(images)
                loss = criterion(outputs, targets)
            if use_scaler:
                with nvtx.range("backward+grad_sync"):
                    scaler.scale(loss).backward()
                with nvtx.range("optimizer"):
                    scaler.step(optimizer)
                    scaler.update()
            else:
                with nvtx.range("backward+grad_sync"):
                    loss.backward()
                with nvtx.range("optimizer"):
                    optimizer.step()
            with torch.no_grad():
                acc[0] += loss.detach().double() * bsz
                acc[1] += (outputs.detach().argmax(1) == targets).sum().double()
                acc[2] += bsz
        if profiler is not None and profiler._ev0 is not None:
            profiler.step_end()
        if check_every and (i + 1) % check_every == 0 and hasattr(model, "verify_replicas"):
            model.verify_replicas()
        if cuda:
            ev_e.record()
            win_events.append((ev_s, ev_e))
        printing = rank == 0 and (i + 1) % args.print_freq == 0
        if printing:
            tot = acc.clone()
            if train_step is not None:
                tot += train_step.metrics
            l, c, t = tot.tolist()  # the only host sync of the window
        accum_time += time.time() - batch_start
        accum_samples += bsz * world_size
        if printing:
            if cuda and win_events:
                accum_time = sum(a.elapsed_time(b) for a, b in win_events) * 1e-3
            throughput = accum_samples / accum_time if accum_time > 0 else 0.0
            print(M.step_line(epoch, i + 1, n_steps, l / max(t, 1.0), 100.0 * c / max(t, 1.0),
                              throughput))
            accum_time, accum_samples = 0.0, 0
            win_events = []
        elif cuda and len(win_events) > 4096:
            win_events = win_events[-4096:]
    tot = acc.clone()
    if train_step is not None:
        tot += train_step.metrics
    l, c, t = tot.tolist()
    global_loss, global_acc = _epoch_reduce(l, c, t, device, rank)
    epoch_time = time.time() - start_epoch
    ddp_comm = getattr(model, "comm", None)
    if ddp_comm is not None:
        ddp_comm.check_health()
    return global_loss, global_acc, epoch_time


@torch.no_grad()
def validate(model, val_loader, criterion, device, rank, world_size, args=None):
    """Full validation set on every rank, never autocast — FP32 storage, TF32 tensor-core
    convolutions — even under ``--amp``, exactly as the reference (train_ddp.py:266-300; cuDNN's
    allow_tf32 default).  ``--eval-amp`` opts into the 16-bit inference kernels instead."""
    model.eval()
    acc = torch.zeros(3, device=device, dtype=torch.float64)
    native_eval = args is not None and getattr(args, "amp", False) and device.type == "cuda" \
        and getattr(args, "eval_amp", False)
    import contextlib
    edt = torch.float16 if (args is not None and getattr(args, "amp_dtype", "bf16") == "fp16") else torch.bfloat16
    ctx = torch.autocast("cuda", dtype=edt) if native_eval else contextlib.nullcontext()
    max_steps = (getattr(args, "steps", 0) or 0) if args is not None else 0
    for i, (images, targets) in enumerate(val_loader):
        if max_steps and i >= max_steps:
            break
        images = _device_batch(val_loader, images, device).to(device, non_blocking=True)
        targets = targets.to(device, non_blocking=True)
        with ctx:
            outputs = model(images)
            loss = criterion(outputs, targets)
        bsz = images.size(0)
        acc[0] += loss.double() * bsz
        acc[1] += (outputs.argmax(1) == targets).sum().double()
        acc[2] += bsz
    l, c, t = acc.tolist()
    return _epoch_reduce(l, c, t, device, rank)
