"""Headline benchmark: training throughput (samples/s) of the reference's default workload —
ResNet-18(num_classes=10) on CIFAR-shaped 128x3x32x32 batches per GPU, AMP, SGD momentum —
at N GPUs of one node (BASELINE.json config 2).

    python bench.py [--gpus N --steps K --warmup W]                 # this framework
    python bench.py --impl reference [--gpus N --steps K --warmup W] # unmodified reference

For N > 1 launch with torchrun (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the env).
Rank 0 prints ONE JSON line.  Timing: W >= 3 warm-up steps, then exactly K steps bracketed by
barrier + cuda synchronize, device-timed with CUDA events, max over ranks.  One step touches
~207 MB (fp32 params + grads + momentum, 16-bit shadow, activations) > 126 MB of L2, so no
explicit flush by default (`--l2-flush` adds an untimed 256 MiB memset between steps); SM clocks /
throttle reasons are sampled through NVML during the timed region.  `e2e` is the same metric
through the public API (TrainStep.__call__) with every step's inputs copied from pinned host
memory and the step's loss read back.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_VALUE = None  # the reference publishes no number


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--arch", default="resnet18")
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--image-size", type=int, default=32)
    p.add_argument("--num-classes", type=int, default=10)
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--amp-dtype", default="bf16", choices=["bf16", "fp16", "fp32"],
                   help="16-bit compute dtype of this framework's arm (fp16 adds the dynamic GradScaler, "
                        "i.e. exactly the reference's --amp numerics)")
    p.add_argument("--l2-flush", action="store_true",
                   help="flush L2 (untimed 256 MiB memset) between timed steps; default is no flush "
                        "because one step touches ~207 MB > 126 MB L2")
    p.add_argument("--fused-optimizer", action="store_true", default=None,
                   help="apply SGD inside the all-reduce kernel (default: on when N > 1)")
    p.add_argument("--no-fused-optimizer", dest="fused_optimizer", action="store_false")
    p.add_argument("--bucket-cap-mb", type=float, default=None)
    p.add_argument("--allreduce-algo", default="auto",
                   choices=["auto", "oneshot", "twoshot", "nvls", "twoshot_mc", "autotune"])
    p.add_argument("--comm-ctas", type=int, default=16,
                   help="CTAs of each bucket all-reduce kernel (fewer = less SM contention with backward)")
    p.add_argument("--eager-buffer-sync", action="store_true",
                   help="broadcast BN buffers from rank 0 every step (torch DDP behaviour) instead "
                        "of lazily before eval/checkpoint")
    p.add_argument("--no-extras", action="store_true",
                   help="skip the cold-path evidence after the timed regions (multi-GPU reduction "
                        "checks, fp16 + GradScaler arm, loss parity vs torchvision + torch DDP)")
    p.add_argument("--extras-timeout", type=float, default=150.0,
                   help="seconds the cold-path extras may take before the bench line is printed without them")
    p.add_argument("--tail-kb", type=float, default=None, help="tail bucket size (KiB; 0 = off)")
    p.add_argument("--oneshot-max-kb", type=float, default=None,
                   help="buckets up to this size use the one-shot push kernel")
    p.add_argument("--profile-sync", action="store_true",
                   help="extra eager pass with per-bucket event timing (reported under grad_sync)")
    return p.parse_args()


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int, period_s: float = 0.05) -> None:
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def env_world():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def max_over_ranks(value: float, device) -> float:
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        t = torch.tensor([value], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return value


def barrier_sync():
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize()


def make_host_batches(n, shape, num_classes, seed):
    import torch
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        x = torch.randn(shape, generator=g).pin_memory()
        t = torch.randint(0, num_classes, (shape[0],), generator=g).pin_memory()
        out.append((x, t))
    return out


# ====================================================================================== ours
def run_ours(a):
    import torch
    import dpt_b200  # noqa: F401
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.ops import _lib
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel
    from dpt_b200.runtime import cleanup_distributed, set_seed, setup_distributed

    rank, world, local = setup_distributed("native")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    set_seed(42, rank)
    model = build_model(device, a.arch, a.num_classes)
    ddp = None
    if world > 1:
        model = ddp = DistributedDataParallel(model, device_ids=[local], backend="native",
                                              bucket_cap_mb=a.bucket_cap_mb, comm_ctas=a.comm_ctas,
                                              allreduce_algo=a.allreduce_algo,
                                              tail_bucket_kb=a.tail_kb,
                                              oneshot_max_kb=a.oneshot_max_kb)
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    if a.fused_optimizer is None:
        a.fused_optimizer = world > 1 and a.amp_dtype != "fp16"
    if ddp is not None and a.fused_optimizer:
        a.fused_optimizer = bool(ddp.enable_fused_sgd(opt))
    fp16 = a.amp_dtype == "fp16"
    if fp16 and a.fused_optimizer:
        a.fused_optimizer = False        # GradScaler skip-step semantics need the separate update
    scaler = GradScaler(device=device, enabled=fp16)  # bf16: no loss scaling needed
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    step = TrainStep(model, opt, scaler, batch_shape=shape, cuda_graph=not a.no_graph,
                     lazy_buffer_sync=not a.eager_buffer_sync,
                     dtype={"fp16": torch.float16, "fp32": torch.float32}.get(a.amp_dtype, torch.bfloat16))
    batches = make_host_batches(8, shape, a.num_classes, seed=1234 + rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if a.l2_flush else None

    # ---- warm-up (untimed) ----------------------------------------------------------------
    for i in range(max(a.warmup, 3)):
        x, t = batches[i % len(batches)]
        step(x, t)
    barrier_sync()

    # ---- device-timed region: K steps, per-step events, L2 flushed (untimed) between steps ----
    K = a.steps
    x0, t0 = batches[0]
    step.x.copy_(x0, non_blocking=True)
    step.t.copy_(t0, non_blocking=True)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(local)
    barrier_sync()
    sampler.start()
    launches_before = _lib.LAUNCHES
    for i in range(K):
        if flush is not None:
            flush.zero_()
        starts[i].record()
        if step.graph is not None:
            step.graph.replay()
        else:
            step._body()
        ends[i].record()
    barrier_sync()
    clocks = sampler.stop()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    dev_ms = max_over_ranks(dev_ms, device)
    launches = step.launches_per_step * K if step.graph is not None else (_lib.LAUNCHES - launches_before) + 2 * K
    if ddp is not None:
        ddp.comm.check_health()

    # ---- end to end through the public API: H2D from pinned memory + loss read every step -----
    barrier_sync()
    t_begin = time.perf_counter()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for i in range(K):
        x, t = batches[i % len(batches)]
        step(x, t)
        v = step.read_loss(lag=1)   # host read of the previous step's loss (D2H every step)
        last = v if v is not None else last
    final = step.read_loss(lag=0)
    e1.record()
    barrier_sync()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1), device)
    wall_ms = max_over_ranks((time.perf_counter() - t_begin) * 1e3, device)
    e2e_ms = max(e2e_ms, wall_ms)

    sync = {}
    gb = a.batch_size * world
    res = {
        "metric": "samples/sec (device-timed, max over ranks)",
        "value": gb * K / (dev_ms * 1e-3),
        "unit": "samples/s",
        "n_gpus": world, "steps": K, "warmup": max(a.warmup, 3),
        "ms_per_step": dev_ms / K,
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None if BASELINE_VALUE is None else gb * K / (dev_ms * 1e-3) / BASELINE_VALUE,
        "dtype": a.amp_dtype + (" + dynamic GradScaler" if fp16 else "") +
                 (" activations, TF32 tensor-core math (no AMP)" if a.amp_dtype == "fp32" else ""),
        "data": "synthetic (random-init weights, random CIFAR-shaped batches)",
        "impl": "ours",
        "config": {"model": f"{a.arch}(num_classes={a.num_classes})",
                   "global_batch": gb, "per_gpu_batch": a.batch_size,
                   "input": f"{a.batch_size}x3x{a.image_size}x{a.image_size} fp32 NCHW per GPU",
                   "seq_len": None, "parallelism": f"dp{world}", "optimizer": "SGD(0.1, m=0.9, wd=5e-4)",
                   "amp": f"{a.amp_dtype} activations/weights, fp32 master+grads+BN stats",
                   "cuda_graph": step.graph is not None,
                   "l2": "flushed between timed steps (untimed 256 MiB memset)" if flush is not None
                         else "not flushed: one step touches ~207 MB (fp32 params+grads+momentum, bf16 shadow, "
                              "activations) > 126 MB L2",
                   "grad_allreduce": (f"native NVLink peer-memory kernels (per bucket: see bucket_algorithms), "
                                      f"{a.comm_ctas} CTAs/bucket, fused 1/world"
                                      + (" + SGD update" if a.fused_optimizer else "") + " epilogue")
                                     if world > 1 else "none (1 GPU)",
                   "timed_inputs": "device-timed `value`: K graph replays on one resident batch (inputs "
                                   "already on the device); `e2e`: 8 distinct pinned host batches, H2D + "
                                   "loss D2H every step",
                   "fused_optimizer": bool(a.fused_optimizer and world > 1),
                   "nvls": bool(getattr(ddp, "nvls", False)) if ddp is not None else False,
                   "grad_buckets_bytes": ddp.bucket_sizes() if ddp is not None else [],
                   "bn_buffer_sync": "lazy (before eval/checkpoint)" if not a.eager_buffer_sync else "every step"},
        "clocks": clocks,
        "e2e": {"value": gb * K / (e2e_ms * 1e-3), "unit": "samples/s",
                "ms_per_step": e2e_ms / K,
                "h2d_bytes_per_step": batches[0][0].numel() * 4 + batches[0][1].numel() * 8,
                "d2h_bytes_per_step": 4,
                "note": "TrainStep.__call__ per step: pinned H2D of images+labels, graph replay, "
                        "async D2H of the loss read by the host one step later"},
        "gpu_launches": int(launches),
        "launches_per_step": int(step.launches_per_step),
        "final_loss": final,
    }
    if sync:
        res["grad_sync"] = sync
    # ---- evidence beside the number (all cold path, after the timed regions) ------------------
    # The bench line must never be lost to an extra: if the extras have not finished after
    # `--extras-timeout` seconds (a rank stuck in a collective, say), every rank stops and rank 0
    # prints the line with what it has.
    printed = threading.Lock()

    def emit(note=None):
        if not printed.acquire(blocking=False):
            return
        if rank == 0:
            for _ in range(3):          # the main thread may be adding a key at this very moment
                try:
                    line = dict(res)
                    if note:
                        line["extras_note"] = note
                    print(json.dumps(line), flush=True)
                    break
                except RuntimeError:
                    time.sleep(0.05)

    def bail():
        emit(f"extras did not finish within {a.extras_timeout:.0f} s and were abandoned")
        os._exit(0)

    timer = threading.Timer(a.extras_timeout, bail)
    timer.daemon = True
    timer.start()

    def attempt(name, fn):
        try:
            res[name] = fn()
        except Exception as e:  # noqa: BLE001 - never lose the bench line to an extra
            res[name] = {"error": f"{type(e).__name__}: {e}"[:300]}

    if ddp is not None:
        res["config"]["bucket_algorithms"] = ddp.bucket_algorithms()
        if not a.no_extras:
            attempt("multi_gpu_checks", lambda: multi_gpu_checks(ddp, device, opt))
        ddp.shutdown()
    del step
    if not a.no_extras and a.amp_dtype == "bf16":
        attempt("fp16_gradscaler", lambda: fp16_scaler_arm(a, device, rank, world, local))
    if not a.no_extras and a.arch in ("resnet18", "resnet34", "resnet50"):
        attempt("loss_parity", lambda: loss_parity(a, device, rank, world, local))
    timer.cancel()
    emit()
    cleanup_distributed()


# ------------------------------------------------------------------------ correctness evidence
def multi_gpu_checks(ddp, device, optimizer=None) -> dict:
    """Cold-path proof, printed with the bench line, that the native gradient path reduces
    CORRECTLY (the scaling run is the only multi-GPU run the driver does): replicas bit-identical,
    every rank issued the same collectives, and a real bucket all-reduced by each native kernel
    equals torch.distributed's NCCL all-reduce of the same data."""
    import torch
    import torch.distributed as dist
    out = {}
    out["replicas_bit_identical"] = bool(ddp.verify_replicas(raise_on_mismatch=False))
    try:
        ep = ddp.comm.validate_epochs()
        out["collective_epochs_agree"] = True
        out["collectives_issued_per_cta0"] = int(ep[0])
    except RuntimeError as e:
        out["collective_epochs_agree"] = False
        out["epoch_error"] = str(e)[:200]
    world = ddp.world_size
    errs = {}
    # the largest bucket through the two-shot kernel (and NVLS when bound), the tail bucket
    # through the one-shot push kernel
    big = max(ddp.plans, key=lambda pl: pl["numel"])
    small = ddp.plans[-1]
    gen = torch.Generator(device=device).manual_seed(4242 + ddp.rank)
    for name, pl in (("twoshot", big), ("nvls", big), ("oneshot", small)):
        n = pl["numel"]
        view = ddp.arena.grad[pl["offset"]: pl["offset"] + n]
        x = torch.randn(n, device=device, generator=gen)
        ref = x.clone()
        dist.all_reduce(ref)
        ref /= world
        view.copy_(x)
        torch.cuda.synchronize()
        dist.barrier()
        if name == "twoshot":
            ddp.comm.allreduce_(ddp.arena.grad_buf, pl["offset"] * 4, n)
        elif name == "nvls":
            if not hasattr(ddp.arena.grad_buf, "mc_ptr"):
                continue
            ddp.comm.allreduce_nvls_(ddp.arena.grad_buf, pl["offset"] * 4, n)
        else:
            if getattr(ddp, "oneshot_buf", None) is None or n * 4 > ddp.oneshot_max_bytes:
                continue
            ddp.comm.allreduce_oneshot_push_(view, ddp.oneshot_buf, ddp.oneshot_cap)
        torch.cuda.synchronize()
        err = torch.tensor([(view - ref).abs().max().item()], device=device, dtype=torch.float64)
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
        errs[name] = {"bytes": n * 4, "max_abs_err_vs_nccl": float(err.item())}
        view.zero_()
    out["allreduce_vs_nccl"] = errs
    # the REAL step-time gradient path (reducer -> per-bucket kernels with the fused SGD update,
    # whatever algorithm each bucket uses by default) on known gradients, against NCCL all-reduce
    # + the SGD formula; parameters / momentum / shadow are restored afterwards
    if optimizer is not None and getattr(ddp, "_fusion_cfg", None) is not None:
        g0 = optimizer.param_groups[0]
        arena = ddp.arena
        m = optimizer._flat_momentum()
        snap_p, snap_m = arena.param.clone(), m.clone()
        m.zero_()
        x = torch.randn(arena.numel, device=device, generator=gen)
        ref = x.clone()
        dist.all_reduce(ref)
        ref /= world
        arena.grad.copy_(x)
        torch.cuda.synchronize()
        dist.barrier()
        ddp.prepare_for_backward()
        for p in list(arena.order):
            ddp._sink([p])
        ddp.finalize_backward()
        torch.cuda.synchronize()
        expect = snap_p - g0["lr"] * (ref + g0["weight_decay"] * snap_p)   # first step: m = d
        worst = 0.0
        for pl in ddp.plans:
            a, b = pl["offset"], pl["offset"] + pl["numel"]
            worst = max(worst, (arena.param[a:b] - expect[a:b]).abs().max().item())
        shadow_ok = bool(torch.equal(arena.shadow, arena.param.to(arena.shadow.dtype)))
        t = torch.tensor([worst, 0.0 if shadow_ok else 1.0], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["fused_update_path"] = {"bucket_algorithms": ddp.bucket_algorithms(),
                                    "param_max_abs_err_vs_nccl_plus_sgd": float(t[0].item()),
                                    "shadow_equals_cast_of_param": bool(t[1].item() == 0.0),
                                    "replicas_bit_identical_after": bool(ddp.verify_replicas(False))}
        arena.param.copy_(snap_p)
        m.copy_(snap_m)
        arena.grad.zero_()
        arena.refresh_shadow()
        torch.cuda.synchronize()
        dist.barrier()
    ddp.comm.check_health()
    return out


def loss_parity(a, device, rank, world, local, steps=8) -> dict:
    """N-step loss parity on the SAME weights and the SAME batches: this
    framework's native model under its native DDP vs torchvision's model with the identical
    initial state_dict under torch DDP + NCCL, torch.autocast(bf16) and torch.optim.SGD — the
    stack the reference drives (train_ddp.py:154, 204-214, 305-310), lr 0.01."""
    import torch
    import torch.nn.functional as F
    import torchvision
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel
    torch.manual_seed(777)                      # same init on every rank and for both arms
    ours = build_model(device, a.arch, a.num_classes)
    init = {k: v.detach().clone() for k, v in ours.state_dict().items()}
    ddp = None
    if world > 1:
        ours = ddp = DistributedDataParallel(ours, device_ids=[local], backend="native")
    opt = SGD(ours.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    if ddp is not None:
        ddp.enable_fused_sgd(opt)
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    batches = make_host_batches(steps, shape, a.num_classes, seed=99 + rank)
    step = TrainStep(ours, opt, None, batch_shape=shape, cuda_graph=True)
    l_ours = []
    for x, t in batches:
        step(x, t)
        l_ours.append(step.read_loss(lag=0))
    torch.cuda.synchronize()
    if ddp is not None:
        ddp.comm.check_health()
        ddp.shutdown()
    tv = getattr(torchvision.models, a.arch)(num_classes=a.num_classes).to(device)
    tv.load_state_dict(init)
    if world > 1:
        tv = torch.nn.parallel.DistributedDataParallel(tv, device_ids=[local])
    topt = torch.optim.SGD(tv.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    l_ref = []
    for x, t in batches:
        x, t = x.to(device, non_blocking=True), t.to(device, non_blocking=True)
        topt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = F.cross_entropy(tv(x), t)
        loss.backward()
        topt.step()
        l_ref.append(float(loss.item()))
    rel = max(abs(o - r) / max(abs(r), 1e-6) for o, r in zip(l_ours, l_ref))
    return {"what": "rank-0 training loss per step, same initial weights and batches, lr 0.01, bf16",
            "loss_ours": [round(v, 4) for v in l_ours],
            "loss_torch_ddp_nccl": [round(v, 4) for v in l_ref],
            "max_rel_diff": rel}


def fp16_scaler_arm(a, device, rank, world, local) -> dict:
    """The reference's LITERAL --amp numerics (fp16 autocast + dynamic GradScaler,
    train_ddp.py:203-209, 346) through the native path: same-precision throughput next to the
    bf16 headline, device-timed the same way."""
    import torch
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel
    model = build_model(device, a.arch, a.num_classes)
    ddp = None
    if world > 1:
        model = ddp = DistributedDataParallel(model, device_ids=[local], backend="native",
                                              bucket_cap_mb=a.bucket_cap_mb, comm_ctas=a.comm_ctas)
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    scaler = GradScaler(device=device, enabled=True)
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    step = TrainStep(model, opt, scaler, batch_shape=shape, cuda_graph=True, dtype=torch.float16)
    x, t = make_host_batches(1, shape, a.num_classes, seed=5 + rank)[0]
    for _ in range(max(a.warmup, 3)):
        step(x, t)
    barrier_sync()
    K = a.steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step.graph.replay()
    e1.record()
    barrier_sync()
    ms = max_over_ranks(e0.elapsed_time(e1), device)
    res = {"value": a.batch_size * world * K / (ms * 1e-3), "unit": "samples/s",
           "ms_per_step": ms / K, "dtype": "fp16 + dynamic GradScaler (device-side skip/update)",
           "loss_scale_after": scaler.get_scale(), "launches_per_step": int(step.launches_per_step)}
    if ddp is not None:
        ddp.comm.check_health()
        ddp.shutdown()
    return res



def run_profile_sync(a):
    """Separate pass (`--profile-sync`): eager (non-graph) steps with per-bucket CUDA events —
    % of step time spent in exposed gradient sync, max over ranks (BASELINE.json metric 2)."""
    import torch
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel
    from dpt_b200.profiler import GradSyncProfiler
    from dpt_b200.runtime import cleanup_distributed, set_seed, setup_distributed
    rank, world, local = setup_distributed("native")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    set_seed(42, rank)
    model = DistributedDataParallel(build_model(device, a.arch, a.num_classes), device_ids=[local],
                                    backend="native", profile=True)
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    step = TrainStep(model, opt, GradScaler(device=device, enabled=False), batch_shape=shape,
                     cuda_graph=False)
    batches = make_host_batches(4, shape, a.num_classes, seed=99 + rank)
    prof = GradSyncProfiler(model, sample_first=10 ** 9)
    for i in range(max(a.warmup, 3)):
        step(*batches[i % 4])
    barrier_sync()
    for i in range(a.steps):
        prof.step_begin()
        step(*batches[i % 4])
        prof.step_end()
    summ = prof.summary()
    last = prof.records[-1].get("buckets", [])
    res = {"metric": "% step time in exposed gradient sync (device events, max over ranks)",
           "impl": "ours", "n_gpus": world, "steps": a.steps, "mode": "eager launches (event timing is "
           "not available inside a CUDA graph)", "grad_sync": summ,
           "buckets_last_step": last, "bucket_bytes": model.bucket_sizes()}
    if rank == 0:
        print(json.dumps(res))
    model.shutdown()
    cleanup_distributed()


# ====================================================================================== reference
def run_reference(a):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "train_ddp.py")):
        print(json.dumps({"impl": "reference",
                          "unavailable": "baseline/_ref/train_ddp.py missing (run baseline/install_reference.sh)"}))
        return
    try:
        sys.path.insert(0, ref_dir)
        import torch
        import torch.nn as nn
        import torch.optim as optim
        import train_ddp as ref  # the unmodified reference script
    except Exception as e:  # e.g. torchvision missing
        print(json.dumps({"impl": "reference", "unavailable": f"import failed: {e!r}"[:200]}))
        return
    import types
    import warnings
    warnings.filterwarnings("ignore")
    # the reference's own bootstrap: env:// NCCL group + set_device (train_ddp.py:53-68)
    rank, world, local = ref.setup_distributed()
    if world == 1:
        torch.cuda.set_device(local)
    device = torch.device(f"cuda:{local}")
    ref.set_seed(42, rank)
    torch.backends.cudnn.benchmark = True            # train_ddp.py:329
    if a.arch == "resnet18":
        model = ref.build_model(device)               # torchvision resnet18(num_classes=10)
    else:
        # the reference only knows resnet18 (train_ddp.py:154); other BASELINE.json configs use
        # the same torchvision family through the reference's own loop
        import torchvision
        model = getattr(torchvision.models, a.arch)(num_classes=a.num_classes).to(device)
    model = ref.maybe_wrap_ddp(model, device, local)  # torch DDP + NCCL when world > 1
    criterion = nn.CrossEntropyLoss().to(device)
    optimizer = optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    ref_amp = a.amp_dtype != "fp32"                   # --amp-dtype fp32 = the reference's FP32 branch
    scaler = torch.cuda.amp.GradScaler(enabled=ref_amp)  # --amp path (fp16 autocast + GradScaler)
    args = types.SimpleNamespace(amp=ref_amp, print_freq=10 ** 9)
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    batches = make_host_batches(8, shape, a.num_classes, seed=1234 + rank)

    class Loader:
        """Stands in for the CIFAR DataLoader (no dataset files offline): yields pinned host
        batches exactly like DataLoader(pin_memory=True) does."""
        def __init__(self, n):
            self.n = n
        def __len__(self):
            return self.n
        def __iter__(self):
            for i in range(self.n):
                yield batches[i % len(batches)]

    W = max(a.warmup, 3)
    ref.train_one_epoch(0, model, Loader(W), optimizer, criterion, device, args, rank, world,
                        scaler=scaler, train_sampler=None)
    barrier_sync()
    sampler = ClockSampler(local)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    sampler.start()
    t0 = time.perf_counter()
    e0.record()
    ref.train_one_epoch(1, model, Loader(a.steps), optimizer, criterion, device, args, rank, world,
                        scaler=scaler, train_sampler=None)
    e1.record()
    barrier_sync()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    ms = max_over_ranks(max(e0.elapsed_time(e1), 0.0), device)
    wall_ms = max_over_ranks(wall_ms, device)
    gb = a.batch_size * world
    val = gb * a.steps / (ms * 1e-3)
    res = {
        "metric": "samples/sec (device-timed, max over ranks)", "value": val, "unit": "samples/s",
        "n_gpus": world, "steps": a.steps, "warmup": W, "ms_per_step": ms / a.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp16 autocast + GradScaler (the reference's --amp path)" if ref_amp
                 else "fp32 (TF32 cuDNN convolutions; the reference's non-AMP branch)",
        "data": "synthetic",
        "impl": "reference",
        "config": {"model": f"torchvision {a.arch}(num_classes={a.num_classes})", "global_batch": gb,
                   "per_gpu_batch": a.batch_size, "seq_len": None, "parallelism": f"dp{world}",
                   "path": "reference train_ddp.train_one_epoch (unmodified) with a synthetic pinned "
                           "loader; torch DDP + NCCL when world > 1; cudnn.benchmark=True",
                   "l2": "not flushed (the reference loop is timed as one region)"},
        "clocks": clocks,
        "e2e": {"value": gb * a.steps / (wall_ms * 1e-3), "unit": "samples/s",
                "h2d_bytes_per_step": batches[0][0].numel() * 4 + batches[0][1].numel() * 8,
                "d2h_bytes_per_step": 12,
                "note": "the reference loop itself copies pinned inputs H2D and reads loss/correct "
                        "with .item() every step, so its device-timed and end-to-end regions coincide"},
        # kernels of THIS repository launched by the reference arm: none by construction (it runs
        # cuDNN / cuBLAS / ATen / NCCL kernels, which are not counted here)
        "gpu_launches": None,
        "gpu_launches_note": "library kernels only (cuDNN/cuBLAS/ATen/NCCL), ~190 per step per "
                             "SURVEY 3.2; none from this repository",
    }
    if rank == 0:
        print(json.dumps(res))
    ref.cleanup_distributed()


def main():
    a = parse()
    _, world, _ = env_world()
    if a.gpus != world and world == 1 and a.gpus > 1:
        # convenience: self-launch one process per GPU
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1", "--master-port", "29531",
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if a.impl == "reference":
        run_reference(a)
    elif a.profile_sync:
        run_profile_sync(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
