"""TrainStep — the whole training step as one static kernel sequence, optionally one CUDA graph.

Reference hot loop (train_ddp.py:195-226): H2D copy, zero_grad, autocast forward, loss,
scaler.scale(loss).backward() (DDP all-reduce overlapped), scaler.step, scaler.update, then two
``.item()`` host syncs for the running metrics — ~190 kernel launches, 3 host syncs, 2 buffer
broadcasts and 3 NCCL all-reduces per step, all issued from Python.

Here the same step is a fixed plan over preallocated buffers:
    [H2D] -> memset(grad arena, BN stats) -> engine.forward -> fused CE (+loss scale, +metrics)
          -> engine.backward (bucket all-reduces on the comm stream as buckets close)
          -> flat optimizer kernel (reads found_inf on the device) -> scaler update kernel
with no allocation and no host synchronisation, so it can be captured once
(``cuda_graph=True``) and replayed with a single launch per step.  Loss / accuracy accumulate in
a device-side ``metrics`` vector that the host reads only when it wants to print.
"""
from __future__ import annotations

import torch

from .models.resnet import ResNet
from .parallel.ddp import DistributedDataParallel


class TrainStep:
    def __init__(self, model, optimizer, scaler=None, *, batch_shape, cuda_graph: bool = True,
                 warmup: int = 2, lazy_buffer_sync: bool = True, dtype=torch.bfloat16,
                 trace_records: int = 0) -> None:
        self.lazy_buffer_sync = lazy_buffer_sync
        # optional device-clock timeline (benchmarks/step_trace.py): phase stamps of this step
        # and, under DDP, one record per collective kernel — valid under CUDA-graph replay
        self.trace = None
        self._trace_records = int(trace_records)
        self.dtype = dtype
        self.ddp = model if isinstance(model, DistributedDataParallel) else None
        self.net = model.module if self.ddp is not None else model
        if not isinstance(self.net, ResNet):
            raise TypeError("TrainStep drives models with a native engine (dpt_b200 ResNet)")
        from .ops import functional as Fn
        self.Fn = Fn
        self.optimizer = optimizer
        self.scaler = scaler if (scaler is not None and scaler.is_enabled()) else None
        self.arena = getattr(optimizer, "arena", None)
        if self.arena is None or not getattr(optimizer, "_flat", False):
            raise RuntimeError("TrainStep needs a flat-arena optimizer (dpt_b200.optim)")
        dev = self.arena.param.device
        self.device = dev
        B = batch_shape[0]
        K = self.net.fc.out_features
        self.batch_shape = tuple(batch_shape)
        self.x = torch.zeros(self.batch_shape, device=dev, dtype=torch.float32)
        self.t = torch.zeros(B, device=dev, dtype=torch.int64)
        self.loss = torch.zeros((), device=dev, dtype=torch.float32)
        self.dlogits = torch.zeros((B, K), device=dev, dtype=torch.float32)
        self.metrics = torch.zeros(3, device=dev, dtype=torch.float64)
        self.net.train()
        from .ops import functional as _Fn
        _Fn.set_lowp_dtype(dtype)
        if dtype != torch.float32:          # FP32 mode reads the fp32 masters, no shadow retyping
            self.arena.set_shadow_dtype(dtype)
        self.eng = self.net.engine_for(self.batch_shape, True, dtype)
        self.fused_scaler = False
        if self.scaler is not None:
            self.scaler._lazy_init(dev)
            if self.ddp is not None and self.ddp.backend == "native":
                self.ddp.bind_scaler(self.scaler)
                self.fused_scaler = True
        # pinned ring for lagged loss reads (no per-step host stall)
        self._ring = 4
        self.loss_host = torch.zeros(self._ring, dtype=torch.float32).pin_memory() \
            if dev.type == "cuda" else torch.zeros(self._ring)
        self._events = [None] * self._ring
        self._slot = 0
        # host->device staging on a copy stream: the next batch's H2D overlaps the current
        # step's compute (the call returns without waiting, so the caller is already copying
        # batch t+1 while step t runs)
        if dev.type == "cuda":
            self.copy_stream = torch.cuda.Stream(device=dev)
            self.x_stage = torch.zeros_like(self.x)
            self.t_stage = torch.zeros_like(self.t)
            self._stage_ready = torch.cuda.Event()
            self._stage_free = torch.cuda.Event()
            self._stage_free.record()
        else:
            self.copy_stream = None
        self.steps = 0
        self.graph = None
        self.launches_per_step = 0
        # debug: poison the gradient arena once the optimizer has consumed it, so
        # anything that reads gradients after their consumption (a late bucket kernel, a stale
        # view) turns into NaNs instead of silently using old data.  The next step's memset
        # clears the poison.
        import os
        self._poison = os.environ.get("DPT_DEBUG_POISON", "0") == "1"
        self._setup_chunked_sgd()
        if self._trace_records > 0 and dev.type == "cuda":
            self.trace = torch.zeros((1 + self._trace_records) * 4, dtype=torch.int64, device=dev)
            if self.ddp is not None and self.ddp.comm is not None:
                _Fn._lib.lib().dpt_comm_set_trace(self.ddp.comm.handle, self.trace.data_ptr(),
                                                  self._trace_records)
        if cuda_graph and dev.type == "cuda":
            self._capture(warmup)

    # ------------------------------------------------------------------ chunked optimizer
    def _setup_chunked_sgd(self) -> None:
        """Single-GPU counterpart of the fused-SGD all-reduce: the SGD update of a chunk of the
        arena is launched on the engine's side stream as soon as the last weight gradient of
        that chunk has been produced, so the 246 MB optimizer sweep overlaps the rest of backward
        instead of trailing it.  Only without a scaler (skip-step needs the global found_inf)."""
        from .optim import SGD
        from .parallel.bucket import plan_arena_buckets
        self._chunks = None
        if self.ddp is not None or self.scaler is not None or not isinstance(self.optimizer, SGD):
            return
        if self.device.type != "cuda" or getattr(self.net, "_grad_sink", None) is not None:
            return
        g0 = self.optimizer.param_groups[0]
        if g0["momentum"] == 0:
            return
        plans = plan_arena_buckets(self.arena, 4 << 20, 1 << 20, 64 << 10)
        index = {id(p): i for i, p in enumerate(self.arena.order)}
        owner = {}
        for b, pl in enumerate(plans):
            for i in pl["params"]:
                owner[i] = b
        self._chunks = {"plans": plans, "index": index, "owner": owner, "pending": [0] * len(plans)}
        self.net._grad_sink = self._chunk_sink

    def _chunk_reset(self) -> None:
        c = self._chunks
        c["pending"] = [len(pl["params"]) for pl in c["plans"]]

    def _chunk_sink(self, params) -> None:
        from .ops import _lib
        c = self._chunks
        g0 = self.optimizer.param_groups[0]
        for p in params:
            b = c["owner"][c["index"][id(p)]]
            c["pending"][b] -= 1
            if c["pending"][b] == 0:
                pl = c["plans"][b]
                lo, hi = pl["offset"], pl["offset"] + pl["numel"]
                a = self.arena
                shadow = a.shadow[lo:hi] if (a.shadow is not None and self.dtype != torch.float32) else None
                _lib.pdl_skip_next()      # reads gradients accumulated with REDs by wgrad kernels
                self.Fn.sgd_flat(a.param[lo:hi], a.grad[lo:hi], self.optimizer._flat_momentum()[lo:hi],
                                 shadow, g0["lr"], g0["momentum"], g0["dampening"], g0["weight_decay"],
                                 g0["nesterov"])

    # ------------------------------------------------------------------ the step
    def _stamp(self, tag: int) -> None:
        if self.trace is not None:
            from .ops import _lib
            _lib.lib().dpt_trace_stamp(self.trace.data_ptr(), self._trace_records, tag, _lib.stream())

    def read_trace(self, reset: bool = True) -> list[dict]:
        """Records appended since the last reset: {"kind": "stamp"|"oneshot"|"twoshot"|"nvls",
        "tag"/"bytes", "t0", "t_barrier", "t1"} in globaltimer ns (device clock of this GPU)."""
        torch.cuda.synchronize()
        raw = self.trace.cpu().tolist()
        n = min(raw[0], self._trace_records)
        names = {1: "oneshot", 2: "twoshot", 3: "nvls", 255: "stamp"}
        out = []
        for i in range(n):
            w0, t0, tb, t1 = raw[4 + 4 * i: 8 + 4 * i]
            kind = (w0 >> 56) & 0xFF
            val = w0 & ((1 << 56) - 1)
            rec = {"kind": names.get(kind, str(kind)), "t0": t0}
            if kind == 255:
                rec["tag"] = val
            else:
                rec.update(bytes=val, t_barrier=tb, t1=t1)
            out.append(rec)
        if reset:
            self.trace.zero_()
        return out

    def _body(self) -> None:
        Fn = self.Fn
        ddp = self.ddp
        Fn.set_lowp_dtype(self.dtype)
        self._stamp(1)
        if ddp is not None:
            ddp.sync_buffers(lazy=self.lazy_buffer_sync)
            ddp.prepare_for_backward()
        self.arena.grad.zero_()
        if self._chunks is not None:
            self._chunk_reset()
        logits = self.eng.forward(self.x)
        Fn.cross_entropy(logits, self.t, loss=self.loss, dlogits=self.dlogits,
                         loss_scale=self.scaler.scale_tensor if self.scaler is not None else None,
                         metrics=self.metrics)
        self._stamp(2)
        self.eng.backward(self.dlogits)
        self._stamp(3)          # every backward kernel (incl. the last weight gradient) is done
        if ddp is not None:
            ddp.finalize_backward()
        self._stamp(4)          # ... and so is every bucket kernel: 4 - 3 = exposed gradient sync
        from .ops import _lib
        _lib.pdl_skip_next()          # the optimizer kernel follows cross-stream joins
        sc = self.scaler
        if self._chunks is not None:
            pass                          # every chunk was updated on the side stream already
        elif sc is None:
            self.optimizer.step()
        else:
            if not self.fused_scaler:
                Fn.amp_unscale(self.arena.grad, sc.inv_scale, sc.found_inf)
            self.optimizer.step(found_inf=sc.found_inf)
            sc.update()
        if self._poison:
            self.arena.grad.fill_(float("nan"))
        self._stamp(5)

    def _snapshot(self):
        snap = {"param": self.arena.param.clone(), "metrics": self.metrics.clone()}
        snap["bufs"] = [b.clone() for b in self.net.buffers()]
        snap["opt"] = {k: v.clone() for k, v in self._opt_state().items()}
        if self.scaler is not None:
            snap["scaler"] = (self.scaler._scale.clone(), self.scaler._growth_tracker.clone(),
                              self.scaler._inv_scale.clone())
        return snap

    def _opt_state(self) -> dict:
        o = self.optimizer
        out = {}
        for name in ("_momentum_flat", "_m", "_v", "_step"):
            t = getattr(o, name, None)
            if torch.is_tensor(t):
                out[name] = t
        return out

    def _restore(self, snap) -> None:
        self.arena.param.copy_(snap["param"])
        self.arena.refresh_shadow()
        self.metrics.copy_(snap["metrics"])
        for b, s in zip(self.net.buffers(), snap["bufs"]):
            b.copy_(s)
        for k, v in self._opt_state().items():
            if k in snap["opt"]:
                v.copy_(snap["opt"][k])
            else:
                v.zero_()
        if self.scaler is not None:
            s, g, i = snap["scaler"]
            self.scaler._scale.copy_(s)
            self.scaler._growth_tracker.copy_(g)
            self.scaler._inv_scale.copy_(i)
            self.scaler.found_inf.zero_()

    def _capture(self, warmup: int) -> None:
        from .ops import _lib
        snap = self._snapshot()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        before = _lib.LAUNCHES
        red = getattr(self.ddp, "reducer", None) if self.ddp is not None else None
        red_before = red.launches() if (red is not None and hasattr(red, "launches")) else 0
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self._body()
        # kernels of this library in one step: Python-launched ones + the reducer's bucket
        # all-reduces (launched from C++) + the memset kernels of the step
        red_n = (red.launches() - red_before) if (red is not None and hasattr(red, "launches")) else 0
        self.launches_per_step = _lib.LAUNCHES - before + red_n + 2
        self.graph = g
        self._restore(snap)
        torch.cuda.synchronize()
        if self.trace is not None:
            self.trace.zero_()              # drop the warm-up steps' records
        if self.ddp is not None and self.ddp.comm is not None:
            self.ddp.comm.check_health()

    # ------------------------------------------------------------------ public call
    def __call__(self, images: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        """Runs one optimisation step on a (host-pinned or device) batch; returns the device
        loss scalar (valid until the next call).  Never blocks the host."""
        if self.copy_stream is not None and not images.is_cuda:
            cs = self.copy_stream
            cs.wait_event(self._stage_free)            # the previous step has consumed the stage
            with torch.cuda.stream(cs):
                self.x_stage.copy_(images, non_blocking=True)
                self.t_stage.copy_(targets, non_blocking=True)
                self._stage_ready.record(cs)
            main = torch.cuda.current_stream()
            main.wait_event(self._stage_ready)
            self.x.copy_(self.x_stage, non_blocking=True)
            self.t.copy_(self.t_stage, non_blocking=True)
            self._stage_free.record(main)
        else:
            self.x.copy_(images, non_blocking=True)
            self.t.copy_(targets, non_blocking=True)
        if self.graph is not None:
            self.graph.replay()
        else:
            from .ops import _lib
            before = _lib.LAUNCHES
            self._body()
            self.launches_per_step = _lib.LAUNCHES - before + 2
        slot = self._slot
        self.loss_host[slot:slot + 1].copy_(self.loss.view(1), non_blocking=True)
        if self.device.type == "cuda":
            ev = self._events[slot]
            if ev is None:
                ev = self._events[slot] = torch.cuda.Event()
            ev.record()
        self._slot = (slot + 1) % self._ring
        self.steps += 1
        return self.loss

    def read_loss(self, lag: int = 1) -> float | None:
        """Host value of the loss from ``lag`` steps ago (waits only for that step)."""
        if self.steps <= lag:
            return None
        slot = (self._slot - 1 - lag) % self._ring
        ev = self._events[slot]
        if ev is not None:
            ev.synchronize()
        return float(self.loss_host[slot])

    def read_metrics(self, reset: bool = False) -> tuple[float, float, float]:
        """(sum of loss*B, #correct, #samples) accumulated on the device since the last reset."""
        m = self.metrics.tolist()
        if reset:
            self.metrics.zero_()
        return m[0], m[1], m[2]
