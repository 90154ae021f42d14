"""GradSyncProfiler — per-step, device-timed gradient-synchronisation metrics.

The reference advertises "gradient sync profiling" (README.md:6,23,35) but ships none; torch
DDP's latent Logger/Timer (logger.hpp:90-110) samples host-side.  Here every bucket kernel is
bracketed by CUDA events on the comm stream and the backward span by events on the compute
stream (csrc/reducer/reducer.cpp), giving per step:

  backward_ms          compute-stream time from start of backward to end of finalize
  comm_busy_ms         sum of bucket kernel durations
  comm_exposed_ms      time the compute stream had to wait for the last bucket after its own
                       work ended (what is NOT hidden behind backward)
  overlap_ms           comm_busy_ms - comm_exposed_ms
  per bucket           bytes, ready/start/end offsets, bus GB/s (S/t * 2(N-1)/N)

Aggregation across ranks is max (BASELINE.json metric convention).  Event timing is not
available inside a captured CUDA graph; the graph-replayed step is profiled from the device-clock
trace instead (``summarize_trace`` below: phase stamps + one record per bucket kernel written by
the kernels themselves, ``TrainStep(trace_records=...)``).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradSyncProfiler:
    KEYS = ("step_ms", "backward_ms", "comm_busy_ms", "comm_exposed_ms", "overlap_ms")

    def __init__(self, ddp=None, sample_first: int = 10, sample_every: int = 100) -> None:
        self.ddp = ddp
        self.sample_first, self.sample_every = sample_first, sample_every
        self.records: list[dict] = []
        self._step = 0
        self._ev0 = self._ev1 = None

    def should_sample(self) -> bool:
        """torch DDP samples the first 10 iterations, then every 100th (reducer.hpp:33,166-170)."""
        return self._step < self.sample_first or self._step % self.sample_every == 0

    def step_begin(self) -> None:
        if torch.cuda.is_available():
            self._ev0 = torch.cuda.Event(enable_timing=True)
            self._ev1 = torch.cuda.Event(enable_timing=True)
            self._ev0.record()

    def step_end(self) -> dict | None:
        self._step += 1
        if self._ev0 is None:
            return None
        self._ev1.record()
        self._ev1.synchronize()
        rec = {"step": self._step, "step_ms": self._ev0.elapsed_time(self._ev1)}
        t = self.ddp.sync_timings() if self.ddp is not None and hasattr(self.ddp, "sync_timings") else {}
        if t and t.get("buckets"):
            world = self.ddp.world_size
            busy = sum(b["end_ms"] - b["start_ms"] for b in t["buckets"])
            last_end = max(b["end_ms"] for b in t["buckets"])
            last_ready = max(b["ready_ms"] for b in t["buckets"])
            # compute stream finished its own backward work at ~last_ready; it then waits
            # until the last bucket kernel completes
            exposed = max(0.0, last_end - last_ready)
            rec.update(backward_ms=t["backward_ms"], comm_busy_ms=busy, comm_exposed_ms=exposed,
                       overlap_ms=max(0.0, busy - exposed), buckets=[])
            for b in t["buckets"]:
                dur = max(b["end_ms"] - b["start_ms"], 1e-6)
                rec["buckets"].append({
                    "bytes": b["bytes"], "ready_ms": b["ready_ms"], "start_ms": b["start_ms"],
                    "end_ms": b["end_ms"],
                    "bus_gbs": b["bytes"] / (dur * 1e-3) / 1e9 * (2 * (world - 1) / world if world > 1 else 1.0),
                })
        self.records.append(rec)
        return rec

    def summary(self) -> dict:
        """Mean over recorded steps, then max over ranks."""
        out = {}
        for k in self.KEYS:
            vals = [r[k] for r in self.records if k in r]
            if vals:
                out[k] = sum(vals) / len(vals)
        if out and dist.is_available() and dist.is_initialized() and torch.cuda.is_available():
            t = torch.tensor([out.get(k, 0.0) for k in self.KEYS], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out = {k: float(v) for k, v in zip(self.KEYS, t.tolist()) if k in out}
        if "step_ms" in out and "comm_exposed_ms" in out and out["step_ms"] > 0:
            out["exposed_sync_pct"] = 100.0 * out["comm_exposed_ms"] / out["step_ms"]
        return out


def summarize_trace(records: list[dict], world: int = 1, reduce_max: bool = True) -> dict:
    """Per-step gradient-sync metrics of the CUDA-graph step from ``TrainStep.read_trace()``
    records (globaltimer ns): stamps 1 step start, 2 forward done, 3 every backward kernel done,
    4 every bucket kernel done, 5 step end; bucket records carry start / first-barrier / end.
    Same keys as ``GradSyncProfiler.summary`` (ms), mean over complete steps, max over ranks."""
    steps, cur = [], None
    for r in sorted(records, key=lambda r: r["t0"]):
        if r["kind"] == "stamp" and r.get("tag") == 1:
            cur = {"t": {}, "buckets": []}
            steps.append(cur)
        if cur is None:
            continue
        if r["kind"] == "stamp":
            cur["t"][r["tag"]] = r["t0"]
        else:
            cur["buckets"].append(r)
    full = [s for s in steps if all(k in s["t"] for k in (1, 2, 3, 4, 5))]
    if not full:
        return {}
    for s in full:      # bucket kernels of eager (unstamped) steps that follow do not belong here
        s["buckets"] = [b for b in s["buckets"] if s["t"][1] <= b["t0"] <= s["t"][5]]

    def mean(f):
        return sum(f(s) for s in full) / len(full) / 1e6      # ns -> ms

    out = {
        "step_ms": mean(lambda s: s["t"][5] - s["t"][1]),
        "backward_ms": mean(lambda s: s["t"][3] - s["t"][2]),
        "comm_busy_ms": mean(lambda s: sum(b["t1"] - b["t0"] for b in s["buckets"])),
        "comm_exposed_ms": mean(lambda s: max(0, s["t"][4] - s["t"][3])),
    }
    out["overlap_ms"] = max(0.0, out["comm_busy_ms"] - out["comm_exposed_ms"])
    out["barrier_wait_ms"] = mean(lambda s: sum(max(0, b["t_barrier"] - b["t0"]) for b in s["buckets"]))
    keys = list(out)
    if reduce_max and world > 1 and dist.is_available() and dist.is_initialized() and torch.cuda.is_available():
        t = torch.tensor([out[k] for k in keys], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out = {k: float(v) for k, v in zip(keys, t.tolist())}
    if out["step_ms"] > 0:
        out["exposed_sync_pct"] = 100.0 * out["comm_exposed_ms"] / out["step_ms"]
    out["steps"] = len(full)
    return out
