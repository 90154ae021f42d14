"""Device-clock timeline of the CUDA-graph-replayed training step (1..8 GPUs, torchrun).

What DDP's Logger/Timer gives the reference (torch reducer_timer.hpp:26-32: forward start,
backward compute start/end, backward comm start/end), but taken on the GPU's own clock from
inside the captured graph: phase stamps of the compute stream plus one record per bucket kernel
(start, after its first cross-GPU barrier, end).  From these:

  exposed_sync_us  = (all bucket kernels done) - (last backward kernel done)     per step
  step_us          = step end - step start
  local_only_us    = the same graph with every bucket kernel replaced by its single-rank stand-in
                     (Reducer::set_local_only): step time without any cross-GPU traffic

Usage:  python -m torch.distributed.run --nproc-per-node N benchmarks/step_trace.py [--steps 40]
Prints one JSON line on rank 0 (max over ranks where it matters) and writes the per-bucket
timeline of the last step to gpurun_out/step_trace_N.json.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed_replays(step, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step.graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--arch", default="resnet18")
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--image-size", type=int, default=32)
    ap.add_argument("--num-classes", type=int, default=10)
    ap.add_argument("--bucket-cap-mb", type=float, default=None)
    ap.add_argument("--allreduce-algo", default="auto")
    ap.add_argument("--comm-ctas", type=int, default=16)
    ap.add_argument("--tail-kb", type=float, default=None)
    ap.add_argument("--oneshot-max-kb", type=float, default=None)
    ap.add_argument("--no-fused", action="store_true")
    ap.add_argument("--strict-barriers", action="store_true",
                    help="closing barrier in every bucket kernel (default: only in the last)")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()

    import torch.distributed as dist
    import dpt_b200  # noqa: F401
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel
    from dpt_b200.runtime import cleanup_distributed, set_seed, setup_distributed

    rank, world, local = setup_distributed("native")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    set_seed(42, rank)
    shape = (a.batch_size, 3, a.image_size, a.image_size)
    x = torch.randn(shape, device=dev)
    t = torch.randint(0, a.num_classes, (a.batch_size,), device=dev)

    def build(local_only, trace):
        model = build_model(dev, a.arch, a.num_classes)
        ddp = None
        if world > 1:
            model = ddp = DistributedDataParallel(
                model, device_ids=[local], backend="native", bucket_cap_mb=a.bucket_cap_mb,
                comm_ctas=a.comm_ctas, allreduce_algo=a.allreduce_algo, tail_bucket_kb=a.tail_kb,
                oneshot_max_kb=a.oneshot_max_kb)
        opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
        if ddp is not None and not a.no_fused:
            ddp.enable_fused_sgd(opt)
        if ddp is not None and local_only:
            ddp.reducer.set_local_only(True)
        if ddp is not None and a.strict_barriers:
            ddp.reducer.set_relaxed_barriers(False)
        step = TrainStep(model, opt, None, batch_shape=shape, cuda_graph=True,
                         trace_records=4096 if trace else 0)
        step.x.copy_(x)
        step.t.copy_(t)
        return ddp, step

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def vmax(v):
        if world == 1:
            return v
        tt = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    out = {"n_gpus": world, "arch": a.arch, "tag": a.tag}
    # ---- 1: the real step, untraced (the number bench.py reports) -----------------------------
    ddp, step = build(False, False)
    for _ in range(10):
        step.graph.replay()
    sync()
    out["step_us"] = vmax(timed_replays(step, a.steps))
    if ddp is not None:
        out["buckets_bytes"] = ddp.bucket_sizes()
        out["bucket_algos"] = ddp.bucket_algorithms()
        ddp.comm.check_health()
        ddp.shutdown()
    del step, ddp
    sync()
    # ---- 2: same graph, bucket kernels replaced by single-rank stand-ins ----------------------
    if world > 1:
        ddp, step = build(True, False)
        for _ in range(10):
            step.graph.replay()
        sync()
        out["local_only_step_us"] = vmax(timed_replays(step, a.steps))
        out["exposed_sync_us_by_difference"] = out["step_us"] - out["local_only_step_us"]
        out["exposed_sync_pct_of_step"] = 100.0 * out["exposed_sync_us_by_difference"] / out["step_us"]
        ddp.shutdown()
        del step, ddp
        sync()
    # ---- 3: traced replays -------------------------------------------------------------------
    ddp, step = build(False, True)
    for _ in range(10):
        step.graph.replay()
    sync()
    step.read_trace(reset=True)
    n = min(a.steps, 20)
    for _ in range(n):
        step.graph.replay()
    recs = step.read_trace()
    steps, cur = [], None
    for r in sorted(recs, key=lambda r: r["t0"]):
        if r["kind"] == "stamp" and r["tag"] == 1:
            cur = {"t": {}, "buckets": []}
            steps.append(cur)
        if cur is None:
            continue
        if r["kind"] == "stamp":
            cur["t"][r["tag"]] = r["t0"]
        else:
            cur["buckets"].append(r)
    full = [s for s in steps if all(k in s["t"] for k in (1, 2, 3, 4, 5))]
    if full:
        def avg(f):
            return sum(f(s) for s in full) / len(full) / 1e3
        out["traced"] = {
            "steps": len(full),
            "fwd_us": avg(lambda s: s["t"][2] - s["t"][1]),
            "bwd_us": avg(lambda s: s["t"][3] - s["t"][2]),
            "exposed_sync_us": vmax(avg(lambda s: s["t"][4] - s["t"][3])),
            "tail_us": avg(lambda s: s["t"][5] - s["t"][4]),
            "step_us": vmax(avg(lambda s: s["t"][5] - s["t"][1])),
        }
        last = full[-1]
        t1 = last["t"][1]
        line = [{"algo": b["kind"], "bytes": b["bytes"], "start_us": (b["t0"] - t1) / 1e3,
                 "barrier_wait_us": (b["t_barrier"] - b["t0"]) / 1e3,
                 "busy_us": (b["t1"] - b["t0"]) / 1e3,
                 "end_us": (b["t1"] - t1) / 1e3,
                 "bus_gbs": (b["bytes"] * 2 * (world - 1) / world) / max(b["t1"] - b["t_barrier"], 1)}
                for b in last["buckets"]]
        out["last_step"] = {"fwd_end_us": (last["t"][2] - t1) / 1e3, "bwd_end_us": (last["t"][3] - t1) / 1e3,
                            "comm_end_us": (last["t"][4] - t1) / 1e3, "step_end_us": (last["t"][5] - t1) / 1e3,
                            "buckets": line}
    if ddp is not None:
        ddp.comm.check_health()
        ddp.shutdown()
    if rank == 0:
        print(json.dumps(out))
        os.makedirs("gpurun_out", exist_ok=True)
        name = f"gpurun_out/step_trace_{world}{('_' + a.tag) if a.tag else ''}.json"
        json.dump(out, open(name, "w"), indent=1)
    cleanup_distributed()


if __name__ == "__main__":
    main()
