"""Per-kernel time breakdown of one training step (diagnostic, not a benchmark).

Runs a few eager (non-graph) steps of the native engine under torch.profiler (CUPTI) and
aggregates device time by kernel name, so the step can be attributed to GEMM-shaped kernels
(igemm_kernel<...>), normalisation / pooling / loss kernels and the optimizer.  Numbers taken
under a profiler are for attribution only; bench.py reports the untraced step time.

    python benchmarks/step_breakdown.py --arch resnet50 --batch 256 --image 224 --classes 1000 \
        --out gpurun_out/breakdown_r50.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet18")
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--image", type=int, default=32)
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    dev = torch.device("cuda")
    model = build_model(dev, args.arch, args.classes)
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    shape = (args.batch, 3, args.image, args.image)
    step = TrainStep(model, opt, None, batch_shape=shape, cuda_graph=False)
    x = torch.randn(shape, device=dev)
    t = torch.randint(0, args.classes, (args.batch,), device=dev)
    for _ in range(3):
        step(x, t)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step(x, t)
    e1.record()
    torch.cuda.synchronize()
    plain_ms = e0.elapsed_time(e1) / args.steps

    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            step(x, t)
        torch.cuda.synchronize()
    rows = {}
    for ev in prof.events():
        if ev.device_type != torch.autograd.DeviceType.CUDA:
            continue
        name = ev.name
        r = rows.setdefault(name, [0, 0.0])
        r[0] += 1
        r[1] += ev.device_time_total
    total = sum(v[1] for v in rows.values())
    table = sorted(({"kernel": k, "calls_per_step": v[0] / args.steps, "us_per_step": v[1] / args.steps,
                     "pct": 100.0 * v[1] / max(total, 1e-9)} for k, v in rows.items()),
                   key=lambda r: -r["us_per_step"])
    out = {"arch": args.arch, "batch": args.batch, "image": args.image,
           "eager_step_ms_untraced": plain_ms, "kernel_time_ms_per_step": total / args.steps / 1e3,
           "kernels": table}
    for r in table[:40]:
        print(f"{r['us_per_step']:10.1f} us {r['pct']:5.1f}% x{r['calls_per_step']:6.1f}  {r['kernel'][:110]}")
    print(f"untraced eager step {plain_ms:.3f} ms; sum of kernel time {out['kernel_time_ms_per_step']:.3f} ms")
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
