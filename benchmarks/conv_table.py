"""Per-layer convolution table: this library's tcgen05 implicit-GEMM kernels against cuDNN (what
the reference runs through torchvision's resnet, train_ddp.py:154, 205-213) on the ResNet-18 /
CIFAR batch-128 shapes of SURVEY 2.4 and a sample of the ResNet-50 @224 batch-256 shapes.

Each timing is a CUDA graph of REPS back-to-back launches replayed several times (device time per
launch without Python / launch gaps; operands stay L2-resident for the small layers on both
sides).  cuDNN: channels_last bf16, cudnn.benchmark=True (autotuned, as the reference sets it at
train_ddp.py:329).  ours/cudnn < 1 means this library is faster.

    python benchmarks/conv_table.py [--r50]     -> gpurun_out/conv_table.json + a text table
"""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dpt_b200.ops import functional as Fn  # noqa: E402

REPS = 20


def graph_time(fn, reps=REPS, replays=5):
    """us per call of fn, from a captured graph of `reps` calls."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * replays) * 1e3


def main():
    torch.backends.cudnn.benchmark = True
    dev = "cuda"
    r18 = [(128, 8, 8, 64, 64, 3, 1, 1), (128, 8, 8, 64, 128, 3, 2, 1), (128, 4, 4, 128, 128, 3, 1, 1),
           (128, 4, 4, 128, 256, 3, 2, 1), (128, 2, 2, 256, 256, 3, 1, 1), (128, 2, 2, 256, 512, 3, 2, 1),
           (128, 1, 1, 512, 512, 3, 1, 1), (128, 8, 8, 64, 128, 1, 2, 0)]
    r50 = [(256, 56, 56, 64, 64, 3, 1, 1), (256, 56, 56, 64, 256, 1, 1, 0), (256, 28, 28, 128, 128, 3, 1, 1),
           (256, 14, 14, 256, 256, 3, 1, 1), (256, 14, 14, 1024, 256, 1, 1, 0), (256, 7, 7, 512, 512, 3, 1, 1)]
    shapes = [("resnet18/cifar bs128", s) for s in r18]
    if "--r50" in sys.argv:
        shapes += [("resnet50/224 bs256", s) for s in r50]
    rows = []
    for family, (N, H, W, Ci, Co, k, st, pd) in shapes:
        x = torch.randn(N, H, W, Ci, device=dev).bfloat16()
        w = (torch.randn(Co, k, k, Ci, device=dev) * 0.05).bfloat16()
        P, Q = Fn.conv_out_hw(H, W, k, k, st, pd)
        y = torch.empty(N, P, Q, Co, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(N, P, Q, Co, device=dev).bfloat16()
        dx = torch.empty_like(x)
        dw = torch.zeros(Co, k, k, Ci, device=dev)
        ss, sq = torch.zeros(Co, device=dev), torch.zeros(Co, device=dev)
        xc = x.permute(0, 3, 1, 2)       # NCHW view of NHWC storage = channels_last
        wc = w.permute(0, 3, 1, 2)
        dyc = dy.permute(0, 3, 1, 2)
        row = {"family": family, "shape": [N, H, W, Ci, Co, k, st, pd],
               "gflop": 2.0 * N * P * Q * Co * Ci * k * k / 1e9}
        row["fprop_us"] = graph_time(lambda: Fn.conv_fprop(x, w, st, pd, out=y, stats=(ss, sq)))
        row["dgrad_us"] = graph_time(lambda: Fn.conv_dgrad(dy, w, tuple(x.shape), st, pd, out=dx))
        row["wgrad_us"] = graph_time(lambda: Fn.conv_wgrad(dy, x, tuple(w.shape), st, pd, out=dw))
        row["cudnn_fprop_us"] = graph_time(lambda: F.conv2d(xc, wc, stride=st, padding=pd))
        row["cudnn_dgrad_us"] = graph_time(
            lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [st, st], [pd, pd], [1, 1],
                                                        False, [0, 0], 1, [True, False, False]))
        row["cudnn_wgrad_us"] = graph_time(
            lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [st, st], [pd, pd], [1, 1],
                                                        False, [0, 0], 1, [False, True, False]))
        for op in ("fprop", "dgrad", "wgrad"):
            row[f"{op}_ours_over_cudnn"] = row[f"{op}_us"] / row[f"cudnn_{op}_us"]
        rows.append(row)
        print(f"{family:22s} {N}x{H}x{W}x{Ci}->{Co} k{k}s{st}: "
              f"fprop {row['fprop_us']:7.1f} / {row['cudnn_fprop_us']:7.1f}  "
              f"dgrad {row['dgrad_us']:7.1f} / {row['cudnn_dgrad_us']:7.1f}  "
              f"wgrad {row['wgrad_us']:7.1f} / {row['cudnn_wgrad_us']:7.1f}  us (ours / cuDNN; fprop "
              f"incl. BN statistics)", flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"reps_per_graph": REPS, "note": "us per launch, CUDA-graph timed; ours fprop also "
               "emits the BatchNorm batch statistics; wgrad accumulates into fp32 (RED)", "rows": rows},
              open("gpurun_out/conv_table.json", "w"), indent=1)


if __name__ == "__main__":
    main()
