"""Per-kernel timings of the tcgen05 implicit-GEMM kernels on the ResNet-18/CIFAR layer shapes
and on large square GEMMs (achieved TFLOP/s vs MEASURED_PEAKS.json)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dpt_b200.ops import functional as Fn  # noqa: E402


def timeit(fn, iters=50, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    dev = "cuda"
    peaks = {}
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks = json.load(open(p))
    peak = peaks.get("bf16_tflops", 1590.0)
    out = {"gemm": [], "conv": []}
    from dpt_b200.ops import _lib
    lib = _lib.lib()
    two_sm = "--two-sm" in sys.argv      # also time the experimental cta_group::2 GEMM
    for n in (2048, 4096, 8192):
        A = torch.randn(n, n, device=dev).bfloat16()
        B = torch.randn(n, n, device=dev).bfloat16()
        C = torch.empty(n, n, device=dev, dtype=torch.bfloat16)
        for op, name in ((0, "NT"), (1, "NN")):
            row = {"n": n, "op": name}
            # default policy (persistent variant from 592 tiles) and the one-tile-per-CTA deep ring
            for label, min_tiles in (("", 0), ("_one_tile", 1 << 40)):
                lib.dpt_igemm_set_persist(min_tiles, 0)
                t = timeit(lambda: Fn.gemm(op, A, B, C, n, n, n, n, n, n, out_mode=0), iters=20)
                row["us" + label] = t
                row["tflops" + label] = 2 * n ** 3 / t / 1e6
            lib.dpt_igemm_set_persist(0, 0)
            t, tf = row["us"], row["tflops"]
            row["frac_of_measured_peak"] = tf / peak
            ref = timeit(lambda: torch.matmul(A, B.t() if op == 0 else B, out=C), iters=20)
            row["cublas_us"] = ref
            if two_sm and op == 0:
                st = torch.cuda.current_stream().cuda_stream
                t2 = timeit(lambda: lib.dpt_gemm2sm_bf16(A.data_ptr(), B.data_ptr(), C.data_ptr(), n, n, n,
                                                         n, n, n, 0, 0, st), iters=20)
                row["two_sm_us"] = t2
                row["two_sm_tflops"] = 2 * n ** 3 / t2 / 1e6
            out["gemm"].append(row)
            print(f"gemm {name} {n}^3: {t:9.1f} us  {tf:7.1f} TFLOP/s ({100*tf/peak:4.1f}% of measured {peak:.0f})  "
                  f"one-tile {row['us_one_tile']:9.1f} us  cuBLAS {ref:9.1f} us"
                  + (f"  two-SM {row['two_sm_us']:9.1f} us" if "two_sm_us" in row else ""), flush=True)
    # (N, H, W, Cin, Cout, k, s, p) ResNet-18 @ 128x3x32x32
    shapes = [(128, 8, 8, 64, 64, 3, 1, 1), (128, 8, 8, 64, 128, 3, 2, 1), (128, 4, 4, 128, 128, 3, 1, 1),
              (128, 4, 4, 128, 256, 3, 2, 1), (128, 2, 2, 256, 256, 3, 1, 1), (128, 2, 2, 256, 512, 3, 2, 1),
              (128, 1, 1, 512, 512, 3, 1, 1)]
    import torch.nn.functional as F
    for (N, H, W, Ci, Co, k, s, pd) in shapes:
        x = torch.randn(N, H, W, Ci, device=dev).bfloat16()
        w = (torch.randn(Co, k, k, Ci, device=dev) * 0.05).bfloat16()
        P, Q = Fn.conv_out_hw(H, W, k, k, s, pd)
        y = torch.empty(N, P, Q, Co, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(N, P, Q, Co, device=dev).bfloat16()
        dx = torch.empty_like(x)
        dw = torch.zeros(Co, k, k, Ci, device=dev)
        ss, sq = torch.zeros(Co, device=dev), torch.zeros(Co, device=dev)
        tf = timeit(lambda: Fn.conv_fprop(x, w, s, pd, out=y, stats=(ss, sq)))
        td = timeit(lambda: Fn.conv_dgrad(dy, w, tuple(x.shape), s, pd, out=dx))
        tw = timeit(lambda: Fn.conv_wgrad(dy, x, tuple(w.shape), s, pd, out=dw))
        xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        tc = timeit(lambda: F.conv2d(xc, wc, stride=s, padding=pd))
        fl = 2.0 * N * P * Q * Co * Ci * k * k
        out["conv"].append({"shape": [N, H, W, Ci, Co, k, s, pd], "fprop_us": tf, "dgrad_us": td,
                            "wgrad_us": tw, "cudnn_fprop_us": tc, "gflop": fl / 1e9})
        print(f"conv {N}x{H}x{W}x{Ci}->{Co} k{k}s{s}: fprop {tf:6.1f}  dgrad {td:6.1f}  wgrad {tw:6.1f} us   "
              f"cuDNN fprop {tc:6.1f} us   ({fl/1e9:.2f} GFLOP)", flush=True)
    json.dump(out, open("gpurun_out/kernel_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
