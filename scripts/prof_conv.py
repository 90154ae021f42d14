"""Runs a few ResNet-50-sized convolutions (fprop with BN statistics, dgrad, wgrad) so one of
them can be captured with ncu:  python scripts/prof_conv.py [shape-name]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dpt_b200.ops import functional as Fn  # noqa: E402

SHAPES = {
    # name: (N, H, W, Cin, Cout, k, stride, pad)
    "l1c3": (256, 56, 56, 64, 256, 1, 1, 0),
    "l1c2": (256, 56, 56, 64, 64, 3, 1, 1),
    "l3c2": (256, 14, 14, 256, 256, 3, 1, 1),
}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "l1c3"
    N, H, W, Cin, Cout, k, s, p = SHAPES[name]
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((N, H, W, Cin), device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn((Cout, k, k, Cin), device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    ssum = torch.zeros(Cout, device="cuda")
    ssq = torch.zeros(Cout, device="cuda")
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = torch.randn((N, P, Q, Cout), device="cuda", generator=g).to(torch.bfloat16)
    for _ in range(3):
        y = Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))
        dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)
        dw = Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for label, fn in (("fprop+stats", lambda: Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))),
                      ("fprop", lambda: Fn.conv_fprop(x, w, s, p)),
                      ("dgrad", lambda: Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)),
                      ("wgrad", lambda: Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p))):
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name} {label}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us")
    print(float(y.float().abs().mean()), float(dx.float().abs().mean()), float(dw.float().abs().mean()))


if __name__ == "__main__":
    main()
