"""Localises wrong outputs of the bulk-copy BatchNorm kernels (DPT_BULK_EW=1): runs the backward
(and forward) of one large shape repeatedly and prints WHERE mismatching 8-channel vectors are."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dpt_b200.ops import functional as Fn  # noqa: E402


def where(bad_vec, tag):
    idx = bad_vec.nonzero().flatten()
    if idx.numel() == 0:
        return
    ch = idx // 256
    print(f"  {tag}: {idx.numel()} bad vectors, first {idx[:4].tolist()} last {idx[-4:].tolist()}; "
          f"{ch.unique().numel()} chunks, chunk ids {ch.unique()[:8].tolist()} ... {ch.unique()[-4:].tolist()}; "
          f"lane-in-chunk min {int((idx % 256).min())} max {int((idx % 256).max())}", flush=True)


def main():
    shape = tuple(int(v) for v in os.environ.get("SHAPE", "5,99,101,256").split(","))
    relu = os.environ.get("RELU", "0") == "1"
    iters = int(os.environ.get("ITERS", "30"))
    N, H, W, C = shape
    g = torch.Generator(device="cuda").manual_seed(1)
    x = (torch.randn(shape, device="cuda", generator=g) * 2 + 0.5).bfloat16()
    dy = torch.randn(shape, device="cuda", generator=g).bfloat16()
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda")
    M = x.numel() // C
    xf = x.float().view(M, C)
    mean = xf.mean(0)
    var = xf.var(0, unbiased=False)
    istd = (var + 1e-5).rsqrt()
    yf = ((xf - mean) * istd * gamma + beta)
    y_ref = yf.relu() if relu else yf
    gf = dy.float().view(M, C)
    if relu:
        gf = gf * (y_ref.bfloat16().float() > 0)
        gf = gf.bfloat16().float()
    xh = (xf - mean) * istd
    dx_ref = gamma * istd * (gf - gf.mean(0) - xh * (gf * xh).mean(0))
    nbad_f = nbad_b = 0
    for it in range(iters):
        ssum = torch.zeros(C, device="cuda")
        ssq = torch.zeros(C, device="cuda")
        Fn.bn_stats(x, ssum, ssq)
        y, m_, i_ = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, relu=relu)
        dgamma = torch.zeros(C, device="cuda")
        dbeta = torch.zeros(C, device="cuda")
        dx0 = torch.full_like(x, float('nan'))
        dx, _ = Fn.bn_bwd(dy, x, m_, i_, gamma, dgamma, dbeta, y=y if relu else None, dx=dx0)
        torch.cuda.synchronize()
        ey = (y.float().view(M, C) - y_ref).abs().view(-1, 8).amax(1)
        ex = (dx.float().view(M, C) - dx_ref).abs().view(-1, 8).amax(1)
        nanv = torch.isnan(dx.float().view(-1, 8)).any(1)
        ex = torch.nan_to_num(ex, nan=0.0)
        by, bx = ey > 0.2, ex > 0.2
        if nanv.any():
            where(nanv, f"iter {it} UNWRITTEN dx")
        if by.any():
            nbad_f += 1
            where(by, f"iter {it} fwd y")
        if bx.any():
            nbad_b += 1
            where(bx, f"iter {it} bwd dx")
            if nbad_b <= 3:
                got8 = dx.float().view(-1, 8)
                ref8 = dx_ref.reshape(-1, 8)
                for v in bx.nonzero().flatten()[[0, 1, bx.sum().item() // 2, -1]].tolist():
                    d = (ref8 - got8[v]).abs().sum(1)
                    j = int(d.argmin())
                    print(f"    vec {v} (chunk {v // 256} lane {v % 256}): got {got8[v, :4].tolist()} ref {ref8[v, :4].tolist()} "
                          f"best match ref vec {j} (delta {j - v}, dist {d[j]:.3g}, own dist {d[v]:.3g})", flush=True)
                # is the bad output what the formula gives with dgamma = dbeta = 0 ?
                alt = (gamma * istd * gf).reshape(-1, 8)
                v = int(bx.nonzero().flatten()[0])
                print(f"    vs a*g: {alt[v, :4].tolist()}", flush=True)
        eg = (dgamma - (gf * xh).sum(0)).abs().max().item() / ((gf * xh).sum(0).abs().max().item() + 1e-6)
        if eg > 2e-2:
            print(f"  iter {it} dgamma rel err {eg:.3g}", flush=True)
    print(f"shape {shape} relu {relu} env BULK={os.environ.get('DPT_BULK_EW')} PDL_EW={os.environ.get('DPT_PDL_EW')}: "
          f"{nbad_f}/{iters} bad forwards, {nbad_b}/{iters} bad backwards", flush=True)


if __name__ == "__main__":
    main()
