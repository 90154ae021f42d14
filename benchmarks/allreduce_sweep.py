"""Gradient-bucket all-reduce sweep, 1 KB - 1 GB, native peer-memory kernels vs NCCL
(BASELINE.json config 5).  Launch with torchrun on 2/4/8 GPUs:

    torchrun --nproc-per-node 8 benchmarks/allreduce_sweep.py [--max-mb 1024] [--json out.json]

Timing: CUDA events around `iters` back-to-back calls after warm-up, max over ranks.
busBW = S/t * 2(N-1)/N (NCCL convention); roofline denominator 770 GB/s measured peer copy per
direction (900 nominal).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timeit(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--json", default="")
    ap.add_argument("--ctas", type=int, default=0)
    a = ap.parse_args()
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    world = dist.get_world_size()
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=60.0)
    sizes = []
    s = 1024
    while s <= a.max_mb * 1024 * 1024:
        sizes.append(int(s))
        s *= 4
    extra = [9461800 // 16 * 16, 26494976, 8769792, 44726568 // 16 * 16]  # ResNet-18 DDP buckets
    sizes = sorted(set(sizes + [e for e in extra if e <= a.max_mb * 1024 * 1024]))
    buf = comm.alloc_multicast(max(sizes)) or comm.alloc(max(sizes))
    has_nvls = hasattr(buf, "mc_ptr")
    out = torch.empty(min(max(sizes), 4 << 20) // 4, device="cuda")
    rows = []
    for nbytes in sizes:
        n = nbytes // 4
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (64 << 20) else 10)
        x = torch.ones(n, device="cuda")
        t_nccl = timeit(lambda: dist.all_reduce(x), iters)
        t_two = timeit(lambda: comm.allreduce_(buf, 0, n, ctas=a.ctas), iters)
        t_one = None
        if nbytes <= (4 << 20):
            o = out[:n]
            t_one = timeit(lambda: comm.allreduce_oneshot(buf, 0, o), iters)
        t_nvls = timeit(lambda: comm.allreduce_nvls_(buf, 0, n, ctas=a.ctas), iters) if has_nvls else None
        f = 2 * (world - 1) / world
        row = {"bytes": nbytes, "nccl_us": t_nccl, "twoshot_us": t_two, "oneshot_us": t_one,
               "nccl_busbw_gbs": nbytes / t_nccl * f / 1e3, "twoshot_busbw_gbs": nbytes / t_two * f / 1e3,
               "nvls_us": t_nvls, "nvls_busbw_gbs": (nbytes / t_nvls * f / 1e3) if t_nvls else None,
               "twoshot_vs_nccl": t_nccl / t_two, "nvls_vs_nccl": (t_nccl / t_nvls) if t_nvls else None,
               "twoshot_frac_of_770": nbytes / t_two * f / 1e3 / 770.0}
        rows.append(row)
        if rank == 0:
            one = f"{t_one:9.1f}" if t_one is not None else "        -"
            nv = f"{t_nvls:9.1f}" if t_nvls is not None else "        -"
            print(f"{nbytes:>12d} B  nccl {t_nccl:9.1f} us  two-shot {t_two:9.1f} us  nvls {nv} us  one-shot {one} us  "
                  f"busBW nccl {row['nccl_busbw_gbs']:7.1f}  ours {row['twoshot_busbw_gbs']:7.1f} GB/s  "
                  f"speedup {row['twoshot_vs_nccl']:5.2f}x", flush=True)
    comm.check_health()
    if rank == 0 and a.json:
        with open(a.json, "w") as fjs:
            json.dump({"world": world, "rows": rows}, fjs, indent=1)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
