"""Command-line interface: the reference's 11 flags with identical names/defaults/types
(reference train_ddp.py:19-46) plus the switches this framework adds."""
from __future__ import annotations

import argparse


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="DDP training of ResNet-18 on CIFAR-10")
    # ---- reference flags (train_ddp.py:22-43) ------------------------------------------
    p.add_argument("--data-dir", default="./data", type=str, help="directory to store CIFAR-10")
    p.add_argument("--epochs", default=10, type=int, help="number of total epochs to run")
    p.add_argument("--batch-size", default=128, type=int, help="mini-batch size *per GPU*")
    p.add_argument("--workers", default=4, type=int,
                   help="number of data loading workers per process")
    p.add_argument("--lr", default=0.1, type=float, help="initial learning rate")
    p.add_argument("--momentum", default=0.9, type=float, help="SGD momentum")
    p.add_argument("--weight-decay", default=5e-4, type=float, help="weight decay")
    p.add_argument("--amp", action="store_true", help="use automatic mixed precision (AMP)")
    p.add_argument("--print-freq", default=50, type=int, help="print frequency (in steps)")
    p.add_argument("--output-dir", default="./experiments", type=str,
                   help="directory to save logs")
    p.add_argument("--seed", default=42, type=int, help="random seed")
    # ---- additions ---------------------------------------------------------------------
    p.add_argument("--backend", default="native", choices=["native", "nccl", "gloo"],
                   help="gradient data plane: native = NVLink peer-memory kernels (CUDA), "
                        "nccl/gloo = torch.distributed collectives (baseline / CPU)")
    p.add_argument("--amp-dtype", default="bf16", choices=["fp16", "bf16"],
                   help="low-precision dtype under --amp (reference autocast default is fp16)")
    p.add_argument("--synthetic", action="store_true",
                   help="synthetic CIFAR-shaped data (no dataset files / network needed)")
    p.add_argument("--synthetic-size", default=50000, type=int,
                   help="number of synthetic training samples")
    p.add_argument("--arch", default="resnet18", choices=["resnet18", "resnet34", "resnet50", "resnet101", "resnet152"],
                   help="model architecture")
    p.add_argument("--num-classes", default=10, type=int)
    p.add_argument("--image-size", default=32, type=int)
    p.add_argument("--steps", default=0, type=int,
                   help="stop every epoch after this many steps (0 = full epoch)")
    p.add_argument("--optimizer", default="sgd", choices=["sgd", "adam", "adamw"])
    p.add_argument("--save-every", default=0, type=int, help="checkpoint every N epochs (0 = off)")
    p.add_argument("--resume", default="", type=str,
                   help="checkpoint file to resume from; 'auto' = <output-dir>/checkpoint.pt if it "
                        "exists (for torchrun --max-restarts)")
    p.add_argument("--bucket-cap-mb", default=None, type=float,
                   help="gradient bucket size cap in MiB (default: 4 for the native backend, whose "
                        "bucket kernels cost microseconds to launch; torch's 25 for nccl / gloo)")
    p.add_argument("--allreduce-algo", default="auto",
                   choices=["auto", "autotune", "oneshot", "twoshot", "nvls", "twoshot_mc"],
                   help="peer-memory all-reduce algorithm for the native backend")
    p.add_argument("--fused-optimizer", dest="fused_optimizer", action="store_true", default=None,
                   help="apply the SGD update inside the all-reduce kernel (default: on whenever "
                        "it applies — native backend, SGD with momentum, no dynamic loss scaling)")
    p.add_argument("--no-fused-optimizer", dest="fused_optimizer", action="store_false")
    p.add_argument("--cuda-graph", dest="cuda_graph", action="store_true", default=None,
                   help="capture the whole training step in a CUDA graph (default: on whenever the "
                        "native step applies — CUDA device, flat-arena optimizer)")
    p.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false")
    p.add_argument("--eval-amp", action="store_true",
                   help="run validation in the --amp dtype too (the reference validates in "
                        "FP32/TF32 even with --amp, train_ddp.py:266-300; that is the default)")
    p.add_argument("--gpu-augment", action="store_true",
                   help="CIFAR files: crop / flip / normalise on the GPU (one kernel per batch) "
                        "instead of per sample in the loader workers")
    p.add_argument("--profile-sync", action="store_true",
                   help="record per-step gradient-sync timing (device events, max over ranks)")
    p.add_argument("--metrics-json", default="", type=str, help="write per-epoch metrics as JSON")
    p.add_argument("--check-sync-every", default=0, type=int,
                   help="assert bit-identical parameters on all ranks every N steps (0 = off)")
    p.add_argument("--nvtx", action="store_true", help="emit NVTX ranges for the step phases")
    p.add_argument("--no-validate", action="store_true", help="skip the validation pass")
    p.add_argument("--comm-timeout", default=30.0, type=float,
                   help="seconds a device-side flag wait may spin before reporting a dead peer")
    return p


def parse_args(argv=None) -> argparse.Namespace:
    return build_parser().parse_args(argv)
