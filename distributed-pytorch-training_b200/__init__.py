"""dpt_b200 — Blackwell-native data-parallel training (B200 / sm_100a).

Capabilities of yamiel-abreu/distributed-pytorch-training's ``train_ddp.py`` (reference
train_ddp.py:1-390) rebuilt B200-first: torchrun-compatible launch, process-group bootstrap,
DistributedSampler sharding, AMP vs FP32, a DDP-compatible wrapper whose gradient all-reduce is
a hand-written NVLink peer-memory kernel with fused epilogues, per-step gradient-sync profiling,
torch ``state_dict`` checkpoints, and ResNet models whose conv/BN/pool/loss kernels are
hand-written sm_100a CUDA (tcgen05 + TMEM + TMA).
"""
__version__ = "0.1.0"

from . import runtime  # noqa: F401
from .runtime import (cleanup_distributed, is_distributed, set_seed,  # noqa: F401
                      setup_distributed)
