"""Drop-in entry point (same name and flags as the reference script):
    python train_ddp.py --synthetic --amp
    torchrun --nproc_per_node=8 train_ddp.py --synthetic --amp --cuda-graph
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from dpt_b200.train_ddp import main  # noqa: E402

if __name__ == "__main__":
    main()
