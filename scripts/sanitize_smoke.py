"""Smallest program that launches every kernel family once (for compute-sanitizer / ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dpt_b200.graph_step import TrainStep
from dpt_b200.models.resnet import build_model
from dpt_b200.optim import SGD
dev = torch.device("cuda")
torch.manual_seed(0)
for dtype in (torch.bfloat16, torch.float32):
    model = build_model(dev, "resnet18", 10)
    opt = SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    step = TrainStep(model, opt, None, batch_shape=(8, 3, 32, 32), cuda_graph=False, dtype=dtype)
    x = torch.randn(8, 3, 32, 32).pin_memory(); t = torch.randint(0, 10, (8,)).pin_memory()
    for _ in range(2):
        step(x, t)
    torch.cuda.synchronize()
    print(dtype, "loss", step.read_loss(lag=0))
print("ok")
