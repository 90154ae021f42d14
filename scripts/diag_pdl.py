import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dpt_b200.graph_step import TrainStep
from dpt_b200.models.resnet import build_model
from dpt_b200.optim import SGD
dev = torch.device("cuda")
shape = (32, 3, 32, 32)
g = torch.Generator().manual_seed(0)
x = torch.randn(shape, generator=g).pin_memory(); t = torch.randint(0, 10, (32,), generator=g).pin_memory()
for graph in (False, True, False):
    torch.manual_seed(7)
    model = build_model(dev, "resnet18", 10)
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    step = TrainStep(model, opt, None, batch_shape=shape, cuda_graph=graph)
    losses = []
    for _ in range(4):
        step(x, t); losses.append(round(step.read_loss(lag=0), 5))
    print("graph" if graph else "eager", losses, "splits", [u.fsplit for u in step.eng.units][:8], flush=True)
