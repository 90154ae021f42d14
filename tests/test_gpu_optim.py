"""Flat-arena optimizers on the GPU: checkpoint round trips continue the trajectory (the flat
kernels keep their state in arenas, so load_state_dict has to copy INTO them), and the
device-side CIFAR transform wired into the entry point."""
import os
import pickle

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mlp(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(24, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8)).cuda()


@pytest.mark.parametrize("name", ["sgd", "adam", "adamw"])
def test_flat_optimizer_resume_matches_uninterrupted_run_and_torch(name):
    from dpt_b200.optim import build_optimizer
    kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-2)
    xs = [torch.randn(16, 24, device="cuda", generator=torch.Generator(device="cuda").manual_seed(i))
          for i in range(6)]

    def step(model, opt, x):
        opt.zero_grad(set_to_none=False)
        model(x).square().mean().backward()
        opt.step()

    # uninterrupted: 6 steps
    ma = _mlp(0)
    oa = build_optimizer(name, ma.parameters(), **kw)
    assert oa._flat, "flat-arena path expected on CUDA"
    for x in xs:
        step(ma, oa, x)
    # interrupted after 3 steps: state_dict -> fresh objects -> 3 more steps
    mb = _mlp(0)
    ob = build_optimizer(name, mb.parameters(), **kw)
    for x in xs[:3]:
        step(mb, ob, x)
    sd_m = {k: v.detach().cpu().clone() for k, v in mb.state_dict().items()}
    sd_o = ob.state_dict()
    sd_o = {"state": {i: {k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                      for i, st in sd_o["state"].items()},
            "param_groups": sd_o["param_groups"]}
    mc = _mlp(1)                      # different init: everything must come from the checkpoint
    oc = build_optimizer(name, mc.parameters(), **kw)
    mc.load_state_dict(sd_m)
    oc.arena.refresh_shadow()
    oc.load_state_dict(sd_o)
    for x in xs[3:]:
        step(mc, oc, x)
    for (n, a), b in zip(ma.named_parameters(), mc.parameters()):
        assert torch.allclose(a, b, atol=1e-6, rtol=1e-5), f"{name}: {n} differs after resume"
    # and the same trajectory as the stock torch optimizer
    mt = _mlp(0)
    ot = {"sgd": lambda p: torch.optim.SGD(p, lr=0.05, momentum=0.9, weight_decay=1e-2),
          "adam": lambda p: torch.optim.Adam(p, lr=0.05, weight_decay=1e-2),
          "adamw": lambda p: torch.optim.AdamW(p, lr=0.05, weight_decay=1e-2)}[name](mt.parameters())
    for x in xs:
        step(mt, ot, x)
    for (n, a), b in zip(ma.named_parameters(), mt.parameters()):
        assert torch.allclose(a, b, atol=2e-5, rtol=1e-4), f"{name}: {n} differs from torch.optim"


def _fake_cifar(root, n_per_batch=40, seed=0):
    base = os.path.join(root, "cifar-10-batches-py")
    os.makedirs(base, exist_ok=True)
    rng = np.random.default_rng(seed)
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        d = {"data": rng.integers(0, 256, size=(n_per_batch, 3072), dtype=np.uint8),
             "labels": rng.integers(0, 10, size=n_per_batch).tolist()}
        with open(os.path.join(base, name), "wb") as f:
            pickle.dump(d, f)
    return base


def test_gpu_augment_loader_matches_host_transform_and_trains(tmp_path):
    """--gpu-augment: the loaders hand out raw uint8 batches and one kernel per batch does
    crop / flip / normalise on the device.  The eval transform must equal the host pipeline
    (ToTensor + Normalize, train_ddp.py:98-101) and the entry point must train with it."""
    from dpt_b200.cli import parse_args
    from dpt_b200.data import CIFAR10Files, get_dataloaders
    from dpt_b200.train_ddp import main
    from dpt_b200.trainer import _device_batch
    _fake_cifar(str(tmp_path / "data"))
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        os.environ.pop(k, None)
    args = parse_args(["--data-dir", str(tmp_path / "data"), "--gpu-augment", "--batch-size", "16",
                       "--workers", "0"])
    train_loader, val_loader, _ = get_dataloaders(args, 0, 1)
    host = CIFAR10Files(str(tmp_path / "data"), train=False, augment=False)
    images, labels = next(iter(val_loader))
    assert images.dtype == torch.uint8
    dev = torch.device("cuda")
    out = _device_batch(val_loader, images, dev)
    ref = torch.stack([host[i][0] for i in range(16)]).to(dev)
    assert out.dtype == torch.float32 and torch.allclose(out, ref, atol=1e-5)
    timg, _ = next(iter(train_loader))
    aug = _device_batch(train_loader, timg, dev)
    assert aug.shape == (16, 3, 32, 32) and torch.isfinite(aug).all()
    out_dir = str(tmp_path / "exp")
    main(["--data-dir", str(tmp_path / "data"), "--gpu-augment", "--epochs", "1", "--batch-size", "16",
          "--workers", "0", "--amp", "--print-freq", "5", "--output-dir", out_dir, "--steps", "6"])
    lines = open(os.path.join(out_dir, "metrics_rank0.csv")).read().splitlines()
    assert len(lines) == 2 and lines[0].startswith("epoch,")
    vals = [float(v) for v in lines[1].split(",")]
    assert all(np.isfinite(vals))
