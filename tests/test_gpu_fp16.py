"""fp16 instantiation of the native path (the reference's literal --amp dtype: autocast defaults to
fp16 + dynamic GradScaler, train_ddp.py:204-209,346)."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp16_mode():
    from dpt_b200.ops import functional as Fn
    Fn.set_lowp_dtype(torch.float16)
    yield
    Fn.set_lowp_dtype(torch.bfloat16)


def _mk(shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(torch.float16)


def _close(got, ref, tol, what):
    err = (got.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"{what}: max abs err {err:.4g} vs ref max {den:.4g}"


@pytest.mark.parametrize("case", [(8, 8, 8, 64, 64, 3, 1, 1), (8, 8, 8, 64, 128, 3, 2, 1), (32, 2, 2, 256, 256, 3, 1, 1)])
def test_conv_fp16(case):
    from dpt_b200.ops import functional as Fn
    N, H, W, Cin, Cout, k, s, p = case
    x = _mk((N, H, W, Cin), 1.0, 1)
    w = _mk((Cout, k, k, Cin), 0.05, 2)
    ssum, ssq = torch.zeros(Cout, device="cuda"), torch.zeros(Cout, device="cuda")
    y = Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))
    assert y.dtype == torch.float16
    xr = x.float().permute(0, 3, 1, 2)
    wr = w.float().permute(0, 3, 1, 2)
    ref = F.conv2d(xr, wr, stride=s, padding=p)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), ref, 2e-3, "fprop fp16")       # fp16 has 3 more mantissa bits than bf16
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stats")
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 3)
    dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)
    dw = Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p)
    torch.cuda.synchronize()
    dyr = dy.float().permute(0, 3, 1, 2)
    _close(dx.permute(0, 3, 1, 2), torch.nn.grad.conv2d_input((N, Cin, H, W), wr, dyr, stride=s, padding=p), 2e-3, "dgrad fp16")
    _close(dw.permute(0, 3, 1, 2), torch.nn.grad.conv2d_weight(xr, (Cout, Cin, k, k), dyr, stride=s, padding=p), 2e-3, "wgrad fp16")


def test_bn_fp16():
    from dpt_b200.ops import functional as Fn
    shape = (16, 8, 8, 64)
    C = 64
    x = (_mk(shape, 1.0, 1).float() * 2 + 0.5).half()
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda")
    ssum, ssq = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    Fn.bn_stats(x, ssum, ssq)
    y, mean, invstd = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, relu=True)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.relu(F.batch_norm(xr, None, None, gamma, beta, training=True))
    dy = _mk(shape, 1.0, 5)
    dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    dx, _ = Fn.bn_bwd(dy, x, mean, invstd, gamma, dg, db, y=y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    torch.cuda.synchronize()
    assert y.dtype == torch.float16
    _close(y.permute(0, 3, 1, 2), yr, 3e-3, "bn fwd fp16")
    _close(dx.permute(0, 3, 1, 2), xr.grad, 1e-2, "bn dx fp16")


def _cos(a, b):
    a, b = a.detach().flatten().float(), b.detach().flatten().float()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-12))


def test_engine_fp16_matches_torch():
    from dpt_b200.models.resnet import build_model
    torch.manual_seed(0)
    dev = torch.device("cuda")
    model = build_model(dev, "resnet18", 10)
    ref = copy.deepcopy(model)
    ref.native = "never"
    x = torch.randn(32, 3, 32, 32, device=dev)
    t = torch.randint(0, 10, (32,), device=dev)
    with torch.autocast("cuda", dtype=torch.float16):     # reference-style autocast: fp16
        out = model(x)
    assert model._last_engine is not None and model._last_engine.dtype == torch.float16
    loss = F.cross_entropy(out.float(), t)
    loss.backward()
    out_r = ref(x)
    loss_r = F.cross_entropy(out_r, t)
    loss_r.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - loss_r.item()) < 0.03 * max(1.0, abs(loss_r.item()))
    assert _cos(out, out_r) > 0.995
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        assert _cos(p.grad, q.grad) > 0.97, (n, _cos(p.grad, q.grad))


def test_trainstep_fp16_dynamic_loss_scaling():
    """fp16 + dynamic GradScaler, whole step in one CUDA graph, no host sync: an overflowing first
    scale is backed off on the device and training then proceeds."""
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    dev = torch.device("cuda")
    torch.manual_seed(3)
    model = build_model(dev, "resnet18", 10)
    opt = SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
    scaler = GradScaler(device=dev, init_scale=2.0 ** 24, growth_interval=1000, enabled=True)
    shape = (32, 3, 32, 32)
    step = TrainStep(model, opt, scaler, batch_shape=shape, cuda_graph=True, dtype=torch.float16)
    x = torch.randn(shape).pin_memory()
    t = torch.randint(0, 10, (32,)).pin_memory()
    p0 = opt.arena.param.clone()
    losses = []
    for _ in range(30):
        step(x, t)
        losses.append(step.read_loss(lag=0))
    torch.cuda.synchronize()
    assert scaler.get_scale() < 2.0 ** 24, "an fp16 overflow must have halved the scale"
    assert torch.isfinite(opt.arena.param).all()
    assert not torch.equal(opt.arena.param, p0)
    assert losses[-1] < losses[0], losses
    assert opt.arena.shadow.dtype == torch.float16
    assert torch.equal(opt.arena.shadow, opt.arena.param.half())
