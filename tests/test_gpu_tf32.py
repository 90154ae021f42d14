"""FP32 mode of the native path: fp32 activations / weights, TF32 tensor-core math (kind::tf32) —
the same numerics the reference gets from cuDNN with allow_tf32 in its non-AMP branch
(train_ddp.py:210-214)."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_mode():
    from dpt_b200.ops import functional as Fn
    Fn.set_lowp_dtype(torch.float32)
    yield
    Fn.set_lowp_dtype(torch.bfloat16)


def _mk(shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(shape, device="cuda", generator=g) * scale


def _close(got, ref, tol, what):
    err = (got.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"{what}: max abs err {err:.4g} vs ref max {den:.4g}"


CASES = [(8, 8, 8, 64, 64, 3, 1, 1), (8, 8, 8, 64, 128, 3, 2, 1), (8, 8, 8, 64, 128, 1, 2, 0),
         (32, 2, 2, 256, 256, 3, 1, 1), (128, 1, 1, 512, 512, 3, 1, 1), (4, 14, 14, 72, 96, 3, 1, 1)]


@pytest.mark.parametrize("case", CASES)
def test_conv_tf32(case):
    from dpt_b200.ops import functional as Fn
    N, H, W, Cin, Cout, k, s, p = case
    x = _mk((N, H, W, Cin), 1.0, 1)
    w = _mk((Cout, k, k, Cin), 0.05, 2)
    ssum, ssq = torch.zeros(Cout, device="cuda"), torch.zeros(Cout, device="cuda")
    y = Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))
    assert y.dtype == torch.float32
    xr, wr = x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2)
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(xr, wr, stride=s, padding=p)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), ref, 3e-3, "fprop tf32")        # 10-bit mantissa operands
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stats")
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 3)
    dw = Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p)
    dyr = dy.permute(0, 3, 1, 2)
    torch.cuda.synchronize()
    _close(dw.permute(0, 3, 1, 2), torch.nn.grad.conv2d_weight(xr, (Cout, Cin, k, k), dyr, stride=s, padding=p), 3e-3, "wgrad tf32")
    if not (s == 2 and (H % 2 or W % 2)):
        dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)
        torch.cuda.synchronize()
        _close(dx.permute(0, 3, 1, 2), torch.nn.grad.conv2d_input((N, Cin, H, W), wr, dyr, stride=s, padding=p), 3e-3, "dgrad tf32")
    torch.backends.cudnn.allow_tf32 = True


def test_bn_and_pool_fp32():
    from dpt_b200.ops import functional as Fn
    shape = (16, 8, 8, 64)
    C = 64
    x = _mk(shape, 2.0, 1) + 0.5
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda")
    ssum, ssq = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    Fn.bn_stats(x, ssum, ssq)
    y, mean, invstd = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, relu=True)
    xr = x.permute(0, 3, 1, 2).clone().requires_grad_(True)
    yr = F.relu(F.batch_norm(xr, None, None, gamma, beta, training=True))
    dy = _mk(shape, 1.0, 5)
    dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    dx, _ = Fn.bn_bwd(dy, x, mean, invstd, gamma, dg, db, y=y)
    yr.backward(dy.permute(0, 3, 1, 2))
    torch.cuda.synchronize()
    assert y.dtype == torch.float32
    _close(y.permute(0, 3, 1, 2), yr, 1e-4, "bn fwd fp32")
    _close(dx.permute(0, 3, 1, 2), xr.grad, 1e-3, "bn dx fp32")
    yp, idx = Fn.maxpool_fwd(x)
    _close(yp.permute(0, 3, 1, 2), F.max_pool2d(x.permute(0, 3, 1, 2), 3, 2, 1), 1e-6, "maxpool fp32")


def _cos(a, b):
    a, b = a.detach().flatten().float(), b.detach().flatten().float()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-12))


def test_engine_fp32_matches_torch():
    from dpt_b200.models.resnet import build_model
    torch.manual_seed(0)
    dev = torch.device("cuda")
    model = build_model(dev, "resnet18", 10)
    ref = copy.deepcopy(model)
    ref.native = "never"
    x = torch.randn(32, 3, 32, 32, device=dev)
    t = torch.randint(0, 10, (32,), device=dev)
    torch.backends.cudnn.allow_tf32 = False         # reference side in IEEE fp32
    torch.backends.cuda.matmul.allow_tf32 = False
    out = model(x)                                  # no autocast: native fp32 / TF32 engine
    assert model._last_engine is not None and model._last_engine.dtype == torch.float32
    F.cross_entropy(out, t).backward()
    out_r = ref(x)
    F.cross_entropy(out_r, t).backward()
    torch.cuda.synchronize()
    torch.backends.cudnn.allow_tf32 = True
    assert _cos(out, out_r) > 0.9995
    # TF32 operand rounding (10-bit mantissa) through 17 layers of a random-init, batch-32 net:
    # the stem gradient is the most sensitive one (cf. the bf16 calibration in test_gpu_engine)
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        assert _cos(p.grad, q.grad) > 0.97, (n, _cos(p.grad, q.grad))


def test_trainstep_fp32_graph_learns():
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    dev = torch.device("cuda")
    torch.manual_seed(3)
    model = build_model(dev, "resnet18", 10)
    opt = SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
    shape = (32, 3, 32, 32)
    step = TrainStep(model, opt, None, batch_shape=shape, cuda_graph=True, dtype=torch.float32)
    x = torch.randn(shape).pin_memory()
    t = torch.randint(0, 10, (32,)).pin_memory()
    losses = []
    for _ in range(8):
        step(x, t)
        losses.append(step.read_loss(lag=0))
    torch.cuda.synchronize()
    assert losses[-1] < losses[0], losses
    assert torch.isfinite(opt.arena.param).all()
