"""BatchNorm / pooling / head / cross-entropy / optimizer kernels vs PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _F():
    from dpt_b200.ops import functional as Fn
    return Fn


def _rand(shape, seed, scale=1.0, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(dtype)


def _close(got, ref, tol, what):
    err = (got.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"{what}: max abs err {err:.4g} vs ref max {den:.4g}"


@pytest.mark.parametrize("shape", [(16, 8, 8, 64), (128, 1, 1, 512), (4, 7, 7, 2048), (3, 5, 5, 32)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
def test_bn_train_fwd_bwd(shape, relu, res):
    _bn_train_fwd_bwd(shape, relu, res)


# ResNet-50-sized tensors (>= 1.2M 8-channel vectors: large enough for the bulk-copy streaming
# kernels; the second shape ends in a partial 256-vector chunk and has C = 256, the case that
# exposed the missing proxy fence in bulk_stream).
@pytest.mark.parametrize("shape", [(16, 112, 112, 64), (5, 99, 101, 256)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
def test_bn_train_fwd_bwd_large(shape, relu, res):
    _bn_train_fwd_bwd(shape, relu, res)


def test_bn_statistics_with_mean_much_larger_than_std():
    """Batch statistics are accumulated as sum / sum of squares in fp32 (not Welford like
    torch): check the cancellation in E[x^2] - E[x]^2 where it matters for this framework's
    inputs — bf16 activations with |mean| = 10 std (at |mean| = 100 std bf16 itself cannot
    represent the data: its spacing there is half a standard deviation)."""
    Fn = _F()
    shape = (64, 8, 8, 64)
    C = shape[-1]
    g = torch.Generator(device="cuda").manual_seed(9)
    x = (torch.randn(shape, device="cuda", generator=g) + 10.0).to(torch.bfloat16)
    ssum = torch.zeros(C, device="cuda")
    ssq = torch.zeros(C, device="cuda")
    Fn.bn_stats(x, ssum, ssq)
    gamma = torch.ones(C, device="cuda")
    beta = torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    y, mean, invstd = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, running_mean=rm, running_var=rv)
    xf = x.float().view(-1, C)
    torch.cuda.synchronize()
    assert torch.allclose(mean, xf.mean(0), rtol=1e-5, atol=1e-4)
    var_ref = xf.var(0, unbiased=False)
    var = 1.0 / invstd.square() - 1e-5
    assert torch.allclose(var, var_ref, rtol=2e-3), ((var - var_ref).abs().max().item())
    ref = (xf - xf.mean(0)) * (var_ref + 1e-5).rsqrt()
    _close(y.view(-1, C), ref, 1.5e-2, "normalised output at mean = 10 std")


def _bn_train_fwd_bwd(shape, relu, res):
    Fn = _F()
    N, H, W, C = shape
    x = _rand(shape, 1) * 2 + 0.5
    x = x.to(torch.bfloat16)
    gamma = _rand((C,), 2, dtype=torch.float32).abs() + 0.5
    beta = _rand((C,), 3, dtype=torch.float32)
    residual = _rand(shape, 4) if res else None
    rm = torch.zeros(C, device="cuda")
    rv = torch.ones(C, device="cuda")
    nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    ssum = torch.zeros(C, device="cuda")
    ssq = torch.zeros(C, device="cuda")
    Fn.bn_stats(x, ssum, ssq)
    y, mean, invstd = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, residual=residual, relu=relu,
                                      running_mean=rm, running_var=rv, num_batches_tracked=nbt)
    # reference (NCHW fp32 on the same bf16-rounded input)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    rm_r = torch.zeros(C, device="cuda")
    rv_r = torch.ones(C, device="cuda")
    yr = F.batch_norm(xr, rm_r, rv_r, gr, br, training=True, momentum=0.1, eps=1e-5)
    if res:
        yr = yr + residual.float().permute(0, 3, 1, 2)
    if relu:
        yr = F.relu(yr)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), yr, 1.5e-2, "bn fwd")
    _close(rm, rm_r, 1e-3, "running_mean")
    _close(rv, rv_r, 1e-3, "running_var")
    assert int(nbt.item()) == 1
    # backward
    dy = _rand(shape, 5)
    dgamma = torch.zeros(C, device="cuda")
    dbeta = torch.zeros(C, device="cuda")
    dx, g = Fn.bn_bwd(dy, x, mean, invstd, gamma, dgamma, dbeta, y=y if relu else None)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    torch.cuda.synchronize()
    # the ReLU mask is taken from the bf16 output: elements whose fp32 value rounds to zero
    # differ marginally, hence the looser tolerance.  Where the pre-activation is within rounding
    # distance of zero the ReLU decision itself is a coin flip between two correct fp32
    # evaluation orders (ours: fma(x, scale, shift); torch: ((x - mean) * invstd) * w + b), and
    # the batch statistics differ in the last bit from run to run (atomic accumulation order).
    # With ~13M elements in the large shapes a few such elements exist in every run; they carry
    # the full incoming gradient on one side and zero on the other, so they are excluded.
    dx_got, dx_ref = dx.permute(0, 3, 1, 2).float(), xr.grad
    if relu:
        with torch.no_grad():
            pre = F.batch_norm(x.float().permute(0, 3, 1, 2), None, None, gamma, beta, training=True,
                               eps=1e-5)
            if res:
                pre = pre + residual.float().permute(0, 3, 1, 2)
            decided = pre.abs() > 1e-4
        assert (~decided).float().mean().item() < 1e-3
        dx_got, dx_ref = dx_got * decided, dx_ref * decided
    _close(dx_got, dx_ref, 3e-2, "bn dx")
    _close(dgamma, gr.grad, 2e-2, "dgamma")
    _close(dbeta, br.grad, 2e-2, "dbeta")


def test_bn_eval():
    Fn = _F()
    shape = (8, 4, 4, 128)
    C = shape[-1]
    x = _rand(shape, 1)
    gamma = _rand((C,), 2, dtype=torch.float32)
    beta = _rand((C,), 3, dtype=torch.float32)
    rm = _rand((C,), 4, dtype=torch.float32)
    rv = _rand((C,), 5, dtype=torch.float32).abs() + 0.1
    y = Fn.bn_fwd_eval(x, gamma, beta, rm, rv, relu=True)
    ref = F.relu(F.batch_norm(x.float().permute(0, 3, 1, 2), rm, rv, gamma, beta, training=False))
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, "bn eval")


@pytest.mark.parametrize("shape", [(8, 16, 16, 64), (2, 112, 112, 64), (3, 7, 9, 8)])
def test_maxpool(shape):
    Fn = _F()
    x = _rand(shape, 1)
    y, idx = Fn.maxpool_fwd(x)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 2, 1)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), yr, 1e-6, "maxpool fwd")
    dy = _rand(tuple(y.shape), 2)
    dx = Fn.maxpool_bwd(dy, idx, shape)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    torch.cuda.synchronize()
    # ties (equal bf16 values in a window) may pick another argmax; compare sums per window
    diff = (dx.float().permute(0, 3, 1, 2) - xr.grad).abs()
    assert (diff > 1e-2).float().mean().item() < 5e-3


@pytest.mark.parametrize("B,HW,C,K", [(128, 1, 512, 10), (16, 49, 2048, 1000), (5, 4, 64, 7)])
def test_head(B, HW, C, K):
    Fn = _F()
    side = int(HW ** 0.5)
    x = _rand((B, side, side, C), 1)
    w = _rand((K, C), 2, 0.05, torch.float32)
    b = _rand((K,), 3, 0.1, torch.float32)
    logits, pooled = Fn.head_fwd(x, w, b)
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    ref = F.linear(xr.mean(dim=(1, 2)), wr, br)
    torch.cuda.synchronize()
    _close(logits, ref, 1e-4, "head fwd")
    dl = _rand((B, K), 4, dtype=torch.float32)
    dw = torch.zeros_like(w)
    db = torch.zeros_like(b)
    dx = Fn.head_bwd(dl, w, pooled, tuple(x.shape), dw, db)
    ref.backward(dl)
    torch.cuda.synchronize()
    _close(dx, xr.grad, 1e-2, "head dx")
    _close(dw, wr.grad, 1e-4, "head dw")
    _close(db, br.grad, 1e-4, "head db")


@pytest.mark.parametrize("B,K", [(128, 10), (256, 1000), (7, 3)])
def test_cross_entropy(B, K):
    Fn = _F()
    z = _rand((B, K), 1, 3.0, torch.float32)
    t = torch.randint(0, K, (B,), device="cuda")
    scale = torch.tensor(1024.0, device="cuda")
    metrics = torch.zeros(3, device="cuda", dtype=torch.float64)
    loss, dz = Fn.cross_entropy(z, t, loss_scale=scale, metrics=metrics)
    zr = z.clone().requires_grad_(True)
    lr = F.cross_entropy(zr, t)
    (lr * 1024.0).backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - lr.item()) < 1e-4 * max(1.0, abs(lr.item()))
    _close(dz, zr.grad, 1e-4, "ce grad")
    assert int(metrics[1].item()) == int((z.argmax(1) == t).sum().item())
    assert int(metrics[2].item()) == B
    assert abs(metrics[0].item() - lr.item() * B) < 1e-3 * B


def test_stem_im2col_gemm():
    """7x7/2 RGB stem = im2col + tcgen05 GEMM, vs conv2d."""
    Fn = _F()
    N = 16
    x = torch.randn(N, 3, 32, 32, device="cuda")
    w = torch.randn(64, 3, 7, 7, device="cuda") * 0.05
    w_krsc = w.permute(0, 2, 3, 1).contiguous()  # [64,7,7,3]
    ldk = 152
    cols = Fn.im2col_nchw(x, 7, 7, 2, 3, ldk)
    wp = Fn.pad_rows_to_bf16(w_krsc.view(64, 147), 64, 147, ldk)
    M = N * 16 * 16
    y = torch.empty((M, 64), device="cuda", dtype=torch.bfloat16)
    ssum = torch.zeros(64, device="cuda")
    ssq = torch.zeros(64, device="cuda")
    Fn.gemm(0, cols, wp, y, M, 64, 147, ldk, ldk, 64, out_mode=0, stats=(ssum, ssq))
    ref = F.conv2d(x.bfloat16().float(), w.bfloat16().float(), stride=2, padding=3)
    torch.cuda.synchronize()
    _close(y.view(N, 16, 16, 64).permute(0, 3, 1, 2), ref, 1e-2, "stem")
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stem stats")
    # wgrad through the same im2col matrix
    dy = _rand((M, 64), 3)
    dw = torch.zeros((64, 147), device="cuda")
    Fn.gemm(2, dy, cols, dw, 64, 147, M, 64, ldk, 147, out_mode=2, splits=32)
    refw = torch.nn.grad.conv2d_weight(x.bfloat16().float(), (64, 3, 7, 7),
                                       dy.float().view(N, 16, 16, 64).permute(0, 3, 1, 2),
                                       stride=2, padding=3)
    torch.cuda.synchronize()
    _close(dw.view(64, 7, 7, 3).permute(0, 3, 1, 2), refw, 2e-3, "stem wgrad")


@pytest.mark.parametrize("n", [1000, 44726568 // 4 // 16, 12345])
def test_sgd_flat_matches_torch(n):
    Fn = _F()
    p0 = torch.randn(n, device="cuda")
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=5e-4)
    p = p0.clone()
    m = torch.zeros_like(p)
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for it in range(3):
        g = torch.randn(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(it))
        ref_p.grad = g.clone()
        opt.step()
        Fn.sgd_flat(p, g, m, shadow, 0.1, 0.9, 0.0, 5e-4, False)
    torch.cuda.synchronize()
    _close(p, ref_p.data, 1e-5, "sgd")
    _close(shadow, p.bfloat16(), 1e-6, "shadow")


def test_sgd_skips_on_found_inf_and_scales():
    Fn = _F()
    n = 4096
    p = torch.randn(n, device="cuda")
    p_before = p.clone()
    g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    found = torch.ones((), device="cuda")
    Fn.sgd_flat(p, g, m, None, 0.1, 0.9, 0.0, 0.0, False, found_inf=found)
    torch.cuda.synchronize()
    assert torch.equal(p, p_before)
    found.zero_()
    inv = torch.tensor(0.5, device="cuda")
    Fn.sgd_flat(p, g, m, None, 0.1, 0.0, 0.0, 0.0, False, grad_scale=inv, found_inf=found)
    torch.cuda.synchronize()
    _close(p, p_before - 0.1 * 0.5 * g, 1e-6, "scaled sgd")


def test_adam_flat_matches_torch():
    Fn = _F()
    n = 5000
    p0 = torch.randn(n, device="cuda")
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref_p], lr=1e-3, weight_decay=1e-2)
    p = p0.clone()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    step = torch.zeros((), device="cuda", dtype=torch.int32)
    for it in range(4):
        g = torch.randn(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(it))
        ref_p.grad = g.clone()
        opt.step()
        Fn.adam_flat(p, g, m, v, None, step, 1e-3, weight_decay=1e-2)
    torch.cuda.synchronize()
    _close(p, ref_p.data, 1e-5, "adam")
    assert int(step.item()) == 4


def test_amp_scale_state_machine():
    Fn = _F()
    scale = torch.tensor(65536.0, device="cuda")
    tracker = torch.zeros((), device="cuda", dtype=torch.int32)
    found = torch.ones((), device="cuda")
    inv = torch.zeros((), device="cuda")
    Fn.amp_update_scale(scale, tracker, found, inv, 2.0, 0.5, 3)
    torch.cuda.synchronize()
    assert scale.item() == 32768.0 and tracker.item() == 0 and found.item() == 0.0
    for _ in range(3):
        Fn.amp_update_scale(scale, tracker, found, inv, 2.0, 0.5, 3)
    torch.cuda.synchronize()
    assert scale.item() == 65536.0 and tracker.item() == 0
    assert abs(inv.item() - 1.0 / 65536.0) < 1e-12
    g = torch.full((1024,), 4.0, device="cuda")
    g[77] = float("inf")
    Fn.amp_unscale(g, inv, found)
    torch.cuda.synchronize()
    assert found.item() == 1.0
    assert abs(g[0].item() - 4.0 / 65536.0) < 1e-9


@pytest.mark.parametrize("shape", [(128, 4, 4, 128), (128, 1, 1, 512), (16, 2, 2, 256)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_bn_small_f32_fused_and_f32_dy_backward(shape, relu, res):
    """Split-K path: fp32 conv accumulator -> single-launch BN; fp32 incoming gradient."""
    Fn = _F()
    N, H, W, C = shape
    x32 = torch.randn(shape, device="cuda") * 2 + 0.3
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda")
    residual = _rand(shape, 4) if res else None
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    y = torch.empty(shape, device="cuda", dtype=torch.bfloat16)
    xb = torch.empty(shape, device="cuda", dtype=torch.bfloat16)
    mean, invstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    Fn.bn_fwd_train_small_f32(x32, gamma, beta, residual=residual, relu=relu, out=y, x_bf16=xb,
                              save_mean=mean, save_invstd=invstd, running_mean=rm, running_var=rv,
                              num_batches_tracked=nbt)
    xr = x32.permute(0, 3, 1, 2).clone().requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    rm_r, rv_r = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    yr = F.batch_norm(xr, rm_r, rv_r, gr, br, training=True, momentum=0.1, eps=1e-5)
    if res:
        yr = yr + residual.float().permute(0, 3, 1, 2)
    if relu:
        yr = F.relu(yr)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), yr, 1.5e-2, "y")
    _close(xb, x32, 1e-2, "bf16 copy of x")
    _close(rm, rm_r, 1e-3, "running_mean")
    _close(rv, rv_r, 1e-3, "running_var")
    assert int(nbt) == 1
    dy32 = torch.randn(shape, device="cuda")
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    dx, g = Fn.bn_bwd(dy32, xb, mean, invstd, gamma, dgamma, dbeta, y=y if relu else None)
    yr.backward(dy32.permute(0, 3, 1, 2))
    torch.cuda.synchronize()
    _close(dx.permute(0, 3, 1, 2), xr.grad, 4e-2, "dx")
    _close(dgamma, gr.grad, 3e-2, "dgamma")
    _close(dbeta, br.grad, 2e-2, "dbeta")


@pytest.mark.parametrize("shape", [(128, 16, 16, 64), (128, 4, 4, 128), (128, 1, 1, 512), (5, 3, 3, 32)])
@pytest.mark.parametrize("relu,two", [(True, True), (True, False), (False, False)])
def test_bn_bwd_fused_single_launch(shape, relu, two):
    """One-launch BN backward (grid barrier) equals the two-kernel version, incl. accumulation."""
    Fn = _F()
    C = shape[-1]
    x = (_rand(shape, 1).float() * 2 + 0.5).to(torch.bfloat16)
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda")
    ssum, ssq = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    Fn.bn_stats(x, ssum, ssq)
    y, mean, invstd = Fn.bn_fwd_train(x, ssum, ssq, gamma, beta, relu=relu)
    dy = _rand(shape, 5)
    dy2 = _rand(shape, 6) if two else None
    ref_dg, ref_db = torch.ones(C, device="cuda"), torch.ones(C, device="cuda")   # pre-existing grads
    ref_dx, ref_g = Fn.bn_bwd(dy, x, mean, invstd, gamma, ref_dg, ref_db, dy2=dy2, y=y if relu else None)
    dg, db = torch.ones(C, device="cuda"), torch.ones(C, device="cuda")
    scratch = torch.zeros(2 * C, device="cuda")
    counter = torch.zeros(1, device="cuda", dtype=torch.int32)
    g = torch.empty_like(x)
    dx = torch.empty_like(x)
    Fn.bn_bwd_fused(dy, x, mean, invstd, gamma, dg, db, scratch, counter, dy2=dy2,
                    y=y if relu else None, g_out=g, dx=dx)
    torch.cuda.synchronize()
    _close(dx, ref_dx, 4e-2, "dx")   # bf16 output, sums reduced in a different order
    _close(dg, ref_dg, 5e-3, "dgamma")
    _close(db, ref_db, 5e-3, "dbeta")


def test_gpu_augment_matches_host_transform():
    Fn = _F()
    from dpt_b200.data import CIFAR_MEAN, CIFAR_STD
    N = 16
    img = torch.randint(0, 256, (N, 3, 32, 32), dtype=torch.uint8)
    params = torch.stack([torch.randint(0, 9, (N,)), torch.randint(0, 9, (N,)),
                          torch.randint(0, 2, (N,))], dim=1).to(torch.int32)
    out = Fn.augment_crop_flip_norm(img.cuda(), params.cuda(), CIFAR_MEAN, CIFAR_STD, 4)
    mean = torch.tensor(CIFAR_MEAN).view(3, 1, 1)
    std = torch.tensor(CIFAR_STD).view(3, 1, 1)
    for n in range(N):
        dy, dx, flip = params[n].tolist()
        padded = torch.zeros(3, 40, 40, dtype=torch.uint8)
        padded[:, 4:36, 4:36] = img[n]
        crop = padded[:, dy:dy + 32, dx:dx + 32]
        if flip:
            crop = crop.flip(2)
        ref = (crop.float() / 255.0 - mean) / std
        assert torch.allclose(out[n].cpu(), ref, atol=1e-5), n
