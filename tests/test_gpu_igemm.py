"""tcgen05 implicit-GEMM conv / GEMM kernels vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _F():
    from dpt_b200.ops import functional as Fn
    return Fn


def _mk(shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(torch.bfloat16)


def _ref_conv(x_nhwc, w_krsc, stride, pad):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    w = w_krsc.float().permute(0, 3, 1, 2)
    return F.conv2d(x, w, stride=stride, padding=pad)  # NCHW fp32


def _close(got, ref, tol, what):
    err = (got.float() - ref).abs().max().item()
    den = ref.abs().max().item() + 1e-6
    assert err / den < tol, f"{what}: max abs err {err:.4g} vs ref max {den:.4g}"


# (N, H, W, Cin, Cout, k, stride, pad)
CONV_CASES = [
    (8, 8, 8, 64, 64, 3, 1, 1),       # layer1
    (8, 8, 8, 64, 128, 3, 2, 1),      # layer2.0.conv1
    (8, 8, 8, 64, 128, 1, 2, 0),      # layer2.0.downsample
    (16, 4, 4, 128, 128, 3, 1, 1),    # layer2
    (32, 2, 2, 256, 256, 3, 1, 1),    # layer3
    (128, 1, 1, 512, 512, 3, 1, 1),   # layer4 (only the centre tap is live)
    (4, 14, 14, 72, 96, 3, 1, 1),     # ragged tiles, Cin/Cout not multiples of 64
    (2, 56, 56, 64, 64, 1, 1, 0),     # resnet50 1x1
    (3, 16, 16, 64, 256, 3, 2, 1),    # batch tail (N not a multiple of the box)
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fprop(case):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    x = _mk((N, H, W, Cin), 1.0, 1)
    w = _mk((Cout, k, k, Cin), 0.05, 2)
    ssum = torch.zeros(Cout, device="cuda")
    ssq = torch.zeros(Cout, device="cuda")
    y = Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))
    torch.cuda.synchronize()
    ref = _ref_conv(x, w, s, p)
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, "fprop")
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stat sum")
    _close(ssq, (ref * ref).sum(dim=(0, 2, 3)), 5e-3, "stat sq")


@pytest.mark.parametrize("case", CONV_CASES[:6])
def test_conv_fprop_splitk(case):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    x = _mk((N, H, W, Cin), 1.0, 3)
    w = _mk((Cout, k, k, Cin), 0.05, 4)
    y = Fn.conv_fprop(x, w, s, p, splits=4)
    torch.cuda.synchronize()
    assert y.dtype == torch.float32
    _close(y.permute(0, 3, 1, 2), _ref_conv(x, w, s, p), 2e-3, "fprop split-k")


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_dgrad(case):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    if s == 2 and (H % 2 or W % 2):
        pytest.skip("odd stride-2 input")
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 5)
    w = _mk((Cout, k, k, Cin), 0.05, 6)
    dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                     dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
    _close(dx.permute(0, 3, 1, 2), ref, 1e-2, "dgrad")


def test_conv_dgrad_fused_downsample():
    """3x3/2 conv and the 1x1/2 downsample that read the same block input, one kernel."""
    Fn = _F()
    N, H, W, Cin, Cout = 8, 8, 8, 64, 128
    dy = _mk((N, 4, 4, Cout), 1.0, 7)
    dy2 = _mk((N, 4, 4, Cout), 1.0, 8)
    w = _mk((Cout, 3, 3, Cin), 0.05, 9)
    w2 = _mk((Cout, 1, 1, Cin), 0.05, 10)
    dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), 2, 1, dy2=dy2, w2=w2, pad2=0)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                     dy.float().permute(0, 3, 1, 2), stride=2, padding=1)
    ref = ref + torch.nn.grad.conv2d_input((N, Cin, H, W), w2.float().permute(0, 3, 1, 2),
                                           dy2.float().permute(0, 3, 1, 2), stride=2, padding=0)
    _close(dx.permute(0, 3, 1, 2), ref, 1e-2, "fused dgrad")


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_wgrad(case):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 11)
    x = _mk((N, H, W, Cin), 1.0, 12)
    dw = Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (Cout, Cin, k, k),
                                      dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
    _close(dw.permute(0, 3, 1, 2), ref, 2e-3, "wgrad")


@pytest.mark.parametrize("mnk", [(256, 64, 147), (300, 200, 520), (128, 1000, 2048), (4096, 64, 152)])
def test_gemm_ops(mnk):
    Fn = _F()
    M, N, K = mnk
    ld = lambda v: (v + 7) // 8 * 8  # noqa: E731
    A = _mk((M, ld(K)), 1.0, 13)
    B = _mk((N, ld(K)), 0.05, 14)
    C = torch.empty((M, N), device="cuda", dtype=torch.float32)
    Fn.gemm(0, A, B, C, M, N, K, ld(K), ld(K), N, out_mode=1)
    torch.cuda.synchronize()
    _close(C, A[:, :K].float() @ B[:, :K].float().t(), 2e-3, "gemm NT")

    B1 = _mk((K, ld(N)), 0.05, 15)
    C1 = torch.empty((M, N), device="cuda", dtype=torch.float32)
    Fn.gemm(1, A, B1, C1, M, N, K, ld(K), ld(N), N, out_mode=1)
    torch.cuda.synchronize()
    _close(C1, A[:, :K].float() @ B1[:, :N].float(), 2e-3, "gemm NN")

    A2 = _mk((K, ld(M)), 1.0, 16)
    C2 = torch.zeros((M, N), device="cuda", dtype=torch.float32)
    Fn.gemm(2, A2, B1, C2, M, N, K, ld(M), ld(N), N, out_mode=2, splits=3)
    torch.cuda.synchronize()
    _close(C2, A2[:, :M].float().t() @ B1[:, :N].float(), 2e-3, "gemm TN split")


@pytest.mark.parametrize("case", CONV_CASES[:6])
def test_conv_dgrad_splitk(case):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 21)
    w = _mk((Cout, k, k, Cin), 0.05, 22)
    dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p, splits=3)
    torch.cuda.synchronize()
    assert dx.dtype == torch.float32
    ref = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                     dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
    _close(dx.permute(0, 3, 1, 2), ref, 2e-3, "dgrad split-k")


# ---------------------------------------------------------------------------------------------
# Persistent variant (CTAs loop over output tiles, double-buffered TMEM accumulator).  Production
# uses it from 4 waves of tiles upwards; here every case is pushed through it with a grid of only
# 3 CTAs so each CTA walks several tiles (accumulator hand-over, barrier phases across tiles,
# producer running ahead over tile boundaries, empty dgrad parity classes).
@pytest.fixture
def force_persistent():
    from dpt_b200.ops import _lib
    lib = _lib.lib()
    lib.dpt_igemm_set_persist(1, 3)
    yield
    lib.dpt_igemm_set_persist(0, 0)


@pytest.mark.parametrize("case", CONV_CASES)
def test_persistent_conv_all_passes(case, force_persistent):
    Fn = _F()
    N, H, W, Cin, Cout, k, s, p = case
    x = _mk((N, H, W, Cin), 1.0, 31)
    w = _mk((Cout, k, k, Cin), 0.05, 32)
    ssum = torch.zeros(Cout, device="cuda")
    ssq = torch.zeros(Cout, device="cuda")
    y = Fn.conv_fprop(x, w, s, p, stats=(ssum, ssq))
    ref = _ref_conv(x, w, s, p)
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, "persistent fprop")
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "persistent stat sum")
    _close(ssq, (ref * ref).sum(dim=(0, 2, 3)), 5e-3, "persistent stat sq")
    y4 = Fn.conv_fprop(x, w, s, p, splits=3)
    torch.cuda.synchronize()
    _close(y4.permute(0, 3, 1, 2), ref, 2e-3, "persistent fprop split-k")
    P, Q = Fn.conv_out_hw(H, W, k, k, s, p)
    dy = _mk((N, P, Q, Cout), 1.0, 33)
    dw = Fn.conv_wgrad(dy, x, (Cout, k, k, Cin), s, p)
    torch.cuda.synchronize()
    refw = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (Cout, Cin, k, k),
                                       dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
    _close(dw.permute(0, 3, 1, 2), refw, 2e-3, "persistent wgrad")
    if not (s == 2 and (H % 2 or W % 2)):
        dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), s, p)
        torch.cuda.synchronize()
        refx = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                          dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
        _close(dx.permute(0, 3, 1, 2), refx, 1e-2, "persistent dgrad")


def test_persistent_gemm_many_tiles(force_persistent):
    Fn = _F()
    M, N, K = 1500, 520, 264
    A = _mk((M, K), 1.0, 41)
    B = _mk((N, K), 0.05, 42)
    C = torch.empty((M, N), device="cuda", dtype=torch.float32)
    Fn.gemm(0, A, B, C, M, N, K, K, K, N, out_mode=1)      # 12 x 5 tiles on 3 CTAs
    torch.cuda.synchronize()
    _close(C, A.float() @ B.float().t(), 2e-3, "persistent gemm NT")


def test_persistent_default_policy_large_grid():
    """Default policy: >= 592 tiles switches to the persistent kernel (2 CTAs per SM)."""
    Fn = _F()
    N, H, W, Cin, Cout = 32, 56, 56, 64, 256          # 784 row tiles x 2 column tiles
    x = _mk((N, H, W, Cin), 1.0, 51)
    w = _mk((Cout, 1, 1, Cin), 0.05, 52)
    ssum = torch.zeros(Cout, device="cuda")
    ssq = torch.zeros(Cout, device="cuda")
    y = Fn.conv_fprop(x, w, 1, 0, stats=(ssum, ssq))
    torch.cuda.synchronize()
    ref = _ref_conv(x, w, 1, 0)
    _close(y.permute(0, 3, 1, 2), ref, 1e-2, "fprop large")
    _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stat sum large")
    dy = _mk((N, H, W, Cout), 1.0, 53)
    dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), 1, 0)
    torch.cuda.synchronize()
    refx = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                      dy.float().permute(0, 3, 1, 2), stride=1, padding=0)
    _close(dx.permute(0, 3, 1, 2), refx, 1e-2, "dgrad large")


# ---------------------------------------------------------------------------------------------
# 128 x 256 tiles (DPT_IGEMM_BN256 / dpt_igemm_set_bn256; not the default tile shape): validated on
# B200 in round 2 and part of the default suite since.  A wrong barrier count in such a kernel
# is a hang, not a wrong number: bounded by a hard timeout.
@pytest.mark.timeout(120, method="thread")
@pytest.mark.parametrize("persistent", [False, True])
def test_block_n_256_all_passes(persistent):
    from dpt_b200.ops import _lib
    lib = _lib.lib()
    Fn = _F()
    lib.dpt_igemm_set_bn256(1)
    lib.dpt_igemm_set_persist(1 if persistent else (1 << 40), 7 if persistent else 0)
    try:
        N, H, W, Cin, Cout = 64, 28, 28, 256, 512            # 392 row tiles x 2 (or 1) column tiles
        x = _mk((N, H, W, Cin), 1.0, 61)
        w = _mk((Cout, 1, 1, Cin), 0.05, 62)
        ssum = torch.zeros(Cout, device="cuda")
        ssq = torch.zeros(Cout, device="cuda")
        y = Fn.conv_fprop(x, w, 1, 0, stats=(ssum, ssq))
        torch.cuda.synchronize()
        ref = _ref_conv(x, w, 1, 0)
        _close(y.permute(0, 3, 1, 2), ref, 1e-2, "fprop bn256")
        _close(ssum, ref.sum(dim=(0, 2, 3)), 5e-3, "stat sum bn256")
        _close(ssq, (ref * ref).sum(dim=(0, 2, 3)), 5e-3, "stat sq bn256")
        dy = _mk((N, H, W, Cout), 1.0, 63)
        dx = Fn.conv_dgrad(dy, w, (N, H, W, Cin), 1, 0)
        dw = Fn.conv_wgrad(dy, x, (Cout, 1, 1, Cin), 1, 0)
        torch.cuda.synchronize()
        refx = torch.nn.grad.conv2d_input((N, Cin, H, W), w.float().permute(0, 3, 1, 2),
                                          dy.float().permute(0, 3, 1, 2), stride=1, padding=0)
        refw = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (Cout, Cin, 1, 1),
                                           dy.float().permute(0, 3, 1, 2), stride=1, padding=0)
        _close(dx.permute(0, 3, 1, 2), refx, 1e-2, "dgrad bn256")
        _close(dw.permute(0, 3, 1, 2), refw, 2e-3, "wgrad bn256")
        # 3x3 with padding: the halo taps and the k loop over 9 taps x 4 chunks
        w3 = _mk((256, 3, 3, Cin), 0.05, 64)
        y3 = Fn.conv_fprop(x, w3, 1, 1)
        torch.cuda.synchronize()
        _close(y3.permute(0, 3, 1, 2), _ref_conv(x, w3, 1, 1), 1e-2, "fprop 3x3 bn256")
        M, Ng, K = 20000, 512, 264
        A = _mk((M, K), 1.0, 65)
        B = _mk((Ng, K), 0.05, 66)
        C = torch.empty((M, Ng), device="cuda", dtype=torch.float32)
        Fn.gemm(0, A, B, C, M, Ng, K, K, K, Ng, out_mode=1)
        torch.cuda.synchronize()
        _close(C, A.float() @ B.float().t(), 2e-3, "gemm NT bn256")
    finally:
        lib.dpt_igemm_set_bn256(0)
        lib.dpt_igemm_set_persist(0, 0)
