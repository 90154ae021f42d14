"""The allocation-free BatchNorm entry point (dpt_bn_fwd_train_noalloc) against the regular
dpt_bn_stats + dpt_bn_fwd_train pair, both called through the C ABI with ctypes.

Some users of the library cannot allocate device memory at the point where they need a
BatchNorm forward: their memory was planned once at start-up, or they sit in a loop that must not
hit the device-wide synchronisation of cudaMalloc, or they are a plugin that only ever receives
pointers and a stream.  The no-alloc entry point serves them from a scratch array compiled into
the library.  This test makes two calls in a row with different shapes (the second must not see
the first one's leftovers in the shared scratch) and checks every output against the regular
path within the tolerance the other BatchNorm tests use.  This file sorts last on purpose, so a
run that stops at the first failure has already exercised everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float


def _lib():
    from dpt_b200.ops import _lib as L
    lib = L.lib()
    fn = lib.dpt_bn_fwd_train_noalloc
    fn.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp]
    fn.restype = i32
    return lib


def _close(got, ref, tol, what):
    err = (got.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"{what}: max abs err {err:.4g} vs ref max {den:.4g}"


def _one_call(lib, shape, seed, relu, with_res, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    N, H, W, Cn = shape
    M = N * H * W
    x = (torch.randn(shape, device="cuda", generator=g) * 2 + 0.5).to(dtype)
    res = torch.randn(shape, device="cuda", generator=g).to(dtype) if with_res else None
    gamma = torch.rand(Cn, device="cuda", generator=g) + 0.5
    beta = torch.randn(Cn, device="cuda", generator=g)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda t: None if t is None else t.data_ptr()  # noqa: E731

    def outputs():
        return dict(y=torch.empty_like(x), mean=torch.empty(Cn, device="cuda"),
                    invstd=torch.empty(Cn, device="cuda"), rm=torch.zeros(Cn, device="cuda"),
                    rv=torch.ones(Cn, device="cuda"),
                    nbt=torch.zeros((), device="cuda", dtype=torch.int64))

    # regular path: the caller owns (and here allocates) the sum / sum-of-squares scratch
    a = outputs()
    ssum = torch.zeros(Cn, device="cuda")
    ssq = torch.zeros(Cn, device="cuda")
    assert lib.dpt_bn_stats(p(x), p(ssum), p(ssq), M, Cn, st) == 0
    assert lib.dpt_bn_fwd_train(p(x), p(ssum), p(ssq), p(gamma), p(beta), p(res), p(a["y"]),
                                p(a["mean"]), p(a["invstd"]), p(a["rm"]), p(a["rv"]), p(a["nbt"]),
                                M, Cn, 0.1, 1e-5, int(relu), st) == 0
    # allocation-free path: same inputs and outputs, scratch lives inside the library
    b = outputs()
    rc = lib.dpt_bn_fwd_train_noalloc(p(x), p(gamma), p(beta), p(res), p(b["y"]), p(b["mean"]),
                                      p(b["invstd"]), p(b["rm"]), p(b["rv"]), p(b["nbt"]), M, Cn,
                                      0.1, 1e-5, int(relu), st)
    assert rc == 0, f"dpt_bn_fwd_train_noalloc returned {rc}"
    torch.cuda.synchronize()
    _close(b["y"], a["y"], 1.5e-2, "y")
    _close(b["mean"], a["mean"], 1e-3, "save_mean")
    _close(b["invstd"], a["invstd"], 1e-3, "save_invstd")
    _close(b["rm"], a["rm"], 1e-3, "running_mean")
    _close(b["rv"], a["rv"], 1e-3, "running_var")
    assert int(b["nbt"]) == int(a["nbt"]) == 1
    # and both agree with PyTorch's fp32 batch_norm on the same bf16-rounded input
    ref = torch.nn.functional.batch_norm(x.float().permute(0, 3, 1, 2), None, None, gamma, beta,
                                         training=True, eps=1e-5)
    if with_res:
        ref = ref + res.float().permute(0, 3, 1, 2)
    if relu:
        ref = torch.relu(ref)
    _close(b["y"].permute(0, 3, 1, 2), ref, 1.5e-2, "y vs torch fp32")


def test_noalloc_bn_matches_regular_entry_point_two_calls_in_a_row():
    lib = _lib()
    # first call: wide tensor fills 2*512 floats of the shared scratch
    _one_call(lib, (16, 4, 4, 512), seed=1, relu=True, with_res=True)
    # second call right after: fewer channels, different statistics — must not see leftovers
    _one_call(lib, (8, 8, 8, 64), seed=2, relu=False, with_res=False)


def test_noalloc_bn_follows_the_element_type_and_reaches_its_channel_limit():
    """The entry point honours dpt_set_lowp_dtype like the regular pair (fp16 is the reference's
    own autocast type, fp32 its non-AMP mode), and its documented maximum, 2048 channels, is a
    size it really accepts."""
    from dpt_b200.ops import _lib as L
    lib = _lib()
    try:
        L.set_lowp_dtype(torch.float16)
        _one_call(lib, (8, 4, 4, 128), seed=3, relu=True, with_res=True, dtype=torch.float16)
        L.set_lowp_dtype(torch.float32)
        _one_call(lib, (8, 4, 4, 64), seed=4, relu=True, with_res=False, dtype=torch.float32)
    finally:
        L.set_lowp_dtype(torch.bfloat16)
    _one_call(lib, (4, 2, 2, 2048), seed=5, relu=False, with_res=False)


def test_noalloc_bn_rejects_sizes_beyond_the_builtin_scratch():
    lib = _lib()
    Cn = 8192  # > DPT_BN_NOALLOC_MAX_CHANNELS (2048)
    x = torch.zeros((2, 1, 1, Cn), device="cuda", dtype=torch.bfloat16)
    y = torch.empty_like(x)
    v = torch.zeros(Cn, device="cuda")
    rc = lib.dpt_bn_fwd_train_noalloc(x.data_ptr(), v.data_ptr(), v.data_ptr(), None, y.data_ptr(),
                                      v.data_ptr(), v.data_ptr(), None, None, None, 2, Cn, 0.1,
                                      1e-5, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == -6  # DPT_ERR_TOO_LARGE
