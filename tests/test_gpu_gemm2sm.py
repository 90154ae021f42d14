"""Two-SM (cta_group::2) GEMM vs a PyTorch fp32 reference.  The kernel is not on a model path yet;
it passed on B200 in round 2 (8 cases) and has been part of the default suite since.  A wrong
barrier count in this kind of kernel is a hang, not a wrong number: bounded by a hard timeout."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120, method="thread")]


@pytest.mark.parametrize("mnk", [(256, 256, 64), (512, 512, 256), (1000, 700, 328), (4096, 4096, 1024)])
@pytest.mark.parametrize("out_f32", [False, True])
def test_gemm2sm_matches_fp32(mnk, out_f32):
    from dpt_b200.ops import _lib
    lib = _lib.lib()
    M, N, K = mnk
    g = torch.Generator(device="cuda").manual_seed(1)
    A = torch.randn((M, K), device="cuda", generator=g).to(torch.bfloat16)
    B = (torch.randn((N, K), device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    C = torch.full((M, N), float("nan"), device="cuda",
                   dtype=torch.float32 if out_f32 else torch.bfloat16)
    rc = lib.dpt_gemm2sm_bf16(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, K, K, N,
                              int(out_f32), 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t()
    err = (C.float() - ref).abs().max().item()
    assert err / (ref.abs().max().item() + 1e-6) < (2e-3 if out_f32 else 1e-2)
