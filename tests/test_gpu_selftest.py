"""Device sanity for the hand-built UMMA descriptors (smem matrix descriptor for K-major and
MN-major SWIZZLE_128B operands, TMEM-sourced A operand, instruction descriptor)."""
import pytest
import torch

from compactor_vllm import _native as N

pytestmark = pytest.mark.gpu


def _run(mode, a, b, lbo, sbo):
    d = torch.zeros((128, 128), dtype=torch.float32, device="cuda")
    rc = N.lib().cvllm_selftest_umma(N.ptr(a), N.ptr(b), N.ptr(d), mode, lbo, sbo, N.dtype_code(a.dtype),
                                     N.stream(a.device))
    N.check(rc, "cvllm_selftest_umma")
    torch.cuda.synchronize()
    return d


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_umma_ss_kmajor(dtype):
    g = torch.Generator().manual_seed(0)
    a = torch.randn((128, 128), generator=g).to(dtype).cuda()
    b = torch.randn((128, 128), generator=g).to(dtype).cuda()
    d = _run(0, a, b, 16, 1024)
    ref = a.float() @ b.float().T
    assert torch.allclose(d, ref, atol=1e-2, rtol=1e-3), (d - ref).abs().max().item()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_umma_ts_mnmajor(dtype):
    g = torch.Generator().manual_seed(1)
    a = torch.randn((128, 128), generator=g).to(dtype).cuda()
    b = torch.randn((128, 128), generator=g).to(dtype).cuda()  # [k, n], n contiguous
    d = _run(1, a, b, 16384, 1024)
    ref = a.float() @ b.float()
    assert torch.allclose(d, ref, atol=1e-2, rtol=1e-3), (d - ref).abs().max().item()
