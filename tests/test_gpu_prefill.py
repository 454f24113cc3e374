"""Prefill attention (tcgen05 kernel) through causal_sparse_varlen_with_cache."""
import pytest
import torch

from compactor_vllm import _native
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from oracle import attention as oa

from conftest import load_golden
from helpers import fused_qkv, make_paged, rel_err, torch_prefill_fp32

pytestmark = pytest.mark.gpu
DEV = "cuda"
ATOL = {torch.float16: 3e-3, torch.bfloat16: 2e-2}  # reference test atol (fp16) / north_star bound (bf16)
REL = {torch.float16: 2e-3, torch.bfloat16: 1e-2}


@pytest.fixture(autouse=True, params=["paired", "single"])
def kernel_variant(request):
    """Every test runs on both tcgen05 kernels: 'paired' = two query heads of a KV head per CTA
    (prefill_attn2.cu, the default for even GQA groups), 'single' = one head per CTA (prefill_attn.cu,
    the path odd groups take), forced through the library's "prefill_v1" test option."""
    _native.set_test_option("prefill_v1", 1 if request.param == "single" else 0)
    yield request.param
    _native.set_test_option("prefill_v1", 0)


def _run(q, k, v, kc, vc, lens, pt, bm, cu, HKV, PS, scale=None):
    seqlens = cu[1:] - cu[:-1]
    out = causal_sparse_varlen_with_cache(q, k, v, kc, vc, lens, pt, bm, cu, int(seqlens.max()), int(lens.max()),
                                          HKV, PS, scale)
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("name", ["prefill_bf16", "prefill_f16_ps256"])
def test_prefill_golden(name):
    g = load_golden(name)
    dtype = g["q"].dtype
    N = g["q"].shape[0]
    # rebuild the fused-buffer striding the fixture was generated with
    HQ, HKV, D = g["q"].shape[1], g["k"].shape[1], g["q"].shape[2]
    qkv = torch.cat([g["q"].reshape(N, -1), g["k"].reshape(N, -1), g["v"].reshape(N, -1)], 1).to(DEV)
    q = qkv[:, : HQ * D].view(N, HQ, D)
    k = qkv[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D)
    v = qkv[:, (HQ + HKV) * D:].view(N, HKV, D)
    out = _run(q, k, v, g["k_cache"].to(DEV), g["v_cache"].to(DEV), g["seq_lens_bh"].to(DEV),
               g["page_table"].to(DEV), g["batch_mapping"].to(DEV), g["cu_seqlens_q"].to(DEV), g["HKV"], g["PS"])
    err = (out.float().cpu() - g["out"].float()).abs().max().item()
    assert err <= ATOL[dtype], err


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("appends,HQ,HKV,PS", [([1], 8, 2, 128), ([128], 8, 2, 128), ([129, 7], 8, 2, 128),
                                              ([300, 1, 255], 32, 8, 128), ([257], 8, 1, 256)])
def test_prefill_small_vs_cpu_oracle(dtype, appends, HQ, HKV, PS):
    """Head-sparse prefixes (0 .. 2.5 pages, ragged per head), scrambled pages, non-identity slots,
    NaN-poisoned unused cache rows, q/k/v strided views of one fused buffer."""
    B, D = len(appends), 128
    g = torch.Generator().manual_seed(sum(appends) + HQ)
    lens = torch.randint(0, int(2.5 * PS), (B, HKV), generator=g)
    lens[0, 0] = 0
    lens[-1, -1] = PS
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=21, poison=True)
    cu = torch.tensor([0] + torch.tensor(appends).cumsum(0).tolist(), dtype=torch.int32)
    q, k, v = fused_qkv(int(cu[-1]), HQ, HKV, D, dtype, DEV, seed=5)
    lens_d = lens.to(torch.int32).to(DEV)
    out = _run(q, k, v, kc, vc, lens_d, pt, bm, cu.to(DEV), HKV, PS, scale=0.1)
    kcc, vcc = torch.nan_to_num(kc.cpu()), torch.nan_to_num(vc.cpu())  # oracle only gathers valid rows anyway
    ref = oa.prefill_attention(q.cpu(), k.cpu(), v.cpu(), kcc, vcc, lens, pt.cpu(), bm.cpu(), cu, HKV, PS, 0.1)
    assert torch.isfinite(out).all()
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= ATOL[dtype], err
    assert rel_err(out.cpu(), ref) <= REL[dtype]


def test_device_restatement_matches_cpu_oracle():
    dtype, B, HQ, HKV, D, PS = torch.float16, 2, 8, 2, 128, 128
    lens = torch.tensor([[70, 0], [129, 300]])
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=4)
    cu = torch.tensor([0, 150, 163], dtype=torch.int32)
    q, k, v = fused_qkv(163, HQ, HKV, D, dtype, DEV, seed=6)
    a = torch_prefill_fp32(q, k, v, kc, vc, lens.to(DEV), pt, bm, cu.to(DEV), HKV, PS)
    b = oa.prefill_attention(q.cpu(), k.cpu(), v.cpu(), kc.cpu(), vc.cpu(), lens, pt.cpu(), bm.cpu(), cu, HKV, PS)
    assert (a.cpu() - b).abs().max().item() < 1e-4


# the reference's own grid (tests/test_triton_attention.py:33-50,200-287): fp16, PS=256, HQ=32, HKV=8
@pytest.mark.parametrize("B", [1, 2, 3, 8])
@pytest.mark.parametrize("cache_len", [0, 1, 70, 128, 8193])
@pytest.mark.parametrize("append_len", [1, 2, 13, 8000])
def test_prefill_reference_grid(B, cache_len, append_len):
    dtype, HQ, HKV, D, PS = torch.float16, 32, 8, 128, 256
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, [[cache_len] * HKV] * B, dtype, DEV, seed=1234)
    cu = (torch.arange(B + 1, dtype=torch.int32) * append_len)
    q, k, v = fused_qkv(B * append_len, HQ, HKV, D, dtype, DEV, seed=B + cache_len + append_len)
    lens = torch.full((B, HKV), cache_len, dtype=torch.int32, device=DEV)
    out = _run(q, k, v, kc, vc, lens, pt, bm, cu.to(DEV), HKV, PS)
    ref = torch_prefill_fp32(q, k, v, kc, vc, lens, pt, bm, cu.to(DEV), HKV, PS)
    err = (out.float() - ref).abs().max().item()
    assert err <= ATOL[dtype], err


def test_prefill_bf16_16k_single_sequence():
    """BASELINE config-2 shape (one layer): B=1, N=16384, HQ=32, HKV=8, no prefix."""
    dtype, HQ, HKV, D, PS, Ntok = torch.bfloat16, 32, 8, 128, 128, 16384
    kc, vc, pt, bm = make_paged(1, HKV, D, PS, [[0] * HKV], dtype, DEV, seed=0)
    cu = torch.tensor([0, Ntok], dtype=torch.int32, device=DEV)
    q, k, v = fused_qkv(Ntok, HQ, HKV, D, dtype, DEV, seed=1)
    lens = torch.zeros((1, HKV), dtype=torch.int32, device=DEV)
    out = _run(q, k, v, kc, vc, lens, pt, bm, cu, HKV, PS)
    ref = torch_prefill_fp32(q, k, v, kc, vc, lens, pt, bm, cu, HKV, PS)
    assert (out.float() - ref).abs().max().item() <= ATOL[dtype]
    assert rel_err(out, ref) <= REL[dtype]


@pytest.fixture
def rescale_threshold():
    yield lambda t: _native.set_test_option("prefill_threshold", t)
    _native.set_test_option("prefill_threshold", -1.0)


@pytest.mark.parametrize("threshold", [0.0, 8.0])
def test_prefill_correction_path(threshold, rescale_threshold):
    """A 1-key first tile makes later tiles raise the running max by far more than the lazy-rescale
    threshold, so the O-correction path runs; threshold 0 (test hook) forces it on every increase."""
    rescale_threshold(threshold)
    dtype, HQ, HKV, D, PS = torch.bfloat16, 8, 2, 128, 128
    lens = [[1, 129], [257, 1]]
    kc, vc, pt, bm = make_paged(2, HKV, D, PS, lens, dtype, DEV, seed=13, poison=True)
    cu = torch.tensor([0, 700, 764], dtype=torch.int32, device=DEV)
    q, k, v = fused_qkv(764, HQ, HKV, D, dtype, DEV, seed=3)
    k[::5] *= 4.0  # a few dominant keys -> large max jumps between tiles
    l = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = _run(q, k, v, kc, vc, l, pt, bm, cu, HKV, PS)
    ref = torch_prefill_fp32(q, k, v, torch.nan_to_num(kc), torch.nan_to_num(vc), l, pt, bm, cu, HKV, PS)
    assert torch.isfinite(out).all()
    assert (out.float() - ref).abs().max().item() <= ATOL[dtype]
