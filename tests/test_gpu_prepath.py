"""GPU parity of the step in front of the hot path: `cvllm_rope_qk` (RoPE, optional
per-head RMSNorm, optional Compactor sketch) and `cvllm_rmsnorm`, through the reference-shaped layers.

Bars: the rotation is bit-exact against the oracle and the reference's golden vectors (same fp32
rounding points, no FMA contraction); the norm differs only in the order of the sum of squares, so
it is held to >= 99.9 % of the elements identical and the rest within 1 ulp of the 16-bit dtype
(2 ulp behind a weight multiply); the sketch is a
48-wide fp32-accumulated contraction: 2e-2 relative; leverage scores from the fused sketch are
bit-identical to the ones derived from the keys."""
import pytest
import torch

from conftest import load_golden
from helpers import fused_qkv

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None


def _bits(t):
    return t.contiguous().view(torch.int16)


def _rotated_close(got, want, min_exact=0.99):
    """Rotated outputs of NORMED inputs: a 1-ulp difference of a normed element (sum order of the
    variance) moves both outputs of its rotation pair by up to 1 ulp OF THE PAIR'S MAGNITUDE, which is
    many ulps of a small output.  Bar: >= 99 % identical, the rest within 2 ulp of the pair norm."""
    got, want = got.cpu(), want.cpu()
    same = _bits(got) == _bits(want)
    assert same.float().mean().item() >= min_exact, same.float().mean().item()
    w = want.float()
    pair = (w[..., :64] ** 2 + w[..., 64:] ** 2).sqrt()
    pair = torch.cat((pair, pair), -1)
    eps = 2.0 ** -7 if got.dtype == torch.bfloat16 else 2.0 ** -10
    assert ((got.float() - w).abs() <= 2 * eps * pair + 1e-6).all()


def _ulp_close(a, b, min_exact=0.999, max_ulp=1):
    """a, b 16-bit tensors: identical in >= min_exact of the elements, never more than 1 ulp apart."""
    a, b = a.cpu(), b.cpu()
    ia, ib = _bits(a).int(), _bits(b).int()
    same = (ia == ib)
    assert same.float().mean().item() >= min_exact, same.float().mean().item()
    assert ((ia - ib).abs() <= max_ulp)[~same].all()


@pytest.mark.parametrize("name", ["rope_bf16_llama3", "rope_f16_plain"])
def test_rope_golden_bit_exact(name):
    from compactor_vllm.layers.rotary_embedding import fused_qk_norm_rope

    g = load_golden(name)
    q, k, pos, cache = (g[x].to(DEV) for x in ("q", "k", "positions", "cos_sin_cache"))
    qo, ko, sk = fused_qk_norm_rope(pos, q, k, cache)
    assert sk is None
    assert torch.equal(_bits(qo.cpu()), _bits(g["q_out"]))
    assert torch.equal(_bits(ko.cpu()), _bits(g["k_out"]))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("N,HQ,HKV", [(1, 32, 8), (53, 32, 8), (300, 8, 1), (17, 64, 8), (33, 4, 4)])
def test_rope_module_strided_and_inplace(dtype, N, HQ, HKV):
    from oracle import prepath as op
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding, fused_qk_norm_rope

    D, max_pos = 128, 2048
    rope = RotaryEmbedding(D, D, max_pos, 500000.0, ("llama3", 8.0, 1.0, 4.0, 512)).to(DEV)
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=N)
    pos = torch.randint(0, max_pos, (N,), generator=torch.Generator().manual_seed(3)).to(DEV)
    want_q, want_k, _ = op.fused_qk_norm_rope(pos.cpu(), q.cpu(), k.cpu(), rope.cos_sin_cache.cpu().view(max_pos, D))
    got_q, got_k = rope(pos, q, k)  # strided views of the fused qkv buffer in, fresh tensors out
    assert got_q.shape == q.shape and got_k.shape == k.shape
    assert torch.equal(_bits(got_q.cpu()), _bits(want_q)) and torch.equal(_bits(got_k.cpu()), _bits(want_k))
    # flattened [N, H*D] inputs (the shape qkv.split hands over) and in-place rotation
    q2, k2 = q.clone().view(N, HQ * D), k.clone().view(N, HKV * D)
    r_q, r_k, _ = fused_qk_norm_rope(pos, q2, k2, rope.cos_sin_cache, inplace=True)
    assert r_q.data_ptr() == q2.data_ptr()
    assert torch.equal(_bits(q2.cpu().view(N, HQ, D)), _bits(want_q))
    assert torch.equal(_bits(k2.cpu().view(N, HKV, D)), _bits(want_k))


def test_rope_empty_and_int32_positions():
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    rope = RotaryEmbedding(128, 128, 64, 10000.0, None).to(DEV)
    q = torch.empty((0, 8, 128), dtype=torch.bfloat16, device=DEV)
    k = torch.empty((0, 2, 128), dtype=torch.bfloat16, device=DEV)
    qo, ko = rope(torch.empty((0,), dtype=torch.int64, device=DEV), q, k)
    assert qo.shape == q.shape and ko.shape == k.shape
    q = torch.randn((5, 8, 128), device=DEV).bfloat16()
    k = torch.randn((5, 2, 128), device=DEV).bfloat16()
    p64 = torch.arange(5, device=DEV)
    a = rope(p64, q, k)
    b = rope(p64.int(), q, k)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # position 0 is the identity rotation
    z = rope(torch.zeros(5, dtype=torch.int64, device=DEV), q, k)
    assert torch.equal(z[0], q) and torch.equal(z[1], k)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("N", [16, 61, 1000])
def test_fused_norm_rope_sketch_vs_oracle(dtype, N):
    from oracle import prepath as op
    from compactor_vllm.layers.layernorm import RMSNorm
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    HQ, HKV, D, max_pos = 32, 8, 128, 4096
    g = torch.Generator().manual_seed(N)
    rope = RotaryEmbedding(D, D, max_pos, 1000000.0, None).to(DEV)
    qn, kn = RMSNorm(D, eps=1e-6), RMSNorm(D, eps=1e-6)
    with torch.no_grad():
        qn.weight.copy_(1 + 0.3 * torch.randn(D, generator=g))
        kn.weight.copy_(1 + 0.3 * torch.randn(D, generator=g))
    qn, kn = qn.to(dtype).to(DEV), kn.to(dtype).to(DEV)
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=7)
    q, k = q * 3, k * 0.5  # scale matters for a norm
    pos = torch.randint(0, max_pos, (N,), generator=g).to(DEV)
    phi = (torch.randn((D, 48), generator=g) / 48 ** 0.5).to(dtype).to(DEV)
    got_q, got_k, got_sk = rope.forward_fused(pos, q, k, q_norm=qn, k_norm=kn, PHI=phi)
    want_q, want_k, want_sk = op.fused_qk_norm_rope(
        pos.cpu(), q.cpu(), k.cpu(), rope.cos_sin_cache.cpu().view(max_pos, D),
        qn.weight.detach().cpu(), kn.weight.detach().cpu(), 1e-6, phi.cpu())
    _rotated_close(got_q, want_q)
    _rotated_close(got_k, want_k)
    # position 0 is the identity rotation: the norm itself is within 1 ulp of the oracle
    zero = torch.zeros_like(pos)
    n_q, n_k, _ = rope.forward_fused(zero, q, k, q_norm=qn, k_norm=kn)
    # (two roundings: a 1-ulp flip of round16(x * rs) times a weight of up to ~2 can land 2 ulp away)
    _ulp_close(n_q, op.rms_norm(q.cpu(), qn.weight.detach().cpu(), 1e-6), max_ulp=2)
    _ulp_close(n_k, op.rms_norm(k.cpu(), kn.weight.detach().cpu(), 1e-6), max_ulp=2)
    err = (got_sk.float().cpu() - want_sk.float()).abs()
    assert (err <= 2e-2 * want_sk.float().abs().clamp_min(1.0)).all()
    # only k normed (q_norm None) and norm without sketch take their own template instances
    only_k = rope.forward_fused(pos, q, k, k_norm=kn)
    assert only_k[2] is None
    _rotated_close(only_k[1], want_k)
    plain_q, _, _ = op.fused_qk_norm_rope(pos.cpu(), q.cpu(), k.cpu(), rope.cos_sin_cache.cpu().view(max_pos, D))
    assert torch.equal(_bits(only_k[0].cpu()), _bits(plain_q))


@pytest.mark.parametrize("lens", [[700], [512, 1, 130], [16]])
def test_leverage_from_fused_sketch_is_bit_identical(lens):
    """K route (the kernel derives X = K.PHI itself) == sketch route (X written by cvllm_rope_qk)."""
    from compactor_vllm.compression.compactor import approximate_leverage_scores
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    dtype, HQ, HKV, D = torch.bfloat16, 8, 8, 128
    N = sum(lens)
    g = torch.Generator().manual_seed(11)
    rope = RotaryEmbedding(D, D, 2048, 500000.0, None).to(DEV)
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=5)
    phi = (torch.randn((D, 48), generator=g) / 48 ** 0.5).to(dtype).to(DEV)
    pos = torch.cat([torch.arange(n) for n in lens]).to(DEV)
    _, _, sk = rope.forward_fused(pos, q, k, PHI=phi)
    from_k = approximate_leverage_scores(k, lens, phi, normalize=True, chunk_size=512)
    from_sk = approximate_leverage_scores(None, lens, phi, normalize=True, chunk_size=512, sketch_rows=sk)
    a, b = _bits(from_k.cpu()), _bits(from_sk.cpu())
    nan = torch.isnan(from_k.cpu().float())  # a 1-token chunk z-scores to NaN in the reference too
    assert torch.equal(nan, torch.isnan(from_sk.cpu().float()))
    assert torch.equal(a[~nan], b[~nan])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_sketch_bits_equal_the_reference_gemm(dtype):
    """The Compactor sketch X = K.PHI is rounded to 16 bit and feeds an ill-conditioned factorisation
    (DESIGN 3.4), so its BITS matter: the kernel accumulates in the natural k order of a library GEMM
    (mma.cuh FragNat128) and must reproduce the reference's own call — torch.matmul on the transposed
    view, compression/compactor.py:158 — at the headline shape.  Tolerates a handful of flipped ulps in case
    another cuBLAS build picks a split-K kernel (measured here: 0 of 12.6 M)."""
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    N, HQ, HKV, D = 32768, 8, 8, 128
    g = torch.Generator(device=DEV).manual_seed(3)
    basis = torch.randn((HKV, 12, D), generator=g, device=DEV)
    coef = torch.randn((N, HKV, 12), generator=g, device=DEV) * torch.linspace(2.0, 0.2, 12, device=DEV)
    k = (torch.einsum("nhr,hrd->nhd", coef, basis) + 0.3 * torch.randn((N, HKV, D), generator=g, device=DEV))
    k = k.to(dtype).contiguous()
    q = torch.randn((N, HQ, D), generator=g, device=DEV).to(dtype)
    phi = (torch.randn((D, 48), generator=g, device=DEV) / 48 ** 0.5).to(dtype)
    rope = RotaryEmbedding(D, D, 8, 500000.0, None).to(DEV)
    _, _, sk = rope.forward_fused(torch.zeros(N, dtype=torch.int64, device=DEV), q, k, PHI=phi)
    ref = torch.matmul(k.transpose(0, 1), phi).transpose(0, 1)  # [N, HKV, 48]
    differ = int((_bits(sk) != _bits(ref)).sum())
    assert differ <= 1e-5 * ref.numel(), f"{differ} of {ref.numel()} sketch entries differ from cuBLAS"


@pytest.mark.parametrize("name", ["rmsnorm_bf16_128", "rmsnorm_bf16_4096", "rmsnorm_f16_1024"])
def test_rmsnorm_golden(name):
    from compactor_vllm.layers.layernorm import RMSNorm

    g = load_golden(name)
    width = g["x"].shape[-1]
    norm = RMSNorm(width, eps=g["eps"])
    with torch.no_grad():
        norm.weight.copy_(g["weight"].float())
    norm = norm.to(g["x"].dtype).to(DEV)
    y = norm(g["x"].to(DEV))
    _ulp_close(y, g["y"])
    res = g["residual"].to(DEV).clone()
    y2, r2 = norm(g["x"].to(DEV), res)
    _ulp_close(y2, g["y_add"])
    assert torch.equal(_bits(r2.cpu()), _bits(g["residual_out"]))  # fp32 add, one rounding: exact


def test_rmsnorm_per_head_strided_rows():
    """q_norm(q.view(-1, HQ, D)) on a strided q (models/qwen3.py:90)."""
    from oracle import prepath as op
    from compactor_vllm.layers.layernorm import RMSNorm

    q, _, _ = fused_qkv(40, 8, 2, 128, torch.bfloat16, DEV, seed=2)
    norm = RMSNorm(128).to(torch.bfloat16).to(DEV)
    y = norm(q)
    assert y.shape == q.shape
    _ulp_close(y, op.rms_norm(q.cpu(), norm.weight.detach().cpu(), 1e-6))


def test_rope_full_size_32k_vs_torch_on_device():
    """BASELINE size (32768 tokens, HQ=32, HKV=8): the same eager torch formula evaluated on the GPU
    (separately rounded products, like the oracle) must agree bit for bit; rotation preserves the
    norm of every (d, d+64) pair up to the 16-bit rounding."""
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    N, HQ, HKV, D = 32768, 32, 8, 128
    rope = RotaryEmbedding(D, D, 40960, 500000.0, ("llama3", 8.0, 1.0, 4.0, 8192)).to(DEV)
    q, k, _ = fused_qkv(N, HQ, HKV, D, torch.bfloat16, DEV, seed=9)
    pos = torch.arange(N, device=DEV)
    got_q, got_k = rope(pos, q, k)
    cs = rope.cos_sin_cache[pos]
    c, s = cs[..., :64], cs[..., 64:]
    for x, got in ((q, got_q), (k, got_k)):
        xf = x.float()
        x1, x2 = xf[..., :64], xf[..., 64:]
        want = torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), -1).to(x.dtype)
        assert torch.equal(got, want)
        n_in = (x1 ** 2 + x2 ** 2)
        gf = got.float()
        n_out = gf[..., :64] ** 2 + gf[..., 64:] ** 2
        assert ((n_out - n_in).abs() <= 2e-2 * n_in + 1e-3).all()
