"""Key scorers (Compactor leverage + chunk attention mass, SnapKV) vs the CPU oracle and the
reference-generated goldens.

Tolerance: finite entries |d| <= 2e-2 * max(1, |s|), +-inf pattern exact.  The
leverage chain is ill-conditioned by construction in the reference (every intermediate lives in
K's 16-bit dtype, and the z-score divides by a chunk's std), so next to the element tolerance it
is judged the way it is used: overlap of the top-50 % kept set."""
import math

import numpy as np
import pytest
import torch

from compactor_vllm.compression.compactor import approximate_leverage_scores, non_causal_attn_scores
from compactor_vllm.compression.snapkv import query_aware_key_scores
from compactor_vllm.utils.arguments import make_phi
from oracle import scoring as osc

from conftest import load_golden
from helpers import fused_qkv

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(out, ref, tol=2e-2):
    out, ref = out.float().cpu(), ref.float().cpu()
    assert torch.equal(torch.isinf(out), torch.isinf(ref)), "inf pattern differs"
    assert torch.equal(torch.isnan(out), torch.isnan(ref)), "nan pattern differs"
    fin = torch.isfinite(ref)
    if fin.any():
        d = (out[fin] - ref[fin]).abs() / ref[fin].abs().clamp_min(1.0)
        assert d.max().item() <= tol, d.max().item()


def _cu(lens):
    return torch.tensor([0] + np.cumsum(lens).tolist(), dtype=torch.int32)


# ------------------------------------------------------------------ chunk-local attention mass
def test_chunk_scores_golden():
    g = load_golden("chunk_scores")
    q, k = g["q"].to(DEV), g["k"].to(DEV)
    cu = g["cu_seqlens"].to(DEV)
    lens = (g["cu_seqlens"][1:] - g["cu_seqlens"][:-1]).tolist()
    raw = non_causal_attn_scores(q, k, None, cu, max(lens), g["chunk"], sm_scale=g["sm_scale"], normalize=False)
    _close(raw, g["raw"])
    full = non_causal_attn_scores(
        q, k, None, cu, max(lens), g["chunk"], sm_scale=g["sm_scale"], normalize=True, context_lens=lens,
        protected_first_tokens=g["prot_first"].tolist(), protected_last_tokens=g["prot_last"].tolist(),
        accum_scores=g["accum"].to(DEV), accum_blending=0.5)
    _close(full, g["full"])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,HQ,HKV", [([1], 32, 8), ([127, 129, 64], 32, 8), ([1000], 8, 1), ([300, 5], 16, 2),
                                         ([513], 32, 8)])
@pytest.mark.parametrize("sm_scale", [1.0, None])
def test_chunk_scores_vs_oracle(dtype, lens, HQ, HKV, sm_scale):
    N, D = sum(lens), 128
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=N + HQ)
    if sm_scale == 1.0:  # the engine's setting (compactor.py:51-58): keep logits in a sane range
        q = (q * 0.3).contiguous()
        k = (k * 0.3).contiguous()
    cu = _cu(lens)
    raw = non_causal_attn_scores(q, k, None, cu.to(DEV), max(lens), 128, sm_scale=sm_scale, normalize=False)
    ref = osc.non_causal_attn_scores(q.cpu(), k.cpu(), cu, 128, sm_scale=sm_scale, normalize=False)
    _close(raw, ref)
    if min(lens) > 1:
        g = torch.Generator().manual_seed(3)
        accum = torch.randn((N, HKV), generator=g).to(dtype)
        pf, pl = [4] * len(lens), [16] * len(lens)
        out = non_causal_attn_scores(q, k, None, cu.to(DEV), max(lens), 128, sm_scale=sm_scale, normalize=True,
                                     context_lens=lens, protected_first_tokens=pf, protected_last_tokens=pl,
                                     accum_scores=accum.to(DEV), accum_blending=0.5)
        ref = osc.non_causal_attn_scores(q.cpu(), k.cpu(), cu, 128, sm_scale=sm_scale, normalize=True,
                                         context_lens=lens, protected_first_tokens=pf, protected_last_tokens=pl,
                                         accum_scores=accum, accum_blending=0.5)
        _close(out, ref)


# ------------------------------------------------------------------------------------ SnapKV
@pytest.mark.parametrize("seg", [32, 64])
def test_snapkv_golden(seg):
    g = load_golden(f"snapkv_seg{seg}")
    cu = g["cu_seqlens"].to(DEV)
    out = query_aware_key_scores(g["q"].to(DEV), g["k"].to(DEV), cu, cu, g["w"], pool_segment=g["pool_segment"])
    ok = ~g["untouched_rows"]
    _close(out.cpu()[ok], g["out"][ok])
    assert (out.cpu()[~ok] == 0).all()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,HQ,HKV,w", [([40], 32, 8, 32), ([33, 500, 32, 10], 32, 8, 32), ([5000], 32, 8, 32),
                                           ([2100, 300], 8, 1, 32), ([700], 16, 2, 8)])
@pytest.mark.parametrize("seg", [64, 128])
def test_snapkv_vs_oracle(dtype, lens, HQ, HKV, w, seg):
    N, D = sum(lens), 128
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=N + w)
    cu = _cu(lens)
    out = query_aware_key_scores(q, k, cu.to(DEV), cu.to(DEV), w, pool_segment=seg)
    ref = osc.query_aware_key_scores(q.cpu(), k.cpu(), cu, cu, w, pool_segment=seg)
    _close(out, ref)


def test_snapkv_accumulates_in_place():
    lens, HQ, HKV, w = [300], 32, 8, 32
    q, k, _ = fused_qkv(sum(lens), HQ, HKV, 128, torch.bfloat16, DEV, seed=1)
    cu = _cu(lens).to(DEV)
    base = query_aware_key_scores(q, k, cu, cu, w)
    acc = torch.ones_like(base)
    ret = query_aware_key_scores(q, k, cu, cu, w, accum_scores=acc, accum_blending=0.25)
    assert ret.data_ptr() == acc.data_ptr()  # the reference mutates its argument (snapkv.py:442-446)
    fin = torch.isfinite(base)
    assert torch.allclose(ret[fin], base[fin] + 0.25)


# ---------------------------------------------------------------------------------- leverage
def _gram_like(mats, rank, noise, seed, dtype=torch.bfloat16):
    """fp32 Grams shaped like the chain's: X^T X of a low-rank + noise sketch, rounded entrywise to 16 bit
    (so possibly indefinite), + 5e-3 I.  |lam| spans ~1e-2 .. 1e5."""
    g = torch.Generator().manual_seed(seed)
    X = torch.randn((mats, 512, rank), generator=g) @ (3.0 * torch.randn((mats, rank, 48), generator=g))
    X = X + noise * torch.randn((mats, 512, 48), generator=g)
    X = X - X.mean(dim=1, keepdim=True)
    G = (X.transpose(1, 2) @ X).to(dtype).float()
    G = torch.triu(G) + torch.triu(G, 1).transpose(1, 2)
    return G + 5e-3 * torch.eye(48)


def test_eigensolver_vs_fp64_eigh():
    """The fp64 CTA-per-matrix tred2/tql2 (csrc/score_leverage.cu) against torch.linalg.eigh in fp64 on
    graded, indefinite Gram matrices plus degenerate cases (reg*I, diagonal, a zero row/column block,
    repeated eigenvalues): eigenvalues to 1e-10 |lam|_max, residual and orthogonality to 1e-9, and the product
    output round16(V |lam|^-1/2) reproduced from the returned (lam, V)."""
    from compactor_vllm import _native as N
    mats = [_gram_like(24, 12, 0.05, 0), _gram_like(8, 3, 1e-3, 1), _gram_like(8, 48, 1.0, 2)]
    special = torch.zeros((4, 48, 48))
    special[0] = 5e-3 * torch.eye(48)
    special[1] = torch.diag(torch.linspace(-3.0, 1e4, 48))
    special[2] = _gram_like(1, 12, 0.05, 3)[0]
    special[2][20:30] = 0
    special[2][:, 20:30] = 0
    special[2][range(20, 30), range(20, 30)] = 5e-3  # decoupled block: H_i = I steps in the reduction
    special[3][:24, :24] = _gram_like(1, 6, 0.1, 4)[0][:24, :24]
    special[3][24:, 24:] = special[3][:24, :24]  # every eigenvalue twice
    G = torch.cat(mats + [special]).contiguous()
    M = G.shape[0]
    Gd = G.to(DEV)
    lam = torch.empty((M, 48), dtype=torch.float64, device=DEV)
    vec = torch.empty((M, 48, 48), dtype=torch.float64, device=DEV)
    sv = torch.empty((M, 48, 48), dtype=torch.float32, device=DEV)
    N.check(N.lib().cvllm_selftest_eig(N.ptr(Gd), N.ptr(lam), N.ptr(vec), N.ptr(sv), M, N.CVLLM_BF16, None, N.stream()),
            "cvllm_selftest_eig")
    torch.cuda.synchronize()
    lam, vec, sv = lam.cpu(), vec.cpu(), sv.cpu()
    G64 = G.double()
    ref = torch.linalg.eigvalsh(G64)
    scale = ref.abs().max(dim=1).values.clamp_min(1e-300)
    assert ((lam.sort(dim=1).values - ref).abs().max(dim=1).values / scale).max().item() <= 1e-10
    resid = (G64 @ vec - vec * lam.unsqueeze(1)).abs().amax(dim=(1, 2)) / scale
    assert resid.max().item() <= 1e-9, resid.max().item()
    orth = (vec.transpose(1, 2) @ vec - torch.eye(48, dtype=torch.float64)).abs().amax(dim=(1, 2))
    assert orth.max().item() <= 1e-9, orth.max().item()
    want = (vec * lam.abs().rsqrt().unsqueeze(1)).float().to(torch.bfloat16).float()
    assert torch.equal(sv, want)


def _structured_keys(N, H, D, dtype, seed):
    """Low-rank + positional drift + noise: leverage scores with real spread (randn keys make them
    near-uniform)."""
    g = torch.Generator().manual_seed(seed)
    basis = torch.randn((H, 12, D), generator=g)
    coef = torch.randn((N, H, 12), generator=g) * torch.linspace(2.0, 0.2, 12)
    pos = torch.linspace(0, 6.28, N)[:, None, None] * torch.ones(1, H, 1)
    x = torch.einsum("nhr,hrd->nhd", coef, basis) + 0.5 * torch.sin(pos + torch.arange(D) * 0.1)
    x = x + 0.3 * torch.randn((N, H, D), generator=g)
    spikes = torch.randint(0, N, (max(1, N // 50),), generator=g)
    x[spikes] += 3.0 * torch.randn((spikes.numel(), H, D), generator=g)
    return x.to(dtype)


def _overlap(a, b, frac=0.5):
    k = max(1, int(a.numel() * frac))
    ia = set(torch.topk(a.float().flatten(), k).indices.tolist())
    ib = set(torch.topk(b.float().flatten(), k).indices.tolist())
    return len(ia & ib) / k


@pytest.mark.parametrize("name", ["leverage_bf16"])
def test_leverage_golden(name):
    g = load_golden(name)
    lens = g["context_lens"].tolist()
    raw = approximate_leverage_scores(g["keys"].to(DEV), lens, g["phi"].to(DEV), regularizer=g["regularizer"],
                                      normalize=False, chunk_size=g["chunk_size"])
    assert raw.dtype == g["keys"].dtype
    d = (raw.float().cpu() - g["raw"].float()).abs()
    # Chunks shorter than the sketch (48 columns) have a rank-deficient Gram whose small
    # eigenvalues are bf16 rounding noise: swapping the reference's own factorisation
    # (fp32 svd -> fp64 svd or eigh) moves those scores by 0.03-0.05 on this very fixture
    # (tools/leverage_sensitivity.py), so they get the noise-floor bound; full-rank chunks the
    # north_star bound.
    _, chunk_lens = osc.split_into_chunks(lens, g["chunk_size"])
    off = 0
    for L in chunk_lens:
        blk, ref_blk = d[off:off + L], g["raw"][off:off + L].float()
        tol = 2e-2 * max(1.0, ref_blk.abs().max().item()) if L >= 48 else 0.1
        assert blk.max().item() <= tol, (L, blk.max().item())
        off += L
    norm = approximate_leverage_scores(g["keys"].to(DEV), lens, g["phi"].to(DEV), regularizer=g["regularizer"],
                                       normalize=True, chunk_size=g["chunk_size"])
    full = torch.zeros(len(norm), dtype=torch.bool)
    off = 0
    for L in chunk_lens:
        full[off:off + L] = L >= 48
        off += L
    # z-scores live in 16-bit storage on both sides: the bound is relative to the
    # score magnitude (2 bf16 ulps at |z| in [4, 8) is already 0.06)
    ref_n = g["norm"].float()[full].flatten()
    dn = (norm.float().cpu()[full].flatten() - ref_n).abs() / ref_n.abs().clamp_min(1.0)
    assert dn.kthvalue(int(0.99 * dn.numel())).values.item() <= 5e-2, dn.max().item()
    assert _overlap(norm.cpu()[full], g["norm"][full]) >= 0.98


def _median_rel_err(x, exact):
    e = ((x.float() - exact.float()).abs() / exact.float().abs().clamp_min(1.0)).flatten()
    return e.median().item()


def _isotropic_keys(N, H, D, dtype, seed):
    """Near-isotropic keys with a few mildly stronger directions: the sketch Gram has a condition
    number < 10, far above the 2^-9 relative rounding noise of its 16-bit entries, so leverage
    scores are well defined and the strict tolerance applies."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((N, H, D), generator=g)
    basis = torch.randn((H, 4, D), generator=g) / D ** 0.5
    coef = torch.randn((N, H, 4), generator=g) * torch.tensor([2.0, 1.5, 1.0, 0.7])
    x = x + torch.einsum("nhr,hrd->nhd", coef, basis) * D ** 0.5 * 0.5
    return x.to(dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,H,chunk", [([512], 8, 512), ([1300, 700], 8, 512), ([2048], 4, 512),
                                          ([384, 256], 8, 128)])
def test_leverage_vs_oracle_well_conditioned(dtype, lens, H, chunk):
    """Strict parity where the score is well defined: |delta| <= 2e-2 * max(1, |s|) on every entry
    against the reference arithmetic (fp32 SVD, all 16-bit rounding points)."""
    N, D = sum(lens), 128
    keys = _isotropic_keys(N, H, D, dtype, seed=N + H)
    phi = make_phi(D, 48, dtype, "cpu")
    raw = approximate_leverage_scores(keys.to(DEV), lens, phi.to(DEV), normalize=False, chunk_size=chunk).cpu()
    ref = osc.approximate_leverage_scores(keys, lens, phi, normalize=False, chunk_size=chunk)
    assert raw.dtype == dtype
    d = (raw.float() - ref.float()).abs()
    assert (d <= 2e-2 * ref.float().abs().clamp_min(1.0)).all(), d.max().item()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,H,chunk", [([512], 8, 512), ([1300, 700], 8, 512), ([100, 49, 600], 2, 512),
                                          ([2048], 4, 512), ([300], 8, 128)])
def test_leverage_low_rank_keys_statistics(dtype, lens, H, chunk):
    """Low-rank keys (what real K looks like): the 16-bit Gram is indefinite, its small |lambda| (~1
    against lambda_max ~1e5) are rounding noise of the Gram itself, and x^T |G|^-1 x is ill-conditioned
    in the reference too: one ulp in one entry of the 16-bit column mean (a summation-order effect,
    tools/debug_leverage2.py) moves ~1 % of a chunk's scores by O(1), and the reference's own fp32 SVD
    sits as far from the exact factorisation ("eigh64") of the same Gram as we do.  What is stable —
    and asserted — is the bulk: median distance to the exact answer no worse than the reference
    arithmetic's, scores finite and >= 0, and (full-rank chunks) the z-scored top half agreeing with the
    exact answer as well as the reference arithmetic does."""
    N, D = sum(lens), 128
    scale = 1.0 if dtype == torch.bfloat16 else 0.1  # fp16: keep the Gram below 65504 (inf otherwise)
    keys = (_structured_keys(N, H, D, torch.float32, seed=N + H) * scale).to(dtype)
    phi = make_phi(D, 48, dtype, "cpu")
    raw = approximate_leverage_scores(keys.to(DEV), lens, phi.to(DEV), normalize=False, chunk_size=chunk).cpu()
    exact = osc.approximate_leverage_scores(keys, lens, phi, normalize=False, chunk_size=chunk,
                                            factorization="eigh64")
    ref = osc.approximate_leverage_scores(keys, lens, phi, normalize=False, chunk_size=chunk,
                                          factorization="svd32")
    assert raw.dtype == dtype and torch.isfinite(raw.float()).all() and (raw.float() >= 0).all()
    assert _median_rel_err(raw, exact) <= 3.0 * _median_rel_err(ref, exact) + 1e-2
    norm = approximate_leverage_scores(keys.to(DEV), lens, phi.to(DEV), normalize=True, chunk_size=chunk).cpu()
    exact_n = osc.approximate_leverage_scores(keys, lens, phi, normalize=True, chunk_size=chunk,
                                              factorization="eigh64")
    ref_n = osc.approximate_leverage_scores(keys, lens, phi, normalize=True, chunk_size=chunk,
                                            factorization="svd32")
    _, chunk_lens = osc.split_into_chunks(lens, chunk)
    if all(c >= 64 for c in chunk_lens):  # shorter chunks: leverage ~1 everywhere, z-score is noise
        assert _overlap(norm, exact_n) >= _overlap(ref_n, exact_n) - 0.03


def _reference_arithmetic_on_device(keys, lens, phi, chunk, reg=5e-3):
    """compression/compactor.py:153-210 re-issued with torch on the SAME device — the reference's own call
    sequence (cuBLAS GEMM on the transposed view, in-place mean subtraction, batched Gram GEMM, 16-bit roundings
    where torch puts them) — with the `gesvda` SVD replaced by the factorisation it approximates: fp64 `eigh`
    of the fp32 Gram, S = |lambda|.  On the B200 the unmodified reference is itself only 99.25 % within 2e-2 of
    this on low-rank keys (profiles/r02c_leverage_selfdistance.json: ref_eigh64_vs_ref) while this repo is at
    100.00 % (ours_vs_eigh64), so this is the tighter pin."""
    X = torch.matmul(keys.transpose(0, 1), phi)  # [H, N, k]
    H, N, k = X.shape
    coalesced, _ = osc.split_into_chunks(lens, chunk)
    pieces = torch.split(X, coalesced, dim=-2)
    out = []
    for L, piece in zip(coalesced, pieces):
        c = piece.view(H, -1, chunk, k) if (chunk > 0 and L % chunk == 0) else piece.unsqueeze(1)
        c = c.clone()
        c.sub_(c.mean(dim=-2, keepdim=True))
        G = torch.matmul(c.transpose(-1, -2), c).to(torch.float32)
        G.diagonal(dim1=-2, dim2=-1).add_(reg)
        lam, V = torch.linalg.eigh(G.double())
        SV = (V * lam.abs().rsqrt().unsqueeze(-2)).to(torch.float32).to(X.dtype)
        U = torch.matmul(c, SV)
        out.append((U * U).sum(dim=-1).clamp_min_(0.0).view(H, -1).transpose(-1, -2))
    return torch.cat(out, dim=0)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,chunk", [([4096], 512), ([1536, 700, 49], 512), ([32768], 512)])
def test_leverage_matches_reference_arithmetic_on_device(dtype, lens, chunk):
    """Low-rank keys, where one flipped ulp of the 16-bit sketch moves a chunk's scores by O(1): the kernel
    chain has to reproduce the reference's cuBLAS bits through the Gram (natural-k sketch chain, sequential
    Gram accumulation) and factorise exactly (fp64 tridiagonal QL).  Bar: >= 99.95 % of the scores within
    2e-2 * max(1, |s|), none beyond 0.1, top-half overlap >= 99.9 %."""
    N, H, D = sum(lens), 8, 128
    scale = 1.0 if dtype == torch.bfloat16 else 0.1  # fp16: keep the Gram below 65504
    keys = (_structured_keys(N, H, D, torch.float32, seed=N) * scale).to(dtype).to(DEV)
    phi = make_phi(D, 48, dtype, DEV)
    got = approximate_leverage_scores(keys, lens, phi, normalize=False, chunk_size=chunk).float()
    ref = _reference_arithmetic_on_device(keys, lens, phi, chunk).float()
    _, chunk_lens = osc.split_into_chunks(lens, chunk)
    full = torch.cat([torch.full((L,), L >= 48, dtype=torch.bool) for L in chunk_lens]).to(DEV)
    err = (got - ref).abs() / ref.abs().clamp_min(1.0)
    assert torch.isfinite(got).all()
    assert (err[full] <= 2e-2).float().mean().item() >= 0.9995, (err[full] <= 2e-2).float().mean().item()
    assert err[full].max().item() <= 0.1, err[full].max().item()
    assert _overlap(got[full], ref[full]) >= 0.999


def test_leverage_one_token_chunk_is_nan():
    """a 1-token tail chunk z-scores to 0/0 = NaN, which the ranking keeps first."""
    lens, H = [513], 8
    keys = _structured_keys(sum(lens), H, 128, torch.bfloat16, seed=5)
    phi = make_phi(128, 48, torch.bfloat16, "cpu")
    out = approximate_leverage_scores(keys.to(DEV), lens, phi.to(DEV), normalize=True, chunk_size=512)
    assert torch.isnan(out[512]).all() and torch.isfinite(out[:512]).all()


def test_plugin_scoring_is_stream_independent():
    """CompactorCompression.pre/post_rope_scoring on the caller's stream, on a high-priority store stream and on
    a default-priority one: same bits, twice in a row (per-stream workspace reuse)."""
    from compactor_vllm.compression.compactor import CompactorCompression
    from compactor_vllm.compression.compression_config import CompressionMethod
    from compactor_vllm.utils.arguments import max_tokens_to_retain, tokens_to_retain
    from compactor_vllm.utils.context import CompressionContext, Context

    lens, HQ, HKV, D, dtype = [1500, 77, 600], 32, 8, 128, torch.bfloat16
    N = sum(lens)
    keys_pre = _structured_keys(N, HKV, D, dtype, seed=21).to(DEV)
    q, k, v = fused_qkv(N, HQ, HKV, D, dtype, DEV, seed=22)
    cu = _cu(lens).to(DEV)
    pf, pl = [16] * 3, [64] * 3
    cc = CompressionContext(CompressionMethod.COMPACTOR, 512,
                            torch.tensor(tokens_to_retain(lens, [0.5] * 3, pf, pl, HKV), dtype=torch.int32, device=DEV),
                            max_tokens_to_retain(lens, HKV), lens, make_phi(D, 48, dtype, DEV), pf, pl)
    results = []
    for store in (None, torch.cuda.Stream(priority=-1), torch.cuda.Stream()):
        ctx = Context(is_prefill=True, do_compression=True, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=max(lens),
                      max_seqlen_k=max(lens), compression_context=cc, STORE_STREAM=store)
        for _ in range(2):
            pre = CompactorCompression.pre_rope_scoring(q, keys_pre, v, ctx)
            out = CompactorCompression.post_rope_scoring(q, k, v, pre, ctx)
            if store is not None:
                torch.cuda.current_stream().wait_stream(store)
            torch.cuda.synchronize()
            results.append(out.clone())
    for r in results[1:]:
        assert torch.equal(torch.isinf(r), torch.isinf(results[0]))
        fin = torch.isfinite(results[0])
        assert torch.equal(r[fin], results[0][fin])


# ---------------------------------------------------------------- end-to-end Compactor scores
def test_compactor_pipeline_kept_set_overlap():
    """pre-RoPE leverage -> post-RoPE chunk scores (+0.5 blend, protection) -> top-50 % kept set:
    overlap with the oracle's kept set, the quantity that matters downstream."""
    lens, HQ, HKV, D, dtype = [1500, 600], 32, 8, 128, torch.bfloat16
    N = sum(lens)
    keys_pre = _structured_keys(N, HKV, D, dtype, seed=11)
    q, k, _ = fused_qkv(N, HQ, HKV, D, dtype, "cpu", seed=12)
    q, k = (q * 0.3).contiguous(), (k * 0.3).contiguous()
    phi = make_phi(D, 48, dtype, "cpu")
    cu = _cu(lens)
    pf, pl = [16] * 2, [64] * 2
    pre = approximate_leverage_scores(keys_pre.to(DEV), lens, phi.to(DEV), normalize=True, chunk_size=512)
    out = non_causal_attn_scores(q.to(DEV), k.to(DEV), None, cu.to(DEV), max(lens), 128, sm_scale=1.0,
                                 normalize=True, context_lens=lens, protected_first_tokens=pf,
                                 protected_last_tokens=pl, accum_scores=pre, accum_blending=0.5)
    refs = {}
    for fact in ("svd32", "eigh64"):
        pre0 = osc.approximate_leverage_scores(keys_pre, lens, phi, normalize=True, chunk_size=512,
                                               factorization=fact)
        refs[fact] = osc.non_causal_attn_scores(q, k, cu, 128, sm_scale=1.0, normalize=True, context_lens=lens,
                                                protected_first_tokens=pf, protected_last_tokens=pl,
                                                accum_scores=pre0, accum_blending=0.5)
    assert torch.equal(torch.isinf(out.cpu()), torch.isinf(refs["svd32"]))
    for b in range(2):
        s, e = int(cu[b]), int(cu[b + 1])
        # the reference arithmetic's own kept set is only this close to the exact factorisation's
        floor = _overlap(refs["svd32"][s:e], refs["eigh64"][s:e])
        assert _overlap(out.cpu()[s:e], refs["eigh64"][s:e]) >= floor - 0.03, floor
