"""Oracle: Compactor (leverage + chunk attention) and SnapKV key scores."""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import torch

LOG2E = 1.4426950408889634
REF_CHUNK_BLOCK_M = 64  # the reference's only autotune config for K11 (compression/compactor.py:338-347)


def split_into_chunks(xs: Sequence[int], chunk_size: int):
    """compression/compactor.py:62-110 — ([coalesced], [chunks]) e.g. [257,127],128 -> ([256,1,127],[128,128,1,127])."""
    coalesced, chunks = [], []
    for n in xs:
        full, tail = divmod(n, chunk_size)
        if full:
            coalesced.append(full * chunk_size)
            chunks.extend([chunk_size] * full)
        if tail:
            coalesced.append(tail)
            chunks.append(tail)
    return coalesced, chunks


def zscore_segments_(scores: torch.Tensor, seg_offsets: Sequence[int]) -> None:
    """compression/compactor.py:224-269: per segment over ALL heads, fp32 sum / sum-of-squares,
    var = max(E[x^2] - mean^2, 0), x <- (x - mean) / sqrt(var)  (no eps: a constant segment gives
    0/0 = NaN, which the ranking treats as largest).  In place, result stored in scores.dtype."""
    H = scores.shape[1]
    for i in range(len(seg_offsets) - 1):
        s, e = int(seg_offsets[i]), int(seg_offsets[i + 1])
        if e <= s:
            continue
        x = scores[s:e].float()
        cnt = float((e - s) * H)
        mean = x.sum() / cnt
        var = torch.clamp(x.pow(2).sum() / cnt - mean * mean, min=0.0)
        scores[s:e] = ((x - mean) * (1.0 / torch.sqrt(var))).to(scores.dtype)


def approximate_leverage_scores(key_states: torch.Tensor, context_lens: List[int], PHI: torch.Tensor,
                                regularizer: float = 5e-3, normalize: bool = False, chunk_size: int = 512,
                                emulate_dtype: bool = True, factorization: str = "svd32") -> torch.Tensor:
    """compression/compactor.py:113-221.

    X = K_h PHI; per chunk: centre, G = X^T X + reg*I, (V,S) = svd(G),
    score_i = || x_i V S^-1/2 ||^2 (== x_i^T G^-1 x_i), clamp >= 0, optional per-chunk z-score.
    emulate_dtype=True keeps the reference's rounding points (X, the centred chunk, the Gram,
    V S^-1/2, U and the row sums all live in K's dtype, :158,165-171,197,209-210);
    emulate_dtype=False evaluates the same formula in fp32 from the same inputs.
    factorization: "svd32" = torch.linalg.svd in fp32, the reference's arithmetic (:177 uses cuSOLVER
    gesvda, an approximate fp32 SVD); "eigh64" = exact symmetric eigendecomposition in fp64 of the SAME
    (rounded) Gram, singular values |lambda| — the noise-free answer both fp32 factorisations
    approximate.  On low-rank keys the 16-bit Gram is indefinite with |lambda| down to ~1e-2 against
    lambda_max ~1e5, so scores move by O(1) between fp32 factorisations; tests measure distances to
    "eigh64" and require ours to be no larger than "svd32"'s.
    Returns [N, H] in K's dtype.
    """
    dt = key_states.dtype if emulate_dtype else torch.float32
    if chunk_size > 0:
        _, chunk_lens = split_into_chunks(context_lens, chunk_size)
    else:
        chunk_lens = list(context_lens)
    X = torch.matmul(key_states.transpose(0, 1).to(dt), PHI.to(dt))  # [H, N, k]
    H, N, k = X.shape
    out = torch.empty((N, H), dtype=dt)
    offs = [0]
    for L in chunk_lens:
        s = offs[-1]
        e = s + L
        offs.append(e)
        chunk = X[:, s:e]
        chunk = chunk - chunk.mean(dim=-2, keepdim=True)
        g = torch.matmul(chunk.transpose(-1, -2), chunk).to(torch.float32)  # [H, k, k]
        g = g + regularizer * torch.eye(k)
        if factorization == "eigh64":
            lam, V = torch.linalg.eigh(g.double())
            S, V = lam.abs().float(), V.float()
        else:
            V, S, _ = torch.linalg.svd(g, full_matrices=False)
        SV = (V * S.rsqrt().unsqueeze(-2)).to(dt)
        U = torch.matmul(chunk, SV)
        out[s:e] = (U * U).sum(dim=-1).clamp_min(0.0).transpose(0, 1)
    if normalize:
        zscore_segments_(out, offs)
    return out.to(key_states.dtype)


def non_causal_attn_scores(q, k, cu_seqlens_qk, chunk_size: int, sm_scale: Optional[float] = None,
                           normalize: bool = True, context_lens: Optional[List[int]] = None,
                           protected_first_tokens: Optional[List[int]] = None,
                           protected_last_tokens: Optional[List[int]] = None,
                           accum_scores: Optional[torch.Tensor] = None,
                           accum_blending: Optional[float] = None) -> torch.Tensor:
    """compression/compactor.py:338-599 -> fp32 [N, HKV].

    Inside each chunk_size-token chunk (chunks restart at every sequence start) and for each kv
    head: P = softmax over the chunk's keys of sm_scale*q.k^T for all G*M query rows;
    score_key = sum_rows P.  Rows of the LAST 64-row query tile that fall beyond the chunk add
    1/chunk_size each (:411-414,477-479: the mask is per BLOCK_M=64 tile, so the number of such
    rows is (ceil(M/64)*64 - M)*G, not (chunk_size - M)*G).  Then z-score per sequence (:581-585),
    + blend*accum (:586-590), protected first/last tokens <- +inf (:591-598; clamped to the
    sequence here — the reference's slice can run into the next sequence).
    """
    N, HQ, D = q.shape
    HKV = k.shape[1]
    G = HQ // HKV
    scale = 1.0 / math.sqrt(D) if sm_scale is None else sm_scale
    out = torch.zeros((N, HKV), dtype=torch.float32)
    cu = [int(x) for x in cu_seqlens_qk]
    for b in range(len(cu) - 1):
        for cs in range(cu[b], cu[b + 1], chunk_size):
            ce = min(cs + chunk_size, cu[b + 1])
            M = ce - cs
            padded_rows = (-(-M // REF_CHUNK_BLOCK_M) * REF_CHUNK_BLOCK_M - M) * G
            for h in range(HKV):
                qh = q[cs:ce, h * G:(h + 1) * G].float().reshape(M * G, D)
                s = (qh @ k[cs:ce, h].float().T) * scale
                p = torch.softmax(s, dim=-1)
                out[cs:ce, h] = p.sum(dim=0) + padded_rows * (1.0 / chunk_size)
    if normalize:
        zscore_segments_(out, cu)
    if accum_scores is not None:
        out += accum_scores.float() * (1.0 if accum_blending is None else accum_blending)
    if protected_first_tokens is not None or protected_last_tokens is not None:
        B = len(cu) - 1
        pf = protected_first_tokens if protected_first_tokens is not None else [0] * B
        pl = protected_last_tokens if protected_last_tokens is not None else [0] * B
        for b in range(B):
            s, e = cu[b], cu[b + 1]
            out[s:min(s + pf[b], e)] = float("inf")
            out[max(e - pl[b], s):e] = float("inf")
    return out


def query_aware_key_scores(q, k, cu_seqlens_q, cu_seqlens_k, w: int, sm_scale: Optional[float] = None,
                           kpool: int = 5, pool_segment: int = 64) -> torch.Tensor:
    """compression/snapkv.py:39-276,332-448 -> fp32 [Nk, HKV].

    Rows = last w queries x G heads of each sequence; keys [0, L-w); softmax over those keys only
    (:127-151); score_key = sum_rows P; causal kpool-tap mean pool that never crosses a
    pool_segment boundary (segments counted from the sequence start; the reference's segment is
    its autotuned BLOCK_K, :160-168,253-262); keys [L-w, L) <- +inf (:267-276).  Sequences with
    L <= w: all zeros here (uninitialised in the reference, :94-95,204-205).
    """
    Nq, HQ, D = q.shape
    Nk, HKV, _ = k.shape
    G = HQ // HKV
    scale = 1.0 / math.sqrt(D) if sm_scale is None else sm_scale
    out = torch.zeros((Nk, HKV), dtype=torch.float32)
    B = len(cu_seqlens_q) - 1
    for b in range(B):
        q_end = int(cu_seqlens_q[b + 1])
        kb, ke = int(cu_seqlens_k[b]), int(cu_seqlens_k[b + 1])
        keff = ke - w
        if w <= 0 or keff <= kb:
            continue
        for h in range(HKV):
            qh = q[q_end - w:q_end, h * G:(h + 1) * G].float().reshape(w * G, D)
            s = (qh @ k[kb:keff, h].float().T) * scale
            raw = torch.softmax(s, dim=-1).sum(dim=0)  # [keff-kb]
            n = raw.numel()
            pooled = torch.empty_like(raw)
            for i in range(n):
                lo = max(i - (kpool - 1), (i // pool_segment) * pool_segment)
                pooled[i] = raw[lo:i + 1].sum() / float(i + 1 - lo)
            out[kb:keff, h] = pooled
            out[keff:ke, h] = float("inf")
    return out
