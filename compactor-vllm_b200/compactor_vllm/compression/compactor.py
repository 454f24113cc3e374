"""COMPACTOR scoring: approximate leverage scores (pre-RoPE) blended into
chunk-local non-causal attention mass (post-RoPE).

Reference: compression/compactor.py:16-599.  The two scorers are single C-ABI
calls (csrc/score_leverage.cu, csrc/score_attn.cu): no Python loops over chunks,
no cuSOLVER.  Each chunk's 16-bit 48x48 Gram matrix gets a symmetric
eigendecomposition in shared memory (fp64 tridiagonal reduction + implicit QL,
one CTA per matrix) and the score is ||x V |lambda|^-1/2||^2 — what the
reference's SVD yields, :173-210.  A Cholesky solve is NOT equivalent: the
rounded Gram of low-rank keys is indefinite.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch
from compactor_vllm import _native as N
from compactor_vllm.compression.common import BaseCompressionMethod
from compactor_vllm.utils.helpers import maybe_execute_in_stream


class CompactorCompression(BaseCompressionMethod):
    chunk_size: int = 128

    @staticmethod
    def pre_rope_scoring(q, k, v, context) -> Optional[torch.Tensor]:
        cc = context.compression_context
        return maybe_execute_in_stream(
            approximate_leverage_scores, k, cc.context_lens, cc.PHI, normalize=True,
            chunk_size=cc.compression_chunk_size, STORE_STREAM=context.STORE_STREAM,
        )

    @staticmethod
    def pre_rope_scoring_from_sketch(sketch_rows, context) -> Optional[torch.Tensor]:
        """pre_rope_scoring for runners that call fused_qk_norm_rope (the sketch replaces k)."""
        cc = context.compression_context
        return maybe_execute_in_stream(
            approximate_leverage_scores, None, cc.context_lens, cc.PHI, normalize=True,
            chunk_size=cc.compression_chunk_size, sketch_rows=sketch_rows,
            STORE_STREAM=context.STORE_STREAM,
        )

    @staticmethod
    def post_rope_scoring(q, k, v, pre_rope_scores, context) -> Optional[torch.Tensor]:
        cc = context.compression_context
        # The reference launches this on the main stream while pre_rope_scores is still being
        # produced on STORE_STREAM without a wait (compactor.py:44-59).  Here the
        # whole scoring chain stays on STORE_STREAM: it is ordered after the leverage kernels by
        # the stream itself, its only consumer (extract_and_store_top_kv in Attention.forward)
        # runs on the same stream, and the attention kernel on the main stream starts without
        # waiting for any of it — scoring, selection and compaction all overlap attention.
        return maybe_execute_in_stream(
            non_causal_attn_scores,
            q, k, v, context.cu_seqlens_q, context.max_seqlen_q,
            chunk_size=CompactorCompression.chunk_size, sm_scale=1.0, normalize=True,
            accum_scores=pre_rope_scores, context_lens=cc.context_lens,
            protected_first_tokens=cc.protected_first_tokens,
            protected_last_tokens=cc.protected_last_tokens, accum_blending=0.5,
            STORE_STREAM=context.STORE_STREAM,
        )


def split_into_chunks(xs, chunk_size):
    """(coalesced_chunks, chunks) decomposition of sequence lengths (reference :62-110).

    xs=[257,127], chunk_size=128 -> ([256,1,127], [128,128,1,127]).
    """
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


_OFFSETS_CACHE: dict = {}


def _chunk_offsets(chunk_lens: tuple, dev: torch.device) -> torch.Tensor:
    """Device int32 prefix offsets of the chunks; cached per shape so a steady-state step issues no
    host->device copy (and the whole scoring chain can be captured in a CUDA graph)."""
    key = (chunk_lens, dev.index)
    t = _OFFSETS_CACHE.get(key)
    if t is None:
        offs = [0]
        for c in chunk_lens:
            offs.append(offs[-1] + c)
        t = torch.tensor(offs, dtype=torch.int32, device=dev)
        if len(_OFFSETS_CACHE) > 64:
            _OFFSETS_CACHE.clear()
        _OFFSETS_CACHE[key] = t
    return t


def _protected_counts(pf: tuple, pl: tuple, dev: torch.device) -> torch.Tensor:
    key = ("prot", pf, pl, dev.index)
    t = _OFFSETS_CACHE.get(key)
    if t is None:
        t = torch.tensor(list(pf) + list(pl), dtype=torch.int32, device=dev)
        if len(_OFFSETS_CACHE) > 64:
            _OFFSETS_CACHE.clear()
        _OFFSETS_CACHE[key] = t
    return t


def approximate_leverage_scores(
    key_states: torch.Tensor,  # [N, H, D]
    context_lens: List[int],  # [B]
    PHI: torch.Tensor,  # [D, k]
    regularizer: float = 5e-3,
    normalize: bool = False,
    chunk_size: int = 512,
    *,
    sketch_rows: Optional[torch.Tensor] = None,  # [N, H, k] = round(K.PHI) from fused_qk_norm_rope
) -> torch.Tensor:  # [N, H] in key_states.dtype
    """Randomised leverage scores of the pre-RoPE keys, per <=chunk_size chunk.

    `sketch_rows` (extension): the sketch X = K.PHI already produced by the fused norm + RoPE
    kernel (layers/rotary_embedding.py: fused_qk_norm_rope); `key_states` may then be None — the
    normed pre-RoPE keys never exist in HBM on that route.  Same bits as the K route.

    Reference: compression/compactor.py:113-221: X = K.PHI, per chunk centre X,
    G = X^T X + reg*I, score_i = ||x_i V S^-1/2||^2 (== x_i^T G^-1 x_i), clamp
    >= 0, optional z-score per chunk across all heads (:215-220,224-269).
    Returned in K's dtype like the reference.
    """
    if sketch_rows is not None:
        return _leverage_from_sketch(sketch_rows, context_lens, regularizer, normalize, chunk_size)
    Ntok, H, D = key_states.shape
    assert key_states.stride(-1) == 1
    sketch = PHI.shape[1]
    assert PHI.shape[0] == D and PHI.dtype == key_states.dtype
    if chunk_size > 0:
        _, chunk_lens = split_into_chunks(context_lens, chunk_size)
    else:
        chunk_lens = list(context_lens)
    assert sum(chunk_lens) == Ntok
    dev = key_states.device
    offsets = _chunk_offsets(tuple(chunk_lens), dev)
    out = torch.empty((Ntok, H), dtype=key_states.dtype, device=dev)
    if Ntok == 0:
        return out
    n_chunks = len(chunk_lens)
    L = N.lib()
    ws_bytes = L.cvllm_leverage_scores_workspace_bytes(Ntok, H, sketch, n_chunks)
    ws = N.workspace(dev, ws_bytes, "leverage")
    phi = PHI.contiguous()
    rc = L.cvllm_leverage_scores(
        N.ptr(key_states), key_states.stride(0), key_states.stride(1), N.ptr(offsets), n_chunks,
        max(chunk_lens), N.ptr(phi), N.ptr(out), Ntok, H, D, sketch, float(regularizer),
        int(bool(normalize)), N.dtype_code(key_states.dtype), N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_leverage_scores")
    return out


def _leverage_from_sketch(sketch_rows, context_lens, regularizer, normalize, chunk_size):
    Ntok, H, sketch = sketch_rows.shape
    assert sketch_rows.is_contiguous()
    if chunk_size > 0:
        _, chunk_lens = split_into_chunks(context_lens, chunk_size)
    else:
        chunk_lens = list(context_lens)
    assert sum(chunk_lens) == Ntok
    dev = sketch_rows.device
    offsets = _chunk_offsets(tuple(chunk_lens), dev)
    out = torch.empty((Ntok, H), dtype=sketch_rows.dtype, device=dev)
    if Ntok == 0:
        return out
    n_chunks = len(chunk_lens)
    L = N.lib()
    ws_bytes = L.cvllm_leverage_scores_workspace_bytes(Ntok, H, sketch, n_chunks)
    ws = N.workspace(dev, ws_bytes, "leverage")
    rc = L.cvllm_leverage_scores_from_sketch(
        N.ptr(sketch_rows), N.ptr(offsets), n_chunks, max(chunk_lens), None, N.ptr(out), Ntok, H, sketch,
        float(regularizer), int(bool(normalize)), N.dtype_code(sketch_rows.dtype), N.ptr(ws), ws.numel(),
        N.stream(dev),
    )
    N.check(rc, "cvllm_leverage_scores_from_sketch")
    return out


def non_causal_attn_scores(
    q: torch.Tensor,  # [N, HQ, D]
    k: torch.Tensor,  # [N, HKV, D]
    v: torch.Tensor,  # [N, HKV, D] (unused, kept for signature parity)
    cu_seqlens_qk: torch.Tensor,  # [B + 1]
    max_seqlen_qk: int,
    chunk_size: int,
    sm_scale: float = None,
    normalize: bool = True,
    context_lens: Optional[List[int]] = None,
    protected_first_tokens: Optional[List[int]] = None,
    protected_last_tokens: Optional[List[int]] = None,
    *,
    accum_scores: torch.Tensor = None,  # [N, HKV]
    accum_blending: float = None,
) -> torch.Tensor:
    """Chunk-local non-causal attention mass per key -> fp32 [N, HKV].

    Reference: compression/compactor.py:338-599: inside every chunk_size-token
    chunk, softmax over the chunk's keys for all G*chunk query rows, column
    sums (short chunks add 1/chunk_size per missing row, :477-479); z-score per
    sequence; + accum_blending*accum_scores; protected first/last tokens <- +inf.
    """
    assert q.ndim == 3 and k.ndim == 3
    assert q.shape[0] == k.shape[0] and q.shape[-1] == k.shape[-1]
    Ntok, HQ, D = q.shape
    HKV = k.shape[1]
    assert HQ % HKV == 0, "Number of query heads must divide number of KV heads"
    assert (D & (D - 1)) == 0, "D must be a power of two"
    assert q.stride(-1) == 1 and k.stride(-1) == 1, "last dim must be contiguous"
    assert q.stride(1) == D and k.stride(1) == D
    B = cu_seqlens_qk.numel() - 1
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(D)
    dev = q.device
    cu = cu_seqlens_qk.to(device=dev, dtype=torch.int32).contiguous()
    out = torch.empty((Ntok, HKV), dtype=torch.float32, device=dev)
    if Ntok == 0:
        return out
    pf = pl = None
    if protected_first_tokens is not None or protected_last_tokens is not None:
        pf_l = list(protected_first_tokens) if protected_first_tokens is not None else [0] * B
        pl_l = list(protected_last_tokens) if protected_last_tokens is not None else [0] * B
        both = _protected_counts(tuple(pf_l), tuple(pl_l), dev)
        pf, pl = both[:B], both[B:]
    blend = 1.0 if accum_blending is None else float(accum_blending)
    acc_dtype, acc_stride = 2, 0
    if accum_scores is not None:
        assert accum_scores.shape == (Ntok, HKV) and accum_scores.stride(1) == 1
        acc_dtype = 2 if accum_scores.dtype == torch.float32 else N.dtype_code(accum_scores.dtype)
        acc_stride = accum_scores.stride(0)
    L = N.lib()
    ws_bytes = L.cvllm_chunk_attn_scores_workspace_bytes(B, Ntok, HKV, int(chunk_size))
    ws = N.workspace(dev, ws_bytes, "chunk_attn")
    rc = L.cvllm_chunk_attn_scores(
        N.ptr(q), N.ptr(k), q.stride(0), k.stride(0), N.ptr(cu), B, Ntok, HQ, HKV, D,
        int(chunk_size), int(max_seqlen_qk), float(sm_scale), int(bool(normalize)),
        N.ptr(accum_scores), acc_dtype, acc_stride, blend, N.ptr(pf), N.ptr(pl), N.ptr(out),
        N.dtype_code(q.dtype), N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_chunk_attn_scores")
    return out
