"""Compression plugin interface + scores -> kept rows pipeline.

Reference: compression/common.py:9-243.  `scores_to_retain_indices` returns the
reference's full descending ranking (int64 global flat ids) from a stable
radix sort; `extract_and_store_top_kv` — the call the engine makes — skips the
sort entirely and goes scores -> radix-select -> compaction in three launches
whenever the ranking it would have built is complete (top_k >= max_k_len*H,
which is what the engine passes, utils/arguments.py:208-211).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Optional

import torch
from compactor_vllm import _native as N
from compactor_vllm.config.constants import TRITON_RESERVED_BATCH as _RESERVED
from compactor_vllm.kv_cache.store_kv_cache import prefill_store_topk_kv


class BaseCompressionMethod(ABC):
    """Two optional scoring phases around RoPE; each returns [N, HKV] scores or None.

    Reference: compression/common.py:9-101.
    """

    @staticmethod
    @abstractmethod
    def pre_rope_scoring(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, context) -> Optional[torch.Tensor]:
        """Scores from pre-RoPE q [N,HQ,D] / k,v [N,HKV,D]; None if this phase is a no-op."""

    @staticmethod
    @abstractmethod
    def post_rope_scoring(
        q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, pre_rope_scores: Optional[torch.Tensor], context
    ) -> Optional[torch.Tensor]:
        """Final scores from post-RoPE q/k (may refine pre_rope_scores); None disables compression."""


class NoCompression(BaseCompressionMethod):
    """Reference: compression/common.py:104-123."""

    @staticmethod
    def pre_rope_scoring(q, k, v, context) -> Optional[torch.Tensor]:
        return None

    @staticmethod
    def post_rope_scoring(q, k, v, pre_rope_scores, context) -> Optional[torch.Tensor]:
        return pre_rope_scores


def scores_to_retain_indices(
    scores: torch.Tensor,
    cu_seqlens_k: torch.Tensor,
    max_k_len: int,
    top_k: int,
    H: int,
    padding: float = -float("inf"),
) -> torch.Tensor:
    """Per-sequence descending ranking of (token, head) scores -> int64 [B, k_eff].

    Reference: compression/common.py:171-243 (pad to [B, max_k_len*H] with -inf,
    torch.topk(sorted=True), add cu[b]*H).  Order: score descending, NaN first
    (torch.topk's order), ties by ascending flat index — torch leaves tie order
    unspecified, this pins it.  Only padding=-inf is supported.
    """
    assert padding == -float("inf"), "only -inf padding is supported"
    assert scores.ndim == 2 and scores.shape[1] == H and scores.dtype == torch.float32
    assert scores.stride(1) == 1
    B, dev = cu_seqlens_k.numel() - 1, scores.device
    max_k_len = int(max_k_len)
    k_eff = min(int(top_k), max_k_len * H)
    out = torch.empty((B, k_eff), dtype=torch.int64, device=dev)
    if k_eff == 0 or B == 0:
        return out
    cu = cu_seqlens_k.to(device=dev, dtype=torch.int32).contiguous()
    L = N.lib()
    ws_bytes = L.cvllm_rank_scores_workspace_bytes(B, H, max_k_len)
    ws = N.workspace(dev, ws_bytes, "rank")
    rc = L.cvllm_rank_scores(
        N.ptr(scores), scores.stride(0), N.ptr(cu), B, H, max_k_len, int(top_k),
        N.ptr(out), N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_rank_scores")
    return out


def extract_and_store_top_kv(
    scores: torch.Tensor,
    cu_seqlens_k: torch.Tensor,
    max_k_len: int,
    top_k: int,
    H: int,
    new_keys: torch.Tensor,  # [N_total, H, D]
    new_vals: torch.Tensor,  # [N_total, H, D]
    num_tokens_to_retain: torch.Tensor,  # [B] int32
    page_table: torch.Tensor,  # [B_total, H, N_LOGICAL_PAGES_MAX] int32
    batch_mapping: torch.Tensor,  # [B] int32
    bh_lens: torch.Tensor,  # [B, H] int32 (contiguous), UPDATED
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    PAGE_SIZE: int,
    PAD_TO_PAGE_SIZE: bool = True,
    K_TILE: int = 16,
    padding: float = -float("inf"),
):
    """Keep the top num_tokens_to_retain[b] (token, head) pairs of every sequence.

    Reference: compression/common.py:126-168.  Fused path when the ranking
    would be complete; otherwise rank + index store (identical kept sets).
    """
    max_k_len = int(max_k_len)
    if int(top_k) < max_k_len * H or scores.dtype != torch.float32 or scores.stride(1) != 1:
        idx = scores_to_retain_indices(
            scores.float().contiguous(), cu_seqlens_k=cu_seqlens_k, max_k_len=max_k_len,
            top_k=top_k, H=H, padding=padding,
        )
        prefill_store_topk_kv(
            new_keys=new_keys, new_vals=new_vals, indices_topk=idx,
            num_tokens_to_retain=num_tokens_to_retain, page_table=page_table,
            batch_mapping=batch_mapping, bh_lens=bh_lens, k_cache=k_cache, v_cache=v_cache,
            cu_seqlens_k=cu_seqlens_k, PAGE_SIZE=PAGE_SIZE, PAD_TO_PAGE_SIZE=PAD_TO_PAGE_SIZE,
            K_TILE=K_TILE,
        )
        return
    assert padding == -float("inf")
    assert new_keys.shape == new_vals.shape
    N_total, Hk, D = new_keys.shape
    assert Hk == H and scores.shape == (N_total, H)
    B = cu_seqlens_k.numel() - 1
    assert bh_lens.shape == (B, H) and bh_lens.is_contiguous() and bh_lens.dtype == torch.int32
    assert page_table.is_contiguous() and page_table.dtype == torch.int32
    assert batch_mapping.is_contiguous() and k_cache.is_contiguous() and v_cache.is_contiguous()
    assert new_keys.stride(-1) == 1 and new_vals.stride(-1) == 1
    dev = new_keys.device
    cu = cu_seqlens_k.to(device=dev, dtype=torch.int32).contiguous()
    bm = batch_mapping.to(torch.int32)
    retain = num_tokens_to_retain.to(device=dev, dtype=torch.int32).contiguous()
    L = N.lib()
    ws_bytes = L.cvllm_select_store_workspace_bytes(B, H, N_total)
    ws = N.workspace(dev, ws_bytes, "select_store")
    sk_n, sk_h, _ = new_keys.stride()
    sv_n, sv_h, _ = new_vals.stride()
    rc = L.cvllm_select_store_topk(
        N.ptr(scores), scores.stride(0), N.ptr(new_keys), N.ptr(new_vals), sk_n, sk_h, sv_n, sv_h,
        N.ptr(retain), N.ptr(cu), N.ptr(bm), N.ptr(bh_lens), N.ptr(page_table),
        N.ptr(k_cache), N.ptr(v_cache), B, N_total, H, D, int(PAGE_SIZE), page_table.shape[2],
        new_keys.element_size(), int(bool(PAD_TO_PAGE_SIZE)), int(_RESERVED),
        N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_select_store_topk")
