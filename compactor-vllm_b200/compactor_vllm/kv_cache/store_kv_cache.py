"""Paged-cache store wrappers — same names, keyword arguments and assertion
behaviour as the reference's kv_cache/store_kv_cache.py, served by
libcompactor_sm100.so (csrc/store_kv.cu, csrc/select.cu).
"""
from __future__ import annotations

import torch
from compactor_vllm import _native as N
from compactor_vllm.config.constants import TRITON_RESERVED_BATCH as _TRITON_RESERVED_BATCH


def _i32(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.int32 else t.to(torch.int32)


def prefill_store_topk_kv(
    *,
    new_keys: torch.Tensor,  # [N_total, H, D]
    new_vals: torch.Tensor,  # [N_total, H, D]
    indices_topk: torch.Tensor,  # [B, MAX_SEL] (global flattened token*H + head)
    num_tokens_to_retain: torch.Tensor,  # [B] int32
    page_table: torch.Tensor,  # [B_total, H, N_LOGICAL_PAGES_MAX] int32
    batch_mapping: torch.Tensor,  # [B] int32
    bh_lens: torch.Tensor,  # [B, H] int32 (contiguous), UPDATED
    k_cache: torch.Tensor,  # [N_PAGES * PAGE_SIZE, D]
    v_cache: torch.Tensor,
    PAGE_SIZE: int,
    PAD_TO_PAGE_SIZE: bool = True,
    cu_seqlens_k: torch.Tensor | None = None,
    K_TILE: int = 16,  # accepted for drop-in compatibility; tiling is internal
    TRITON_RESERVED_BATCH: int = None,
):
    """Reference: kv_cache/store_kv_cache.py:81-175 (K4 :9-78, K5 :178-248).

    Differences kept deliberate: rows of a head land in ascending token order
    (the reference's atomicAdd order is nondeterministic), and `cu_seqlens_k` is
    needed in both modes here (it maps packed tokens to sequences); when it is
    None and B == 1 it is synthesised.
    """
    assert new_keys.shape == new_vals.shape
    N_total, H, D = new_keys.shape
    B = indices_topk.shape[0]
    assert page_table.shape[1] == H
    assert bh_lens.shape == (B, H)
    assert new_keys.device == k_cache.device == v_cache.device
    assert page_table.is_contiguous(), "page table must be contiguous."
    assert bh_lens.is_contiguous(), "bh_lens must be contiguous."
    assert batch_mapping.is_contiguous(), "batch mapping must be contiguous."
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    assert new_keys.stride(-1) == 1 and new_vals.stride(-1) == 1, (
        "new_keys/new_vals last dim must be contiguous."
    )
    assert (D & (D - 1)) == 0, "D must be a power of 2"
    assert bh_lens.dtype == torch.int32, "bh_lens is updated in place and must be int32"
    if PAD_TO_PAGE_SIZE:
        assert cu_seqlens_k is not None
    if cu_seqlens_k is None:
        assert B == 1, "cu_seqlens_k is required for B > 1"
        cu_seqlens_k = torch.tensor([0, N_total], dtype=torch.int32, device=new_keys.device)
    page_table = _i32(page_table)
    batch_mapping = _i32(batch_mapping)
    cu_seqlens_k = _i32(cu_seqlens_k)
    num_tokens_to_retain = _i32(num_tokens_to_retain).contiguous()
    if indices_topk.dtype not in (torch.int32, torch.int64):
        indices_topk = indices_topk.to(torch.int64)
    indices_topk = indices_topk.contiguous()
    if TRITON_RESERVED_BATCH is None:
        TRITON_RESERVED_BATCH = _TRITON_RESERVED_BATCH
    MAX_SEL = indices_topk.shape[-1]
    dev = new_keys.device
    L = N.lib()
    ws_bytes = L.cvllm_store_topk_workspace_bytes(B, H, N_total, MAX_SEL)
    ws = N.workspace(dev, ws_bytes, "store_topk")
    sk_n, sk_h, _ = new_keys.stride()
    sv_n, sv_h, _ = new_vals.stride()
    rc = L.cvllm_prefill_store_topk_idx(
        N.ptr(new_keys), N.ptr(new_vals), sk_n, sk_h, sv_n, sv_h,
        N.ptr(indices_topk), indices_topk.element_size(), int(MAX_SEL),
        N.ptr(num_tokens_to_retain), N.ptr(cu_seqlens_k), N.ptr(batch_mapping), N.ptr(bh_lens),
        N.ptr(page_table), N.ptr(k_cache), N.ptr(v_cache),
        B, N_total, H, D, int(PAGE_SIZE), page_table.shape[2], new_keys.element_size(),
        int(bool(PAD_TO_PAGE_SIZE)), int(TRITON_RESERVED_BATCH),
        N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_prefill_store_topk_idx")


def prefill_store_all_kv(
    *,
    new_keys: torch.Tensor,
    new_values: torch.Tensor,  # [N, H_kv, D]
    cu_seqlens_k: torch.Tensor,  # [B + 1] int32
    max_seqlen_k: int,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    page_table: torch.Tensor,  # [B_total, H_kv, N_LOGICAL_PAGES_MAX] int32
    bh_lens: torch.Tensor,  # [B, H_kv] int32 (UPDATED)
    batch_mapping: torch.Tensor,  # [B] int32
    PAGE_SIZE: int,
    K_TILE: int = 32,  # accepted for drop-in compatibility
):
    """Reference: kv_cache/store_kv_cache.py:322-371 (kernel :251-319)."""
    assert new_keys.stride(-1) == 1 and new_values.stride(-1) == 1, "last dim must be contiguous"
    assert page_table.is_contiguous(), "page table must be contiguous"
    assert bh_lens.is_contiguous(), "bh_lens must be contiguous"
    assert batch_mapping.is_contiguous(), "batch mapping must be contiguous"
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    Ntok, HKV, D = new_keys.shape
    B = batch_mapping.shape[0]
    assert (D & (D - 1)) == 0, "D must be a power of 2"
    assert bh_lens.dtype == torch.int32
    sk_n, sk_h, _ = new_keys.stride()
    sv_n, sv_h, _ = new_values.stride()
    dev = new_keys.device
    rc = N.lib().cvllm_prefill_store_all(
        N.ptr(new_keys), N.ptr(new_values), sk_n, sk_h, sv_n, sv_h,
        N.ptr(_i32(cu_seqlens_k)), N.ptr(_i32(batch_mapping)), N.ptr(bh_lens),
        N.ptr(_i32(page_table)), N.ptr(k_cache), N.ptr(v_cache),
        B, Ntok, HKV, D, int(PAGE_SIZE), page_table.shape[-1], new_keys.element_size(),
        N.stream(dev),
    )
    N.check(rc, "cvllm_prefill_store_all")


def decode_store_kv(
    *,
    key: torch.Tensor,  # [B, HKV, D]
    value: torch.Tensor,  # [B, HKV, D]
    batch_mapping: torch.Tensor,  # [B] int32
    bh_lens: torch.Tensor,  # [B, HKV] or flattened [B*HKV] int32
    page_table: torch.Tensor,  # [B_total, HKV, N_LOGICAL_PAGES_MAX] int32
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,  # [N_PAGES*PAGE_SIZE, D]
    PAGE_SIZE: int,
    TRITON_RESERVED_BATCH: int = None,
):
    """Reference: kv_cache/store_kv_cache.py:419-466 (kernel :374-416)."""
    assert key.shape == value.shape and key.ndim == 3, "key/value must be [B, HKV, D]"
    B, HKV, D = key.shape
    assert key.stride(-1) == 1 and value.stride(-1) == 1, "key/value last dim must be contiguous."
    assert page_table.is_contiguous(), "page table must be contiguous."
    assert bh_lens.is_contiguous(), "bh_lens must be contiguous."
    assert batch_mapping.is_contiguous(), "batch mapping must be contiguous."
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    assert (D & (D - 1)) == 0, "D must be a power of 2"
    assert bh_lens.dtype == torch.int32 and batch_mapping.dtype == torch.int32
    assert page_table.dtype == torch.int32
    sk_b, sk_h, _ = key.stride()
    sv_b, sv_h, _ = value.stride()
    dev = key.device
    rc = N.lib().cvllm_decode_store_kv(
        N.ptr(key), N.ptr(value), sk_b, sk_h, sv_b, sv_h,
        N.ptr(batch_mapping), N.ptr(bh_lens), N.ptr(page_table), N.ptr(k_cache), N.ptr(v_cache),
        int(batch_mapping.shape[0]), HKV, D, int(PAGE_SIZE), page_table.shape[2], key.element_size(),
        int(TRITON_RESERVED_BATCH if TRITON_RESERVED_BATCH is not None else _TRITON_RESERVED_BATCH),
        N.stream(dev),
    )
    N.check(rc, "cvllm_decode_store_kv")
