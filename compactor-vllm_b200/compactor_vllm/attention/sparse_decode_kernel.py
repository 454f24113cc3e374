"""Decode-time head-sparse attention (reference: attention/sparse_decode_kernel.py).

`head_sparse_decode_attention` keeps the reference signature; `key_split` is
accepted and ignored except for validation — the sm_100a kernel balances work
over 128-token units of every (batch, kv-head) itself and merges in-kernel, so
the result does not depend on it.  `num_splits_heuristic` is kept
because the reference's model runner imports it (core/model_runner.py:7,503).
"""
from __future__ import annotations

import functools
import math

import torch
from compactor_vllm import _native as N


def head_sparse_decode_attention(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    seq_lens_bh: torch.Tensor,
    global_page_table: torch.Tensor,
    batch_mapping: torch.Tensor,
    HKV: int,
    PAGE_SIZE: int,
    sm_scale: float = None,
    key_split: int = None,
):
    """q [B,HQ,D] or [B,HQ,1,D]; k, v flat caches [CACHE_SIZE, D]; returns [B,HQ,D].

    Reference: attention/sparse_decode_kernel.py:10-165.  CUDA-graph capturable:
    no host sync (the reference syncs on seq_lens_bh.max() when key_split is
    None, :94-103), fixed grid, scratch from torch's caching allocator.  Rows
    whose head is empty are zeros (uninitialised in the reference, :281-283).
    """
    if q.ndim != 3:
        assert q.ndim == 4
        B, HQ, S, D = q.shape
        assert S == 1, "head_sparse_decode_attention only supports q_len=1"
        q = q.squeeze(-2)
    B, HQ, D = q.shape
    assert PAGE_SIZE % 32 == 0, "PAGE_SIZE must be divisible by 32"
    assert HQ % HKV == 0, "HQ must be divisible by H_kv"
    assert B <= 32767, "too many batches"
    assert global_page_table.shape[1] == HKV
    assert q.is_contiguous()
    assert (D & (D - 1)) == 0, "D must be a power of 2"
    assert k.is_contiguous() and v.is_contiguous() and k.dtype == q.dtype == v.dtype
    assert global_page_table.is_contiguous() and global_page_table.dtype == torch.int32
    assert batch_mapping.dtype == torch.int32 and batch_mapping.is_contiguous()
    seq_lens_bh = seq_lens_bh.to(torch.int32).contiguous()
    sm_scale = 1 / math.sqrt(D) if sm_scale is None else sm_scale
    dev = q.device
    out = torch.empty_like(q)
    if B == 0:
        return out
    L = N.lib()
    ws_bytes = L.cvllm_decode_attention_workspace_bytes(B, HQ, HKV, D, 0)
    # one workspace per (B, HQ, HKV): the ticket region's size depends on B*HKV, and the "zero on
    # entry, zero on exit" contract only holds while the layout stays the same
    ws = N.workspace(dev, ws_bytes, f"decode_attn:{B}:{HQ}:{HKV}", zero=True)
    rc = L.cvllm_decode_attention(
        N.ptr(q), N.ptr(out), N.ptr(k), N.ptr(v), N.ptr(seq_lens_bh), N.ptr(global_page_table),
        N.ptr(batch_mapping), N.ptr(ws), ws.numel(), B, HQ, HKV, D, int(PAGE_SIZE),
        global_page_table.shape[-1], float(sm_scale), N.dtype_code(q.dtype), 0, N.stream(dev),
    )
    N.check(rc, "cvllm_decode_attention")
    return out


def decode_append_and_attend(
    q: torch.Tensor,  # [B, HQ, D]
    key: torch.Tensor,  # [B, HKV, D] new row per (b, h)
    value: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    bh_seq_lens: torch.Tensor,  # the cache's [Bmax+1, HKV] table, updated in place (+1 per live (b, h))
    global_page_table: torch.Tensor,
    batch_mapping: torch.Tensor,
    HKV: int,
    PAGE_SIZE: int,
    sm_scale: float = None,
    reserved_batch: int = 0,
):
    """The decode branch of layers/attention.py:133-160 as one native call (cvllm_decode_step):
    decode_store_kv + head_sparse_decode_attention, both reading / bumping `bh_seq_lens` through
    `batch_mapping` — no gathered length copy, no index_copy_ back.  Graph-capturable, two kernel nodes."""
    if q.ndim == 4:
        assert q.shape[2] == 1
        q = q.squeeze(-2)
    B, HQ, D = q.shape
    assert key.shape == value.shape == (B, HKV, D)
    assert key.stride(-1) == 1 and value.stride(-1) == 1 and q.is_contiguous()
    assert k_cache.is_contiguous() and v_cache.is_contiguous() and k_cache.dtype == q.dtype == key.dtype
    assert bh_seq_lens.dtype == torch.int32 and bh_seq_lens.is_contiguous() and bh_seq_lens.shape[-1] == HKV
    assert global_page_table.dtype == torch.int32 and global_page_table.is_contiguous()
    assert batch_mapping.dtype == torch.int32 and batch_mapping.is_contiguous()
    assert HQ % HKV == 0 and PAGE_SIZE % 32 == 0
    sm_scale = 1 / math.sqrt(D) if sm_scale is None else sm_scale
    dev = q.device
    out = torch.empty_like(q)
    if B == 0:
        return out
    L = N.lib()
    ws_bytes = L.cvllm_decode_attention_workspace_bytes(B, HQ, HKV, D, 0)
    ws = N.workspace(dev, ws_bytes, f"decode_attn:{B}:{HQ}:{HKV}", zero=True)
    rc = L.cvllm_decode_step(
        N.ptr(key), N.ptr(value), key.stride(0), key.stride(1), value.stride(0), value.stride(1),
        N.ptr(q), N.ptr(out), N.ptr(k_cache), N.ptr(v_cache), N.ptr(bh_seq_lens), N.ptr(global_page_table),
        N.ptr(batch_mapping), N.ptr(ws), ws.numel(), B, HQ, HKV, D, int(PAGE_SIZE),
        global_page_table.shape[-1], float(sm_scale), N.dtype_code(q.dtype), int(reserved_batch), 0,
        N.stream(dev),
    )
    N.check(rc, "cvllm_decode_step")
    return out


@functools.lru_cache(maxsize=128)
def num_splits_heuristic(total_mblocks: int, max_seq_len: int, num_sms: int, max_splits: int) -> int:
    """Split count the reference engine fixes at graph-capture time (:168-192).

    1 split if the (b, kv-head) blocks already fill >= 80 % of the SMs or the
    context is <= 1024; otherwise the smallest split count whose wave efficiency
    reaches 75 % of the best, each split longer than 512 keys.
    """
    if total_mblocks >= 0.8 * num_sms or max_seq_len <= 1024:
        return 1
    effs = []
    for s in range(1, min(max_splits, num_sms) + 1):
        if max_seq_len / s <= 512:
            break
        waves = float(total_mblocks * s) / float(num_sms)
        effs.append(waves / math.ceil(waves) if waves > 0 else 0.0)
    best = max(effs, default=0.0)
    for i, e in enumerate(effs, start=1):
        if e >= 0.75 * best:
            return i
    return 1
