"""Prefill: causal head-sparse varlen attention over [paged prefix || new K/V].

Reference: attention/sparse_varlen_kernel.py:11-197 (Triton kernel :265-519).
Served by the tcgen05/TMEM/TMA kernel in csrc/prefill_attn.cu.
"""
from __future__ import annotations

import math

import torch
from compactor_vllm import _native as N


def causal_sparse_varlen_with_cache(
    q,
    k,
    v,
    k_cache,
    v_cache,
    seq_lens_bh,
    global_page_table,
    batch_mapping,
    cu_seqlens_q,
    max_seqlen_q: int,
    max_seqlen_k_cache: int,
    HKV: int,
    PAGE_SIZE: int,
    sm_scale=None,
):
    """q [N,HQ,D], k/v [N,HKV,D] (last dim contiguous, any token stride) -> [N,HQ,D].

    For every sequence b, kv head h and query head g of h:
    softmax(scale * q . [Kcache_bh || Kapp_b]^T, causal on the appended part) .
    [Vcache_bh || Vapp_b]; the prefix length is per (b, h).
    `max_seqlen_k_cache` is accepted for signature parity (the reference only
    uses it as an autotune key, :164).
    """
    assert q.ndim == 3, "q should be [N, HQ, D]"
    Ntok, HQ, D = q.shape
    assert (D & (D - 1)) == 0, "D must be power of two"
    B = cu_seqlens_q.numel() - 1
    assert B > 0
    assert HQ % HKV == 0, "Number of query heads must divide number of keys heads"
    assert k.shape == (Ntok, HKV, D) and v.shape == (Ntok, HKV, D)
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, (
        "final dimension must be contiguous"
    )
    assert q.stride(1) == D and k.stride(1) == D and v.stride(1) == D, "heads must be packed"
    assert k_cache.shape[1] == D and v_cache.shape[1] == D
    CACHE_SIZE = k_cache.shape[0]
    assert v_cache.shape[0] == CACHE_SIZE
    assert PAGE_SIZE > 0 and CACHE_SIZE % PAGE_SIZE == 0
    assert k_cache.is_contiguous() and v_cache.is_contiguous()
    assert q.dtype == k.dtype == v.dtype == k_cache.dtype == v_cache.dtype
    dev = q.device
    cu = cu_seqlens_q.to(dtype=torch.int32, device=dev).contiguous()
    lens = seq_lens_bh.to(dtype=torch.int32, device=dev).contiguous()
    bm = batch_mapping.to(dtype=torch.int32, device=dev).contiguous()
    pt = global_page_table
    assert pt.dtype == torch.int32 and pt.is_contiguous()
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(D)
    out = torch.empty((Ntok, HQ, D), dtype=q.dtype, device=dev)
    rc = N.lib().cvllm_prefill_attention(
        N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(out), N.ptr(k_cache), N.ptr(v_cache),
        N.ptr(lens), N.ptr(pt), N.ptr(bm), N.ptr(cu),
        B, Ntok, HQ, HKV, D, int(PAGE_SIZE), pt.shape[-1], CACHE_SIZE,
        q.stride(0), k.stride(0), v.stride(0), out.stride(0), int(max_seqlen_q),
        float(sm_scale), N.dtype_code(q.dtype), N.stream(dev),
    )
    N.check(rc, "cvllm_prefill_attention")
    return out
