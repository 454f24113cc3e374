"""Oracle: prefill and decode attention in fp32 on the CPU."""
from __future__ import annotations

import math

import torch

from .store import head_rows


def _attend(qh: torch.Tensor, K: torch.Tensor, V: torch.Tensor, scale: float, mask=None) -> torch.Tensor:
    """softmax(scale * q K^T + mask) V in fp32; qh [M, D], K/V [T, D]."""
    s = (qh.float() @ K.float().T) * scale
    if mask is not None:
        s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    return p @ V.float()


def prefill_attention(q, k, v, k_cache, v_cache, seq_lens_bh, page_table, batch_mapping, cu_seqlens_q,
                      HKV, PS, sm_scale=None) -> torch.Tensor:
    """attention/sparse_varlen_kernel.py:11-197,265-519.

    For sequence b, kv head h, query head g of h: keys/values are the cached
    prefix of (b,h) (length seq_lens_bh[b,h], rows through the page table)
    followed by the new tokens of b; new token i sees the whole prefix and new
    tokens <= i.  Returns fp32 [N, HQ, D].  (The reference masks with -1e6
    rather than -inf, :386,444,489 — identical after exp in fp32.)
    """
    N, HQ, D = q.shape
    G = HQ // HKV
    scale = 1.0 / math.sqrt(D) if sm_scale is None else sm_scale
    out = torch.zeros((N, HQ, D), dtype=torch.float32)
    B = cu_seqlens_q.numel() - 1
    for b in range(B):
        s, e = int(cu_seqlens_q[b]), int(cu_seqlens_q[b + 1])
        La = e - s
        if La <= 0:
            continue
        bt = int(batch_mapping[b])
        for h in range(HKV):
            Lc = int(seq_lens_bh[b, h])
            rows = torch.from_numpy(head_rows(page_table, bt, h, 0, Lc, PS))
            K = torch.cat([k_cache[rows], k[s:e, h]], dim=0)
            V = torch.cat([v_cache[rows], v[s:e, h]], dim=0)
            mask = torch.ones((La, Lc + La), dtype=torch.bool)
            mask[:, Lc:] = torch.tril(torch.ones((La, La), dtype=torch.bool))
            for g in range(G):
                hq = h * G + g
                out[s:e, hq] = _attend(q[s:e, hq], K, V, scale, mask)
    return out


def decode_attention(q, k_cache, v_cache, seq_lens_bh, page_table, batch_mapping, HKV, PS, sm_scale=None):
    """attention/sparse_decode_kernel.py:10-165,247-435: one query per sequence against the
    L = seq_lens_bh[b,h] cached rows of each kv head.  Rows with L == 0 are zeros here
    (uninitialised in the reference, :281-283).  Returns fp32 [B, HQ, D]."""
    B, HQ, D = q.shape
    G = HQ // HKV
    scale = 1.0 / math.sqrt(D) if sm_scale is None else sm_scale
    out = torch.zeros((B, HQ, D), dtype=torch.float32)
    for b in range(B):
        bt = int(batch_mapping[b])
        for h in range(HKV):
            L = int(seq_lens_bh[b, h])
            if L == 0:
                continue
            rows = torch.from_numpy(head_rows(page_table, bt, h, 0, L, PS))
            out[b, h * G:(h + 1) * G] = _attend(q[b, h * G:(h + 1) * G], k_cache[rows], v_cache[rows], scale)
    return out
