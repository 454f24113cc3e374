"""Test-only builders: paged caches with scrambled pages, non-identity slots and (optionally)
NaN-poisoned unused rows, plus a device-side fp32 restatement for sizes the CPU oracle is too
slow for (it is checked against the oracle in test_gpu_prefill.py at small sizes)."""
from __future__ import annotations

import math

import torch


def make_paged(B, HKV, D, PS, lens_bh, dtype, device, seed=0, poison=False, extra_pages=1, n_extra_slots=2):
    """lens_bh: [B][HKV] valid rows per head.  Returns (k_cache, v_cache, page_table, batch_mapping).

    Physical pages are a seeded permutation, slot 0 is reserved and the B sequences sit in
    shuffled slots 1.. ; rows past each head's length hold NaN when poison=True (the real cache is
    torch.empty — kernels must never let them reach a multiply)."""
    g = torch.Generator().manual_seed(seed)
    lens = torch.as_tensor(lens_bh, dtype=torch.int64).view(B, HKV)
    NLP = int((int(lens.max()) + PS - 1) // PS + extra_pages)
    n_slots = B + n_extra_slots
    n_pages = n_slots * HKV * NLP
    page_table = torch.randperm(n_pages, generator=g).to(torch.int32).view(n_slots, HKV, NLP)
    batch_mapping = (torch.randperm(n_slots - 1, generator=g)[:B] + 1).to(torch.int32)
    k_cache = torch.randn((n_pages * PS, D), generator=g).to(dtype)
    v_cache = torch.randn((n_pages * PS, D), generator=g).to(dtype)
    if poison:
        valid = torch.zeros(n_pages * PS, dtype=torch.bool)
        for b in range(B):
            for h in range(HKV):
                L = int(lens[b, h])
                pos = torch.arange(L)
                rows = page_table[int(batch_mapping[b]), h, pos // PS].long() * PS + pos % PS
                valid[rows] = True
        k_cache[~valid] = float("nan")
        v_cache[~valid] = float("nan")
    return (k_cache.to(device), v_cache.to(device), page_table.to(device).contiguous(),
            batch_mapping.to(device))


def fused_qkv(N, HQ, HKV, D, dtype, device, seed=0):
    """q, k, v as strided views into one [N, (HQ+2HKV)*D] buffer, as the models hand them over."""
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn((N, (HQ + 2 * HKV) * D), generator=g).to(dtype).to(device)
    q = qkv[:, : HQ * D].view(N, HQ, D)
    k = qkv[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D)
    v = qkv[:, (HQ + HKV) * D:].view(N, HKV, D)
    return q, k, v


def head_rows_dev(page_table, bt, h, L, PS):
    pos = torch.arange(L, device=page_table.device)
    return page_table[bt, h, pos // PS].long() * PS + pos % PS


@torch.no_grad()
def torch_prefill_fp32(q, k, v, k_cache, v_cache, seq_lens_bh, page_table, batch_mapping, cu, HKV, PS,
                       sm_scale=None, row_block=2048):
    """Same math as oracle.attention.prefill_attention, on the tensors' device, blocked over rows."""
    N, HQ, D = q.shape
    G = HQ // HKV
    scale = 1.0 / math.sqrt(D) if sm_scale is None else sm_scale
    out = torch.zeros((N, HQ, D), dtype=torch.float32, device=q.device)
    cu_l = cu.tolist()
    lens = seq_lens_bh.tolist()
    bm = batch_mapping.tolist()
    for b in range(len(cu_l) - 1):
        s, e = cu_l[b], cu_l[b + 1]
        La = e - s
        if La <= 0:
            continue
        for h in range(HKV):
            Lc = lens[b][h]
            rows = head_rows_dev(page_table, bm[b], h, Lc, PS)
            K = torch.cat([k_cache[rows], k[s:e, h]], 0).float()
            V = torch.cat([v_cache[rows], v[s:e, h]], 0).float()
            kpos = torch.arange(Lc + La, device=q.device)
            for r0 in range(0, La, row_block):
                r1 = min(La, r0 + row_block)
                qpos = torch.arange(r0, r1, device=q.device)
                mask = kpos[None, :] <= (Lc + qpos)[:, None]
                for g_ in range(G):
                    hq = h * G + g_
                    sc = (q[s + r0:s + r1, hq].float() @ K.T) * scale
                    sc = sc.masked_fill(~mask, float("-inf"))
                    out[s + r0:s + r1, hq] = torch.softmax(sc, -1) @ V
    return out


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()
