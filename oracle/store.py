"""Oracle: paged-cache addressing and the three plain store ops."""
from __future__ import annotations

import numpy as np
import torch


def head_rows(page_table: torch.Tensor, b_true: int, h: int, start: int, count: int, PS: int) -> np.ndarray:
    """Physical cache rows of logical positions [start, start+count) of head (b_true, h).

    row = page_table[b_true, h, pos // PS] * PS + pos % PS  (reference:
    kv_cache/store_kv_cache.py:58-64, attention/sparse_decode_kernel.py:30-35).
    """
    pos = np.arange(start, start + count, dtype=np.int64)
    pt = page_table[b_true, h].cpu().numpy().astype(np.int64)
    return pt[pos // PS] * PS + pos % PS


def decode_store_kv(key, value, batch_mapping, bh_lens, page_table, k_cache, v_cache, PS, reserved_batch=0):
    """kv_cache/store_kv_cache.py:374-416: append one row per (b,h); len += 1; skip reserved rows."""
    B, H, _ = key.shape
    for b in range(B):
        bt = int(batch_mapping[b])
        if bt == reserved_batch:
            continue
        for h in range(H):
            L = int(bh_lens[b, h])
            row = int(head_rows(page_table, bt, h, L, 1, PS)[0])
            k_cache[row] = key[b, h]
            v_cache[row] = value[b, h]
            bh_lens[b, h] = L + 1


def prefill_store_all_kv(new_keys, new_values, cu_seqlens_k, k_cache, v_cache, page_table, bh_lens,
                         batch_mapping, PS):
    """kv_cache/store_kv_cache.py:251-371: row (t,h) of sequence b -> position bh_lens[b,h]+t;
    then bh_lens[b,:] += len_b (:371)."""
    B = batch_mapping.shape[0]
    H = new_keys.shape[1]
    cu = [int(x) for x in cu_seqlens_k]
    for b in range(B):
        s, e = cu[b], cu[b + 1]
        n = e - s
        if n <= 0:
            continue
        bt = int(batch_mapping[b])
        for h in range(H):
            rows = torch.from_numpy(head_rows(page_table, bt, h, int(bh_lens[b, h]), n, PS))
            k_cache[rows] = new_keys[s:e, h]
            v_cache[rows] = new_values[s:e, h]
    bh_lens += torch.tensor([cu[b + 1] - cu[b] for b in range(B)], dtype=bh_lens.dtype)[:, None]


def scatter_to_page_table(add_pages, new_phys_pages, curr_pages, page_table, max_pages_per_head):
    """kv_cache/write_page_table.py:6-95: append add[l,h] ids (row-major (l,h) order) at curr[l,h]."""
    L, H = add_pages.shape
    base = 0
    for l in range(L):
        for h in range(H):
            ap = int(add_pages[l, h])
            if ap <= 0:
                continue
            cp = int(curr_pages[l, h])
            for i in range(ap):
                if cp + i < max_pages_per_head:
                    page_table[l, h, cp + i] = int(new_phys_pages[base + i])
            base += ap
