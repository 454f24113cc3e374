"""Oracle: score ranking, top-k eviction and the compacting store."""
from __future__ import annotations

import numpy as np
import torch

from .store import head_rows


def sort_keys(scores: np.ndarray) -> np.ndarray:
    """Order-preserving uint32 key of fp32 scores; NaN ranks above +inf.

    torch.topk treats NaN as the largest value (compression/common.py:240-242
    relies on it for the 0/0 z-scores).  -0.0 orders just below +0.0.
    """
    bits = np.ascontiguousarray(scores, dtype=np.float32).view(np.uint32)
    key = np.where(bits & 0x80000000, ~bits, bits | np.uint32(0x80000000)).astype(np.uint32)
    key[np.isnan(scores)] = np.uint32(0xFFFFFFFF)
    return key


def rank_scores(scores: torch.Tensor, cu_seqlens_k, max_k_len: int, top_k: int, H: int) -> torch.Tensor:
    """compression/common.py:171-243: pad each sequence to max_k_len with -inf, full descending
    sort of the max_k_len*H entries, keep the first k_eff = min(top_k, max_k_len*H), add cu[b]*H.
    Tie rule (torch leaves it unspecified): ascending flat index."""
    sc = scores.detach().cpu().numpy().astype(np.float32)
    B = len(cu_seqlens_k) - 1
    total = max_k_len * H
    k_eff = min(int(top_k), total)
    out = np.zeros((B, k_eff), dtype=np.int64)
    for b in range(B):
        s, e = int(cu_seqlens_k[b]), int(cu_seqlens_k[b + 1])
        flat = np.full(total, -np.inf, dtype=np.float32)
        flat[: (e - s) * H] = sc[s:e].reshape(-1)
        order = np.argsort(~sort_keys(flat), kind="stable")  # ascending of inverted key
        out[b] = order[:k_eff] + s * H
    return torch.from_numpy(out)


def kept_positions_from_ranking(indices: np.ndarray, retain: int, H: int, cu_b: int, len_b: int,
                                L_prefix: np.ndarray, PS: int, pad: bool):
    """Per-head kept global token ids for one sequence, following the reference kernels.

    K4 (kv_cache/store_kv_cache.py:34-78): the first `retain` ranked entries go to their heads.
    K5 (:202-248): for each head whose length L (prefix + stored) is not a page multiple, walk the
    ranking from position `retain`, appending entries of that head until `need = PS - L%PS` are
    written, the ranking ends, or `ctx_len - L` are written.  Entries outside the sequence's
    own id range (padded slots) are skipped.
    """
    lo, hi = cu_b * H, (cu_b + len_b) * H
    max_sel = len(indices)
    kept = [[] for _ in range(H)]
    k_total = min(int(retain), max_sel)
    for i in range(k_total):
        sel = int(indices[i])
        if lo <= sel < hi:
            kept[sel % H].append(sel // H)
    if pad:
        for h in range(H):
            L = int(L_prefix[h]) + len(kept[h])
            mod = L % PS
            if mod == 0:
                continue
            need = PS - mod
            max_additional = len_b - L
            written, idx = 0, k_total
            while written < need and idx < max_sel and written < max_additional:
                cand = int(indices[idx])
                if lo <= cand < hi and cand % H == h:
                    kept[h].append(cand // H)
                    written += 1
                idx += 1
    return [sorted(x) for x in kept]


def store_topk(new_keys, new_vals, indices_topk, num_retain, cu_seqlens_k, page_table, batch_mapping,
               bh_lens, k_cache, v_cache, PS, pad=True, reserved_batch=0):
    """prefill_store_topk_kv (kv_cache/store_kv_cache.py:81-175) with rows written in ascending
    token order per head (the reference's atomicAdd slot order is nondeterministic, :54-57).
    Returns the kept token lists [b][h] (global token ids)."""
    B = indices_topk.shape[0]
    H = new_keys.shape[1]
    idx = indices_topk.cpu().numpy()
    kept_all = []
    for b in range(B):
        bt = int(batch_mapping[b])
        s, e = int(cu_seqlens_k[b]), int(cu_seqlens_k[b + 1])
        if bt == reserved_batch:
            kept_all.append([[] for _ in range(H)])
            continue
        Lp = bh_lens[b].cpu().numpy().astype(np.int64)
        kept = kept_positions_from_ranking(idx[b], int(num_retain[b]), H, s, e - s, Lp, PS, pad)
        for h in range(H):
            toks = kept[h]
            if not toks:
                continue
            rows = torch.from_numpy(head_rows(page_table, bt, h, int(Lp[h]), len(toks), PS))
            t = torch.tensor(toks, dtype=torch.long)
            k_cache[rows] = new_keys[t, h]
            v_cache[rows] = new_vals[t, h]
            bh_lens[b, h] = int(Lp[h]) + len(toks)
        kept_all.append(kept)
    return kept_all


def extract_and_store_top_kv(scores, cu_seqlens_k, max_k_len, top_k, H, new_keys, new_vals, num_retain,
                             page_table, batch_mapping, bh_lens, k_cache, v_cache, PS, pad=True,
                             reserved_batch=0):
    """compression/common.py:126-168 = rank_scores + store_topk."""
    idx = rank_scores(scores, cu_seqlens_k, max_k_len, top_k, H)
    return store_topk(new_keys, new_vals, idx, num_retain, cu_seqlens_k, page_table, batch_mapping,
                      bh_lens, k_cache, v_cache, PS, pad, reserved_batch)
