"""The two formulas of utils/arguments.py the hot path depends on.

The reference's PackedTensorArguments (utils/arguments.py:43-386) packs batch
metadata and broadcasts it over NCCL — engine plumbing, out of scope.  What the
path needs from it is the sketch matrix PHI (:81-86) and the per-sequence
retain budget (:105-121,208-211); synthetic runs use these so they see the
engine's numbers.
"""
import math
from typing import List, Sequence

import torch


def make_phi(head_dim: int, sketch_dim: int, dtype: torch.dtype, device, seed: int = 42) -> torch.Tensor:
    """PHI = randn(D, k, generator(seed) on `device`).to(dtype) / sqrt(k)  (:81-86)."""
    gen = torch.Generator(device=device).manual_seed(seed)
    phi = torch.randn((head_dim, sketch_dim), device=device, generator=gen)
    return phi.to(dtype) * (1 / math.sqrt(sketch_dim))


def tokens_to_retain(
    prompt_lens: Sequence[int], ratios: Sequence[float], protected_first: Sequence[int],
    protected_last: Sequence[int], num_kv_heads_local: int,
) -> List[int]:
    """retain_b = max(round(ratio*(L - first - last)*HKV_local), 1)  (:109-121)."""
    return [
        max(int(round(r * (L - s - e) * num_kv_heads_local)), 1)
        for L, r, s, e in zip(prompt_lens, ratios, protected_first, protected_last)
    ]


def max_tokens_to_retain(prompt_lens: Sequence[int], num_kv_heads_local: int) -> int:
    """top_k handed to the ranking = max(context_lens)*HKV_local  (:208-211)."""
    return max(prompt_lens) * num_kv_heads_local


def kv_head_shard(rank: int, world_size: int, num_heads: int, num_kv_heads: int):
    """Heads owned by tensor-parallel rank `rank`: KV head h and its G query heads live on rank
    h // (HKV / tp)  (reference: models/llama3.py:41-50, core/memory_manager.py:29-32).
    Returns (kv_start, kv_end, q_start, q_end), end-exclusive."""
    assert num_kv_heads % world_size == 0, "world size needs to divide num_kv_heads"
    assert num_heads % num_kv_heads == 0
    per = num_kv_heads // world_size
    group = num_heads // num_kv_heads
    return rank * per, (rank + 1) * per, rank * per * group, (rank + 1) * per * group
