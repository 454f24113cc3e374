"""Paged KV cache state + host page allocator.

Same public surface and allocation semantics as the reference's
kv_cache/page_table.py:28-313: one flat K and one flat V buffer per layer
(`kv_cache[2, L, n_pages*PS, D]`), a page table `[L, Bmax+1, HKV, P]`, per-
(layer, slot, head) lengths and page counts, per-layer min-heaps of free
physical pages (so pops hand out ascending page ids), slot 0 reserved.

Layout note for B200: a page is PS*D*2 bytes contiguous (32 KB at PS=128,
D=128, bf16), i.e. exactly one TMA box row-block for the prefill kernel and 256
full 128-byte lines for the decode kernel's vector loads.
"""
from __future__ import annotations

import heapq
import logging
from enum import Enum, auto
from typing import List, Optional, Union

import torch
from compactor_vllm.config.constants import RESERVED_BATCH
from compactor_vllm.kv_cache import write_page_table as _wpt

logger = logging.getLogger(__name__)


def cdiv(a, b):
    return (a + b - 1) // b


def next_multiple(a, b):
    return cdiv(a, b) * b


class KVAllocationStatus(Enum):
    EXCEEDS_MAX_SEQUENCE_LENGTH = auto()
    EXCEEDS_CURRENTLY_AVAILABLE_PAGES = auto()
    EXCEEDS_MAX_NUM_BATCHES = auto()
    SUCCESS = auto()


class PagedKVCache(torch.nn.Module):
    def __init__(
        self,
        num_layers: int,
        max_logical_pages_per_head: int,
        num_pages: int,
        page_size: int,
        H_kv: int,
        head_dim: int,
        max_num_batches: int,
        dtype: torch.dtype,
        device: Union[str, torch.device, int] = "cuda",
    ):
        super().__init__()
        self.n_pages = num_pages
        self.num_layers = num_layers
        self.page_size: int = int(page_size)
        self.H_kv = int(H_kv)
        self.max_pages_per_head = max_logical_pages_per_head
        self.max_num_batches = max_num_batches + 1  # + the reserved slot
        self.head_dim = head_dim
        self.kv_cache = torch.empty(
            (2, num_layers, num_pages * self.page_size, head_dim), dtype=dtype, device=device
        )
        self.page_table = torch.empty(
            (num_layers, self.max_num_batches, self.H_kv, self.max_pages_per_head),
            device=device, dtype=torch.int32,
        )
        self.bh_seq_lens = torch.zeros(
            (num_layers, self.max_num_batches, self.H_kv), device=device, dtype=torch.int32
        )
        self.bh_num_pages = torch.zeros_like(self.bh_seq_lens)

        # per-layer min-heap of free physical pages; list(range(n)) is already a heap
        self.free_pages: List[List[int]] = [list(range(num_pages)) for _ in range(num_layers)]
        # pop() hands out slot 1 first; slot RESERVED_BATCH never enters the list
        self.free_batches: List[int] = [
            s for s in range(self.max_num_batches - 1, -1, -1) if s != RESERVED_BATCH
        ]
        self.pages_indices_per_batch: List[List[set]] = [
            [set() for _ in range(num_layers)] for _ in range(self.max_num_batches)
        ]

    # -- slots ----------------------------------------------------------------
    def new_batch(self) -> Optional[int]:
        """Reserve a slot if one is free and every layer still has >= H_kv free pages (:125-142)."""
        if self.free_batches and all(len(fp) >= self.H_kv for fp in self.free_pages):
            return self.free_batches.pop()
        return None

    # -- pages ----------------------------------------------------------------
    def reserve_tokens(self, batch_index: int, add_tokens: int) -> KVAllocationStatus:
        """Make every (layer, head) of the slot able to hold len + add_tokens (:144-198)."""
        lens = self.bh_seq_lens[:, batch_index]  # [L, H]
        pages = self.bh_num_pages[:, batch_index]  # [L, H]
        need = lens + add_tokens
        cap = pages * self.page_size
        if bool((need <= cap).all()):
            return KVAllocationStatus.SUCCESS
        # clamp at 0: heads with a spare page would otherwise get a negative count (the
        # reference does not clamp, :164-166; harmless there because the engine only ever
        # reserves for fresh slots where all heads are equal) -- deliberate fix, see DESIGN.md
        add_pages = cdiv(need - cap, self.page_size).clamp_min(0)
        new_total = pages + add_pages
        if bool((new_total > self.max_pages_per_head).any()):
            return KVAllocationStatus.EXCEEDS_MAX_SEQUENCE_LENGTH
        per_layer = add_pages.sum(dim=-1).tolist()
        if any(n > len(fp) for n, fp in zip(per_layer, self.free_pages)):
            return KVAllocationStatus.EXCEEDS_CURRENTLY_AVAILABLE_PAGES
        flat: List[int] = []
        for layer, n in enumerate(per_layer):
            heap = self.free_pages[layer]
            got = [heapq.heappop(heap) for _ in range(n)]
            self.pages_indices_per_batch[batch_index][layer].update(got)
            flat.extend(got)
        new_phys = torch.tensor(flat, dtype=torch.int32, device=self.page_table.device)
        _wpt.scatter_to_page_table(
            add_pages=add_pages,
            new_phys_pages=new_phys,
            curr_pages=pages,
            page_table=self.page_table[:, batch_index],
            max_pages_per_head=self.max_pages_per_head,
        )
        self.bh_num_pages[:, batch_index, :] = new_total.to(self.bh_num_pages.dtype)
        return KVAllocationStatus.SUCCESS

    def reclaim_pages(self, batch_index: int, future_reserve_tokens: int = 0) -> int:
        """Free pages beyond ceil((len + future)/PS) for the slot (:200-267)."""
        L, B, H = self.bh_seq_lens.shape
        assert 0 <= batch_index < B
        P = self.max_pages_per_head
        alloc = self.bh_num_pages[:, batch_index, :]
        used = torch.minimum(
            cdiv(self.bh_seq_lens[:, batch_index, :] + future_reserve_tokens, self.page_size), alloc
        )
        p = torch.arange(P, device=alloc.device, dtype=torch.int32).view(1, 1, P)
        free_mask = (p < alloc.unsqueeze(-1)) & (p >= used.unsqueeze(-1))  # [L, H, P]
        if not bool(free_mask.any()):
            return 0
        slots = free_mask.reshape(-1).nonzero(as_tuple=False).squeeze(-1)
        freed = self.page_table[:, batch_index].reshape(-1)[slots].tolist()
        layers = (slots // (H * P)).tolist()
        self.bh_num_pages[:, batch_index, :] = used
        for page, layer in zip(freed, layers):
            self.pages_indices_per_batch[batch_index][layer].remove(page)
            heapq.heappush(self.free_pages[layer], page)
        return len(freed) * self.page_size * self.head_dim * self.kv_cache.element_size() * 2

    def _free_batch_layer(self, layer_index: int, batch_index: int) -> None:
        owned = self.pages_indices_per_batch[batch_index][layer_index]
        for phys in owned:
            heapq.heappush(self.free_pages[layer_index], int(phys))
        self.pages_indices_per_batch[batch_index][layer_index] = set()

    def free_batch(self, batch_index: int) -> None:
        """Return every page of the slot and the slot itself (:279-291)."""
        for layer in range(self.num_layers):
            self._free_batch_layer(layer, batch_index)
        self.bh_seq_lens[:, batch_index].zero_()
        self.bh_num_pages[:, batch_index].zero_()
        self.free_batches.append(batch_index)

    def layer_slices(self, layer: int):
        """(k_cache, v_cache, page_table, bh_seq_lens) views of one layer (:293-313)."""
        assert 0 <= layer < self.num_layers
        return (
            self.kv_cache[0, layer], self.kv_cache[1, layer],
            self.page_table[layer], self.bh_seq_lens[layer],
        )
