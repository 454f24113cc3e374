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


class DevicePagedKVCache(PagedKVCache):
    """PagedKVCache whose page allocator lives on the GPU (csrc/page_alloc.cu).

    Same constructor, methods, return values and — bit for bit — the same page ids as the host
    allocator above (and the reference's heaps): the free set of a layer is a bitmap, "pop the n
    smallest pages" a popcount scan.  reserve / reclaim / free are stream-ordered launches:

    * `reserve_tokens_device(slots, add_tokens, status=None)` / `reclaim_pages_device(slots, future)`
      take a DEVICE int32 vector of slots (entries equal to RESERVED_BATCH are skipped), never touch
      the host and can be captured in a CUDA graph next to the decode step — the reference pays two
      device->host syncs, a Python heap loop and a host->device copy per reservation
      (kv_cache/page_table.py:160-182) and cannot reclaim per layer under TP at all
      (core/model_runner.py:287-291);
    * the reference-shaped `reserve_tokens` / `reclaim_pages` / `new_batch` wrap them and read back
      4-8 bytes (the status / freed count / free counts) because their signatures return host values.

    Slots (`free_batches`) stay a host list: the engine hands them out on the host.
    """

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        from compactor_vllm import _native as N
        self._N = N
        dev = self.page_table.device
        assert dev.type == "cuda", "the device allocator has no CPU fallback"
        L = self.num_layers
        self._words = int(N.lib().cvllm_pages_bitmap_words(self.n_pages))
        self.free_bits = torch.empty((L, self._words), dtype=torch.int32, device=dev)
        self.free_count = torch.empty((L,), dtype=torch.int32, device=dev)
        self._freed = torch.zeros((1,), dtype=torch.int64, device=dev)
        self._status1 = torch.zeros((1,), dtype=torch.int32, device=dev)
        self._slot1 = torch.zeros((1,), dtype=torch.int32, device=dev)
        N.check(N.lib().cvllm_pages_init(N.ptr(self.free_bits), N.ptr(self.free_count), L, self.n_pages,
                                         N.stream(dev)), "cvllm_pages_init")
        # the host structures of the base class are not maintained
        self.pages_indices_per_batch = None

    # -- device-resident API (no host sync, graph-capturable) -------------------
    def reserve_tokens_device(self, slots: torch.Tensor, add_tokens: int,
                              status: Optional[torch.Tensor] = None) -> torch.Tensor:
        """slots int32 [nb] on the device -> status int32 [nb] (0 ok, 1 too long, 2 out of pages)."""
        N = self._N
        nb = slots.numel()
        assert slots.dtype == torch.int32 and slots.is_contiguous()
        if status is None:
            status = torch.empty((nb,), dtype=torch.int32, device=slots.device)
        L, Bs, H, P = self.page_table.shape
        lib = N.lib()
        ws_bytes = lib.cvllm_pages_workspace_bytes(L, H, P, self.n_pages, nb)
        ws = N.workspace(slots.device, ws_bytes, "pages")
        rc = lib.cvllm_pages_reserve(
            N.ptr(self.free_bits), N.ptr(self.free_count), N.ptr(self.page_table), N.ptr(self.bh_seq_lens),
            N.ptr(self.bh_num_pages), N.ptr(slots), nb, int(add_tokens), N.ptr(status), L, Bs, H, P,
            self.page_size, self.n_pages, RESERVED_BATCH, N.ptr(ws), ws.numel(), N.stream(slots.device))
        N.check(rc, "cvllm_pages_reserve")
        return status

    def _release(self, slots: torch.Tensor, future: int, free_all: bool, count: bool) -> None:
        N = self._N
        L, Bs, H, P = self.page_table.shape
        rc = N.lib().cvllm_pages_release(
            N.ptr(self.free_bits), N.ptr(self.free_count), N.ptr(self.page_table), N.ptr(self.bh_seq_lens),
            N.ptr(self.bh_num_pages), N.ptr(slots), slots.numel(), int(future), int(free_all),
            N.ptr(self._freed) if count else None, L, Bs, H, P, self.page_size, self.n_pages, RESERVED_BATCH,
            N.stream(slots.device))
        N.check(rc, "cvllm_pages_release")

    def reclaim_pages_device(self, slots: torch.Tensor, future_reserve_tokens: int = 0) -> None:
        assert slots.dtype == torch.int32 and slots.is_contiguous()
        self._release(slots, future_reserve_tokens, False, False)

    # -- reference-shaped API -----------------------------------------------------
    def new_batch(self) -> Optional[int]:
        if self.free_batches and int(self.free_count.min()) >= self.H_kv:
            return self.free_batches.pop()
        return None

    def reserve_tokens(self, batch_index: int, add_tokens: int) -> KVAllocationStatus:
        self._slot1.fill_(int(batch_index))
        code = int(self.reserve_tokens_device(self._slot1, add_tokens, self._status1))
        return (KVAllocationStatus.SUCCESS, KVAllocationStatus.EXCEEDS_MAX_SEQUENCE_LENGTH,
                KVAllocationStatus.EXCEEDS_CURRENTLY_AVAILABLE_PAGES)[code]

    def reclaim_pages(self, batch_index: int, future_reserve_tokens: int = 0) -> int:
        assert 0 <= batch_index < self.max_num_batches
        self._slot1.fill_(int(batch_index))
        self._freed.zero_()
        self._release(self._slot1, future_reserve_tokens, False, True)
        return int(self._freed) * self.page_size * self.head_dim * self.kv_cache.element_size() * 2

    def free_batch(self, batch_index: int) -> None:
        self._slot1.fill_(int(batch_index))
        self._release(self._slot1, 0, True, False)
        self.free_batches.append(batch_index)

    @property
    def free_pages(self) -> List[List[int]]:
        """Sorted free page ids per layer, decoded from the bitmap (tests / introspection: syncs)."""
        bits = self.free_bits.cpu().numpy().view("uint32")
        out = []
        for l in range(self.num_layers):
            ids = []
            for w, word in enumerate(bits[l].tolist()):
                while word:
                    low = word & -word
                    ids.append(w * 32 + low.bit_length() - 1)
                    word ^= low
            out.append(ids)
        return out

    @free_pages.setter
    def free_pages(self, value) -> None:  # the base constructor assigns its host heaps; dropped here
        pass
