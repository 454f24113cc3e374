"""KV cache manager: how many pages fit, which slot a sequence lives in, and the
wiring of the per-layer cache views into the attention modules.

The reference's class (core/memory_manager.py:13-182) mixes three things in its methods: reading the CUDA
allocator, two sizing formulas, and slot bookkeeping.  Here they are separate pieces:

* `MemorySnapshot`                 what the device allocator reports right now (the only CUDA reads);
* `kv_page_budget` / `activation_token_budget`   the two sizing formulas as pure integer functions of a
  snapshot — they have to give the reference's numbers (same cache size => same page ids => same page
  tables), and are pinned to the reference's own methods in tests/test_host_logic.py;
* `SlotMap`                        seq_id -> cache slot, handed out by the paged cache;
* `KVCacheManager`                 the reference's surface (attribute and method names, return values) on top.

Each tensor-parallel rank owns HKV / tp heads and its own cache, so nothing here communicates.
"""
from __future__ import annotations

import logging
from typing import Dict, Iterable, List, NamedTuple, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from compactor_vllm.config.engine_config import LLMConfig
from compactor_vllm.kv_cache.page_table import DevicePagedKVCache, KVAllocationStatus, PagedKVCache
from torch import nn

logger = logging.getLogger(__name__)

_INT32 = 4
_SAFETY = 0.9        # fraction of the user's memory fraction the cache may take (reference :97)
_ACT_SAFETY = 0.95   # same for the activation budget (reference :152)


class MemorySnapshot(NamedTuple):
    total: int      # device memory
    free: int       # cudaMemGetInfo free
    peak: int       # torch allocator, allocated_bytes.all.peak
    current: int    # torch allocator, allocated_bytes.all.current

    @classmethod
    def capture(cls, device=None) -> "MemorySnapshot":
        free, total = torch.cuda.mem_get_info(device)
        stats = torch.cuda.memory_stats(device)
        return cls(int(total), int(free), int(stats.get("allocated_bytes.all.peak", 0)),
                   int(stats.get("allocated_bytes.all.current", 0)))

    @property
    def used(self) -> int:
        return self.total - self.free

    @property
    def transient(self) -> int:
        """Bytes the forward pass needed on top of what is allocated now (peak - current)."""
        return self.peak - self.current


def table_bytes(num_layers: int, max_batches: int, heads: int, logical_pages: int) -> int:
    """page_table [L, B, H, P] + bh_seq_lens [L, B, H], int32 (reference :106-114)."""
    return num_layers * max_batches * heads * (logical_pages + 1) * _INT32


def kv_page_budget(snap: MemorySnapshot, frac: float, *, num_layers: int, heads: int, head_dim: int,
                   page_size: int, elem_size: int, max_batches: int, logical_pages: int) -> int:
    """Pages per layer that fit: (total * frac * 0.9 - used - transient - tables) // bytes of one page in
    every layer (K and V).  Reference :92-126; raises RuntimeError where it does."""
    budget = int(snap.total * frac * _SAFETY) - snap.used - snap.transient
    if budget <= 0:
        raise RuntimeError(f"Insufficient memory for KV cache: gpu_memory_utilization={frac:.2f} leaves "
                           f"{budget} bytes after {snap.used} used and {snap.transient} transient")
    net = budget - table_bytes(num_layers, max_batches, heads, logical_pages)
    if net <= 0:
        raise RuntimeError(f"page-table footprint exceeds KV cache budget ({budget} bytes): reduce max_num_seqs "
                           f"({max_batches}) or raise gpu_memory_utilization ({frac:.2f})")
    return max(1, net // (num_layers * 2 * page_size * head_dim * elem_size))


def activation_token_budget(snap: MemorySnapshot, frac: float, *, warmup_tokens: int, used_before: int,
                            peak_after: int, num_pages: int, page_size: int, heads: int) -> int:
    """Tokens one step may carry: bounded by activations (bytes per token measured on the warm-up pass against
    what is left of the memory fraction) and by what the cache can hold, both rounded down to whole pages
    (reference :128-161)."""
    assert warmup_tokens > 0, "warmup_tokens must be > 0"
    per_token = max(1, -(-max(0, int(peak_after) - int(used_before)) // warmup_tokens))
    room = max(0, int(snap.total * frac) - snap.used - max(0, snap.transient))
    whole_pages = lambda tokens: tokens // page_size * page_size  # noqa: E731
    return min(whole_pages(num_pages * page_size // heads), whole_pages(int(room * _ACT_SAFETY) // per_token))


class SlotMap:
    """seq_id -> cache slot.  Slots come from (and go back to) the paged cache's free list."""

    def __init__(self):
        self.by_seq: Dict[int, int] = {}

    def acquire(self, cache: PagedKVCache, seq_id: int) -> Optional[int]:
        slot = self.by_seq.get(seq_id)
        if slot is None:
            slot = cache.new_batch()
            if slot is None:
                return None
            self.by_seq[seq_id] = slot = int(slot)
        return slot

    def release(self, cache: PagedKVCache, seq_id: int) -> None:
        cache.free_batch(self.by_seq.pop(seq_id, None))


class KVCacheManager:
    def __init__(self, rank: int, config: LLMConfig):
        hf = config.hf_config
        tp = dist.get_world_size() if dist.is_available() and dist.is_initialized() else config.tensor_parallel_size
        assert hf.num_key_value_heads % tp == 0, "world size needs to divide num_kv_heads"
        self.rank, self.world_size = rank, config.tensor_parallel_size
        self.gpu_frac = config.gpu_memory_utilization
        self.page_size = config.kvcache_page_size
        self.max_num_batches, self.max_model_len = config.max_num_seqs, config.max_model_len
        self.num_layers, self.model_dtype = hf.num_hidden_layers, hf.torch_dtype
        self.head_dim = getattr(hf, "head_dim", None)
        self.num_kv_heads = hf.num_key_value_heads // tp          # this rank's shard (reference :29-32)
        self.max_pages_per_batch = -(-self.max_model_len // self.page_size)
        self.num_pages: Optional[int] = None
        self.paged_cache: Optional[PagedKVCache] = None
        self.max_batched_tokens: Optional[int] = None
        self._slots = SlotMap()

    # the reference exposes the dict itself (model_runner reads it)
    @property
    def seq_id_to_batch(self) -> Dict[int, int]:
        return self._slots.by_seq

    # ---- sizing ---------------------------------------------------------------------------------
    def _shape(self) -> dict:
        return dict(num_layers=self.num_layers, heads=self.num_kv_heads, head_dim=self.head_dim,
                    page_size=self.page_size, elem_size=torch.empty((), dtype=self.model_dtype).element_size(),
                    max_batches=self.max_num_batches)

    def page_table_bytes(self, n_logical_pages_max: int) -> int:
        return table_bytes(self.num_layers, self.max_num_batches, self.num_kv_heads, n_logical_pages_max)

    def get_num_pages(self, frac: float, n_logical_pages_max: int) -> int:
        return kv_page_budget(MemorySnapshot.capture(), frac, logical_pages=n_logical_pages_max, **self._shape())

    def estimate_max_batched_tokens(self, warmup_tokens: int, bytes_used_before_warmup: int,
                                    bytes_peak_after_warmup: int) -> int:
        self.max_batched_tokens = activation_token_budget(
            MemorySnapshot.capture(), self.gpu_frac, warmup_tokens=warmup_tokens,
            used_before=bytes_used_before_warmup, peak_after=bytes_peak_after_warmup,
            num_pages=self.num_pages, page_size=self.page_size, heads=self.num_kv_heads)
        return self.max_batched_tokens

    # ---- the cache ------------------------------------------------------------------------------
    def init_cache(self, model: nn.Module, num_pages: Optional[int] = None, device_allocator: bool = False):
        """Size the cache (unless `num_pages` is given), allocate it and hand every attention layer its views.
        device_allocator=True: the GPU-resident page allocator (same surface, same page ids)."""
        if num_pages is None:
            num_pages = self.get_num_pages(self.gpu_frac, self.max_pages_per_batch)
        self.num_pages = int(num_pages)
        cache_type = DevicePagedKVCache if device_allocator else PagedKVCache
        self.paged_cache = cache_type(
            num_layers=self.num_layers, max_logical_pages_per_head=int(self.max_pages_per_batch),
            num_pages=self.num_pages, page_size=self.page_size, H_kv=self.num_kv_heads, head_dim=self.head_dim,
            max_num_batches=self.max_num_batches, dtype=self.model_dtype, device=f"cuda:{self.rank}")
        for index, layer in enumerate(model.model.layers):
            attn = layer.self_attn.attn
            attn.k_cache, attn.v_cache, attn.page_table, attn.bh_seq_lens = self.paged_cache.layer_slices(index)
            attn.page_size = self.page_size

    @property
    def num_free_batches(self) -> int:
        return len(self.paged_cache.free_batches)

    @property
    def num_free_pages(self) -> int:
        """Free pages of the fullest layer (layers fill unevenly after compression)."""
        cache = self.paged_cache
        if isinstance(cache, DevicePagedKVCache):
            return int(cache.free_count.min())
        return min(map(len, cache.free_pages))

    # ---- sequences ------------------------------------------------------------------------------
    def allocate_sequences(self, seq_ids: Sequence[int], max_positions: Sequence[int]
                           ) -> Tuple[bool, Optional[torch.Tensor]]:
        """Give every sequence a slot (a new one the first time it is seen) and room for `max_positions[i]`
        more tokens.  Returns (True, int32 slot vector on the cache's device), or (False, None) as soon as the
        cache runs out of slots or pages; like the reference (:40-60) what was granted before the failure stays
        with its sequence — the scheduler retries or frees."""
        slots: List[int] = []
        for seq_id, tokens in zip(seq_ids, max_positions):
            slot = self._slots.acquire(self.paged_cache, seq_id)
            if slot is None:
                logger.warning("no free cache slot for sequence %s", seq_id)
                return False, None
            status = self.paged_cache.reserve_tokens(slot, tokens)
            if status != KVAllocationStatus.SUCCESS:
                logger.warning("no room for %d tokens of sequence %s: %s", tokens, seq_id, status)
                return False, None
            slots.append(slot)
        return True, torch.tensor(slots, dtype=torch.int32, device=self.paged_cache.page_table.device)

    def free_sequences(self, seq_ids: Iterable[int]) -> None:
        for seq_id in seq_ids:
            self._slots.release(self.paged_cache, seq_id)

    def reclaim_pages(self, seq_ids_to_reclaim: Iterable[int], future_reserved_buffer) -> int:
        """Return the pages each sequence holds beyond len + future_reserved_buffer[i] tokens; approximate bytes."""
        return sum(self.paged_cache.reclaim_pages(self._slots.by_seq[seq_id], future_reserved_buffer[i])
                   for i, seq_id in enumerate(seq_ids_to_reclaim))
