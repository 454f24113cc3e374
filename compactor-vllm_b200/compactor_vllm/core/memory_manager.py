"""KV cache manager: cache sizing, seq_id -> slot map, wiring into layers.

Reference: core/memory_manager.py:13-182.  `world_size` comes from the config
(the reference asks torch.distributed, which forces a process group to exist
even for one GPU); each tensor-parallel rank owns HKV/tp heads and its own
cache, so nothing here communicates.
"""
from __future__ import annotations

import logging
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist
from compactor_vllm.config.engine_config import LLMConfig
from compactor_vllm.kv_cache.page_table import DevicePagedKVCache, KVAllocationStatus, PagedKVCache
from torch import nn

logger = logging.getLogger(__name__)


def _world_size(config: LLMConfig) -> int:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size()
    return config.tensor_parallel_size


class KVCacheManager:
    def __init__(self, rank: int, config: LLMConfig):
        hf = config.hf_config
        self.rank = rank
        self.gpu_frac = config.gpu_memory_utilization
        self.page_size = config.kvcache_page_size
        self.world_size = config.tensor_parallel_size
        self.max_num_batches = config.max_num_seqs
        self.max_model_len = config.max_model_len
        self.num_layers = hf.num_hidden_layers
        self.model_dtype = hf.torch_dtype
        self.head_dim = getattr(hf, "head_dim", None)
        self.max_pages_per_batch = (self.max_model_len + self.page_size - 1) // self.page_size
        ws = _world_size(config)
        assert hf.num_key_value_heads % ws == 0, "world size needs to divide num_kv_heads"
        self.num_kv_heads = hf.num_key_value_heads // ws
        self.num_pages = None
        self.paged_cache: Optional[PagedKVCache] = None
        self.max_batched_tokens = None
        self.seq_id_to_batch = {}

    def allocate_sequences(self, seq_ids: List[int], max_positions: List[int]):
        """(ok, batch_mapping) — reference :40-60."""
        mapping = []
        for seq_id, want in zip(seq_ids, max_positions):
            if seq_id not in self.seq_id_to_batch:
                slot = self.paged_cache.new_batch()
                if slot is None:
                    logger.warning("Failed to allocate batch!")
                    return False, None
                self.seq_id_to_batch[seq_id] = int(slot)
            slot = self.seq_id_to_batch[seq_id]
            mapping.append(slot)
            status = self.paged_cache.reserve_tokens(slot, want)
            if status != KVAllocationStatus.SUCCESS:
                logger.warning(f"Failed to allocate pages ({status})!")
                return False, None
        dev = self.paged_cache.page_table.device
        return True, torch.as_tensor(mapping, dtype=torch.int32, device=dev)

    def free_sequences(self, seq_ids: Iterable[int]):
        for seq_id in seq_ids:
            slot = self.seq_id_to_batch.pop(seq_id, None)
            self.paged_cache.free_batch(slot)

    def init_cache(self, model: nn.Module, num_pages: Optional[int] = None, device_allocator: bool = False):
        """Size and allocate the cache, then wire it into every attention layer (:67-90).
        device_allocator=True: the GPU-resident page allocator (same surface, same page ids)."""
        self.num_pages = (
            int(num_pages) if num_pages is not None
            else self.get_num_pages(self.gpu_frac, self.max_pages_per_batch)
        )
        self.paged_cache = (DevicePagedKVCache if device_allocator else PagedKVCache)(
            num_layers=self.num_layers,
            H_kv=self.num_kv_heads,
            head_dim=self.head_dim,
            page_size=self.page_size,
            num_pages=int(self.num_pages),
            max_num_batches=self.max_num_batches,
            device=f"cuda:{self.rank}",
            dtype=self.model_dtype,
            max_logical_pages_per_head=int(self.max_pages_per_batch),
        )
        self._assign_cache_to_layers(model)

    def _assign_cache_to_layers(self, model) -> None:
        for layer_index, layer in enumerate(model.model.layers):
            attn = layer.self_attn.attn
            k, v, pt, bh = self.paged_cache.layer_slices(layer_index)
            attn.k_cache, attn.v_cache = k, v
            attn.page_table, attn.bh_seq_lens = pt, bh
            attn.page_size = self.page_size

    def page_table_bytes(self, n_logical_pages_max: int) -> int:
        per_layer = self.max_num_batches * self.num_kv_heads * (n_logical_pages_max + 1) * 4
        return self.num_layers * per_layer

    def get_num_pages(self, frac: float, n_logical_pages_max: int) -> int:
        """(total*frac*0.9 - used - peak + current - page_table_bytes) / bytes_per_page (:92-126)."""
        free, total = torch.cuda.mem_get_info()
        stats = torch.cuda.memory_stats()
        peak = int(stats["allocated_bytes.all.peak"])
        current = int(stats["allocated_bytes.all.current"])
        budget = int(total * frac * 0.9) - (total - free) - peak + current
        if budget <= 0:
            raise RuntimeError(
                "Insufficient memory for KV cache."
                f"Try increasing gpu_memory_utilization (currently {frac:.2f})."
            )
        net = budget - self.page_table_bytes(n_logical_pages_max)
        if net <= 0:
            raise RuntimeError(
                "page-table footprint exceeds KV cache budget. "
                f"reduce max_num_seqs ({self.max_num_batches}) "
                f"or increase kv_cache_mem_fraction (currently {frac:.2f})."
            )
        elem = torch.empty((), dtype=self.model_dtype).element_size()
        per_page = self.num_layers * 2 * self.page_size * self.head_dim * elem
        return max(1, net // per_page)

    def estimate_max_batched_tokens(
        self, warmup_tokens: int, bytes_used_before_warmup: int, bytes_peak_after_warmup: int
    ) -> int:
        """Activation-bound token budget (:128-161)."""
        assert warmup_tokens > 0, "warmup_tokens must be > 0"
        delta = max(0, int(bytes_peak_after_warmup) - int(bytes_used_before_warmup))
        per_token = max(1, (delta + warmup_tokens - 1) // warmup_tokens)
        free, total = torch.cuda.mem_get_info()
        stats = torch.cuda.memory_stats()
        cushion = max(
            0, int(stats.get("allocated_bytes.all.peak", 0)) - int(stats.get("allocated_bytes.all.current", 0))
        )
        act_budget = int(max(0, int(total * self.gpu_frac) - int(total - free) - cushion) * 0.95)
        by_act = (act_budget // per_token) // self.page_size * self.page_size
        by_cache = ((self.num_pages * self.page_size) // self.num_kv_heads) // self.page_size * self.page_size
        self.max_batched_tokens = min(by_cache, by_act)
        return self.max_batched_tokens

    @property
    def num_free_batches(self) -> int:
        return len(self.paged_cache.free_batches)

    @property
    def num_free_pages(self) -> int:
        if isinstance(self.paged_cache, DevicePagedKVCache):
            return int(self.paged_cache.free_count.min())
        return min(len(fp) for fp in self.paged_cache.free_pages)

    def reclaim_pages(self, seq_ids_to_reclaim: Iterable[int], future_reserved_buffer) -> int:
        freed = 0
        for i, seq_id in enumerate(seq_ids_to_reclaim):
            freed += self.paged_cache.reclaim_pages(self.seq_id_to_batch[seq_id], future_reserved_buffer[i])
        return freed
