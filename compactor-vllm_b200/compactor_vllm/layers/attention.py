"""Attention module — the drop-in boundary (reference: layers/attention.py:17-161).

Prefill: store of the new K/V (all rows, or only the top-scoring (token, head)
pairs) on STORE_STREAM concurrently with causal attention over
[cached prefix || new K/V] on the main stream.  Decode: append one row per
(b, h), then split-KV paged attention.  The KV-cache manager injects k_cache,
v_cache, page_table, bh_seq_lens and page_size (core/memory_manager.py).
"""
from typing import Optional

import torch
from compactor_vllm.attention.sparse_decode_kernel import decode_append_and_attend, head_sparse_decode_attention
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from compactor_vllm.compression.common import extract_and_store_top_kv
from compactor_vllm.config.constants import RESERVED_BATCH
from compactor_vllm.config.engine_config import AttentionBackend
from compactor_vllm.kv_cache.store_kv_cache import decode_store_kv, prefill_store_all_kv
from compactor_vllm.utils.context import Context, get_context
from compactor_vllm.utils.helpers import maybe_execute_in_stream
from torch import nn


class Attention(nn.Module):
    def __init__(self, num_heads, head_dim, scale, num_kv_heads):
        super().__init__()
        self.num_heads: int = num_heads
        self.head_dim = head_dim
        self.scale: float = scale
        self.num_kv_heads = int(num_kv_heads)
        self.k_cache: Optional[torch.Tensor] = None
        self.v_cache: Optional[torch.Tensor] = None
        self.page_table: Optional[torch.Tensor] = None
        self.bh_seq_lens: Optional[torch.Tensor] = None
        self.page_size: Optional[int] = None

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scores: Optional[torch.Tensor] = None):
        context: Context = get_context()
        batch_mapping = context.batch_mapping
        if not context.is_prefill and self.bh_seq_lens is not None and self.bh_seq_lens.is_contiguous():
            # decode: append + attend index bh_seq_lens through batch_mapping themselves — the
            # reference's index_select / index_copy_ round trip (:46-50,152-160) is not needed
            assert self.k_cache is not None, "KV Cache must be initialized for decoding"
            return decode_append_and_attend(
                q, k, v, self.k_cache, self.v_cache, self.bh_seq_lens, self.page_table, batch_mapping,
                int(self.num_kv_heads), self.page_size, self.scale, reserved_batch=RESERVED_BATCH,
            )
        seq_lens = (
            None if self.bh_seq_lens is None
            else self.bh_seq_lens.index_select(0, batch_mapping).contiguous()
        )
        if context.is_prefill:
            # attention must see the PRE-store lengths (reference :52,115)
            seq_lens_copy = seq_lens.clone() if seq_lens is not None else None
            if self.k_cache is not None and context.do_compression and scores is not None:
                cc = context.compression_context
                assert cc is not None
                maybe_execute_in_stream(
                    extract_and_store_top_kv,
                    scores=scores, cu_seqlens_k=context.cu_seqlens_k, max_k_len=context.max_seqlen_k,
                    top_k=cc.max_tokens_to_retain, H=int(self.num_kv_heads), new_keys=k, new_vals=v,
                    num_tokens_to_retain=cc.batch_tokens_to_retain, page_table=self.page_table,
                    batch_mapping=batch_mapping, bh_lens=seq_lens, k_cache=self.k_cache,
                    v_cache=self.v_cache, PAGE_SIZE=self.page_size, PAD_TO_PAGE_SIZE=True,
                    STORE_STREAM=context.STORE_STREAM,
                )
            elif self.k_cache is not None:
                maybe_execute_in_stream(
                    prefill_store_all_kv,
                    new_keys=k, new_values=v, cu_seqlens_k=context.cu_seqlens_k,
                    max_seqlen_k=context.max_seqlen_k, k_cache=self.k_cache, v_cache=self.v_cache,
                    page_table=self.page_table, bh_lens=seq_lens, batch_mapping=batch_mapping,
                    PAGE_SIZE=self.page_size, STORE_STREAM=context.STORE_STREAM,
                )
            if context.attention_backend == AttentionBackend.FLASH_ATTENTION:
                # The reference dispatches to flash-attn here (:96-107).  This build has one
                # backend by design (no multi-backend dispatch); fail loudly rather than fall back.
                raise NotImplementedError(
                    "AttentionBackend.FLASH_ATTENTION is not provided; use COMPACTOR_TRITON "
                    "(served by the sm_100a CUDA kernels)"
                )
            elif context.attention_backend == AttentionBackend.COMPACTOR_TRITON:
                o = causal_sparse_varlen_with_cache(
                    q, k, v, self.k_cache, self.v_cache,
                    seq_lens_bh=seq_lens_copy, global_page_table=self.page_table,
                    batch_mapping=batch_mapping, cu_seqlens_q=context.cu_seqlens_q,
                    max_seqlen_q=context.max_seqlen_q, max_seqlen_k_cache=context.max_bh_len,
                    HKV=int(self.num_kv_heads), PAGE_SIZE=self.page_size, sm_scale=self.scale,
                )
            else:
                raise NotImplementedError
        else:
            assert self.k_cache is not None, "KV Cache must be initialized for decoding"
            decode_store_kv(
                key=k, value=v, batch_mapping=batch_mapping, bh_lens=seq_lens,
                page_table=self.page_table, k_cache=self.k_cache, v_cache=self.v_cache,
                PAGE_SIZE=self.page_size,
            )
            o = head_sparse_decode_attention(
                q, self.k_cache, self.v_cache, seq_lens, self.page_table, batch_mapping,
                int(self.num_kv_heads), self.page_size, self.scale, key_split=context.key_split,
            )
        if self.bh_seq_lens is not None:
            maybe_execute_in_stream(
                self.bh_seq_lens.index_copy_, 0, batch_mapping.to(torch.long), seq_lens,
                STORE_STREAM=context.STORE_STREAM if context.is_prefill else None,
            )
        return o
