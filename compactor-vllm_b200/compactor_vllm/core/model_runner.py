"""Single-rank model runner of the engine shell (reference: core/model_runner.py:32-555).

What it keeps of the reference: prefill batches packed by a token budget (core/scheduler.py:65-108),
per-sequence retain budgets and PHI (utils/arguments.py:81-121), the global `Context` set before each
forward (:185-207,437-442), scoring + store on a side stream joined before decoding (:322-323),
page reclaim after a compressed prefill (:288-291), decode as CUDA-graph replays per padded batch
size with padding rows mapped to the reserved slot (:469-545), Gumbel-max sampling.

What it leaves out (control plane, out of scope): tokenizer / chat templates, HF weight loading,
multi-process tensor parallelism (`tensor_parallel_size` must be 1), preemption.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from compactor_vllm.api_types import BatchCompressionParams, CompressionMethod
from compactor_vllm.config.constants import RESERVED_BATCH
from compactor_vllm.config.engine_config import LLMConfig
from compactor_vllm.core.memory_manager import KVCacheManager
from compactor_vllm.layers.sampler import Sampler
from compactor_vllm.models import MODEL_REGISTRY
from compactor_vllm.utils import arguments as A
from compactor_vllm.utils.context import CompressionContext, reset_context, set_context
from compactor_vllm.utils.sequence import Sequence, SequenceStatus


def _pad_batch(n: int) -> int:
    """Captured batch sizes: powers of two (core/model_runner.py:99-107)."""
    b = 1
    while b < n:
        b *= 2
    return b


def _check_shape_envelope(hf, config: LLMConfig) -> None:
    """Fail at construction, with the reason, for geometries the sm_100a kernels do not serve (INTEGRATION.md
    section 3: the per-call answer would be CVLLM_E_UNSUPPORTED in the middle of the first prefill)."""
    head_dim = getattr(hf, "head_dim", None) or hf.hidden_size // hf.num_attention_heads
    group = hf.num_attention_heads // hf.num_key_value_heads
    problems = []
    if head_dim != 128:
        problems.append(f"head_dim {head_dim} (only 128)")
    if group > 8:
        problems.append(f"GQA group {group} (decode serves <= 8)")
    if hf.num_key_value_heads > 16:
        problems.append(f"{hf.num_key_value_heads} KV heads per rank (selection serves <= 16)")
    if config.kvcache_page_size % 128 != 0:
        problems.append(f"kvcache_page_size {config.kvcache_page_size} (prefill needs a multiple of 128)")
    if config.leverage_sketch_size != 48:
        problems.append(f"leverage_sketch_size {config.leverage_sketch_size} (only 48)")
    if problems:
        raise ValueError("model / config outside the kernels' shape envelope: " + "; ".join(problems))


class KVCacheExhausted(RuntimeError):
    """The paged cache cannot hold the tokens a decode wave may still produce."""


class ModelRunner:
    def __init__(self, config: LLMConfig, *, num_pages: Optional[int] = None, seed: int = 0,
                 device_allocator: bool = True, max_batched_tokens: Optional[int] = None):
        assert config.tensor_parallel_size == 1, "the engine shell is single-rank (TP is SURVEY 8(f)-4)"
        assert torch.cuda.is_available(), "no CPU fallback"
        self.config = config
        hf = config.hf_config
        _check_shape_envelope(hf, config)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.dtype = hf.torch_dtype if isinstance(hf.torch_dtype, torch.dtype) else torch.bfloat16
        kind = "qwen3" if "qwen3" in str(getattr(hf, "model_type", "llama")).lower() else "llama"
        prev = torch.get_default_dtype()
        torch.set_default_dtype(self.dtype)  # parameters are born in the model dtype on the GPU
        try:
            with torch.device(self.device):
                model = MODEL_REGISTRY[kind](hf)
        finally:
            torch.set_default_dtype(prev)
        model.init_synthetic_weights(seed)
        self.model = model.to(self.device).eval()
        self.num_kv_heads = hf.num_key_value_heads
        self.head_dim = getattr(hf, "head_dim", None) or hf.hidden_size // hf.num_attention_heads
        self.kv = KVCacheManager(self.device.index, config)
        self.kv.head_dim = self.head_dim
        self.kv.model_dtype = self.dtype
        self.kv.init_cache(self.model, num_pages=num_pages, device_allocator=device_allocator)
        self.store_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.PHI = A.make_phi(self.head_dim, config.leverage_sketch_size, self.dtype, self.device)
        self.sampler = Sampler()
        self.max_batched_tokens = int(max_batched_tokens or config.max_model_len)
        self.eos = config.eos
        self._graphs: dict = {}
        self.stats = {"prefill_tokens": 0, "decode_steps": 0, "graph_replays": 0, "bytes_reclaimed": 0}

    # ------------------------------------------------------------------ prefill
    def _prefill_batches(self, seqs: List[Sequence]) -> List[List[Sequence]]:
        batches, cur, tokens = [], [], 0
        for s in seqs:
            n = s.num_prompt_tokens
            assert 0 < n <= self.config.max_model_len, "prompt longer than max_model_len"
            if cur and (tokens + n > self.max_batched_tokens or len(cur) >= self.config.max_num_seqs):
                batches.append(cur)
                cur, tokens = [], 0
            cur.append(s)
            tokens += n
        if cur:
            batches.append(cur)
        return batches

    @torch.no_grad()
    def run_prefill(self, seqs: List[Sequence], params: BatchCompressionParams) -> torch.Tensor:
        dev = self.device
        lens = [s.num_prompt_tokens for s in seqs]
        ok, batch_mapping = self.kv.allocate_sequences([s.seq_id for s in seqs], lens)
        assert ok, "KV cache exhausted (no preemption in the engine shell)"
        ids = torch.tensor([t for s in seqs for t in s.prompt_token_ids], dtype=torch.int64, device=dev)
        positions = torch.cat([torch.arange(n, device=dev) for n in lens])
        cu = torch.zeros(len(seqs) + 1, dtype=torch.int32, device=dev)
        cu[1:] = torch.tensor(lens, dtype=torch.int32, device=dev).cumsum(0)
        ratios = [s.compression_params.compression_ratio for s in seqs]
        do_compression = (params.compression_method is not CompressionMethod.NONE
                          and any(r < 1.0 for r in ratios))
        cc = None
        if do_compression:
            first = [s.compression_params.protected_first_tokens for s in seqs]
            last = [s.compression_params.protected_last_tokens for s in seqs]
            retain = A.tokens_to_retain(lens, ratios, first, last, self.num_kv_heads)
            cc = CompressionContext(
                compression_method=params.compression_method,
                compression_chunk_size=params.chunk_size if params.do_chunked_compression else -1,
                batch_tokens_to_retain=torch.tensor(retain, dtype=torch.int32, device=dev),
                max_tokens_to_retain=A.max_tokens_to_retain(lens, self.num_kv_heads),
                context_lens=lens, PHI=self.PHI, protected_first_tokens=first, protected_last_tokens=last)
        set_context(is_prefill=True, do_compression=do_compression, cu_seqlens_q=cu, cu_seqlens_k=cu,
                    max_seqlen_q=max(lens), max_seqlen_k=max(lens), batch_mapping=batch_mapping,
                    max_bh_len=0, compression_context=cc, STORE_STREAM=self.store_stream)
        hidden = self.model(ids, positions)
        logits = self.model.compute_logits(hidden)
        reset_context()
        # scoring / selection / compaction ran beside attention; join before anything reads the cache
        torch.cuda.current_stream().wait_stream(self.store_stream)
        self.stats["prefill_tokens"] += sum(lens)
        if do_compression:  # give back the pages the evicted pairs no longer need (:288-291)
            self.stats["bytes_reclaimed"] += self.kv.reclaim_pages(
                [s.seq_id for s in seqs], [0] * len(seqs))
        return logits

    # ------------------------------------------------------------------- decode
    @torch.no_grad()
    def _decode_forward(self, ids, positions, batch_mapping, temps):
        set_context(is_prefill=False, batch_mapping=batch_mapping)
        hidden = self.model(ids, positions)
        logits = self.model.compute_logits(hidden)
        reset_context()
        return self.sampler(logits, temps)

    def _graph_for(self, bs: int):
        g = self._graphs.get(bs)
        if g is not None:
            return g
        dev = self.device
        st = {"ids": torch.zeros(bs, dtype=torch.int64, device=dev),
              "pos": torch.zeros(bs, dtype=torch.int64, device=dev),
              "bm": torch.full((bs,), RESERVED_BATCH, dtype=torch.int32, device=dev),
              "temps": torch.zeros(bs, dtype=torch.float32, device=dev)}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            # warm-up on the capture stream with every row on the reserved slot: allocates the per-stream
            # workspaces and touches no sequence's cache
            self._decode_forward(st["ids"], st["pos"], st["bm"], st["temps"])
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                st["out"] = self._decode_forward(st["ids"], st["pos"], st["bm"], st["temps"])
        torch.cuda.current_stream().wait_stream(side)
        self._graphs[bs] = (graph, st)
        return self._graphs[bs]

    @torch.no_grad()
    def run_decode_loop(self, seqs: List[Sequence]) -> None:
        dev = self.device
        live = [s for s in seqs if s.status is SequenceStatus.RUNNING]
        if not live:
            return
        # capacity for every token this wave can still produce, reserved once (the reference reserves per
        # scheduler step on the host; `DevicePagedKVCache.reserve_tokens_device` is the in-graph alternative)
        for s in live:
            left = s.sampling_params.max_new_tokens - s.num_generated_tokens
            # never ask for more than the model length allows (generation stops there, see _accept): the tail
            # would be refused as "too long" although those tokens are never produced
            room = self.config.max_model_len - (s.num_prompt_tokens + s.num_generated_tokens)
            want = max(min(left, room), 0) + 1
            ok, _ = self.kv.allocate_sequences([s.seq_id], [want])
            if not ok:
                raise KVCacheExhausted(f"sequence {s.seq_id}: no room for {want} more tokens "
                                       f"({self.kv.num_free_pages} free pages; the engine shell has no preemption)")
        bs = _pad_batch(len(live))
        slots = [self.kv.seq_id_to_batch[s.seq_id] for s in live]
        bm = torch.full((bs,), RESERVED_BATCH, dtype=torch.int32, device=dev)
        bm[:len(live)] = torch.tensor(slots, dtype=torch.int32, device=dev)
        temps = torch.zeros(bs, dtype=torch.float32, device=dev)
        temps[:len(live)] = torch.tensor([s.sampling_params.temperature for s in live], device=dev)
        ids = torch.zeros(bs, dtype=torch.int64, device=dev)
        pos = torch.zeros(bs, dtype=torch.int64, device=dev)
        use_graph = not self.config.enforce_eager
        if use_graph:
            graph, st = self._graph_for(bs)
        while any(s.status is SequenceStatus.RUNNING for s in live):
            for i, s in enumerate(live):
                if s.status is SequenceStatus.RUNNING:
                    ids[i] = s.completion_token_ids[-1]
                    pos[i] = s.num_prompt_tokens + s.num_generated_tokens - 1
                else:
                    bm[i] = RESERVED_BATCH  # finished rows are skipped by the store and the attention
            if use_graph:
                st["ids"].copy_(ids); st["pos"].copy_(pos); st["bm"].copy_(bm); st["temps"].copy_(temps)
                graph.replay()
                out = st["out"]
                self.stats["graph_replays"] += 1
            else:
                out = self._decode_forward(ids, pos, bm, temps)
            self.stats["decode_steps"] += 1
            toks = out[:len(live)].tolist()
            for s, t in zip(live, toks):
                if s.status is SequenceStatus.RUNNING:
                    self._accept(s, t)

    def _accept(self, s: Sequence, token: int) -> None:
        s.add_new_token(token)
        full = s.num_prompt_tokens + s.num_generated_tokens >= self.config.max_model_len  # RoPE table / reservation
        if s.num_generated_tokens >= s.sampling_params.max_new_tokens or token == self.eos or full:
            s.status = SequenceStatus.FINISHED

    # ----------------------------------------------------------------- generate
    @torch.no_grad()
    def generate(self, seqs: List[Sequence], params: BatchCompressionParams) -> None:
        """Fill `completion_token_ids` of every sequence (reference: core/model_runner.py:230-338)."""
        wave = self.config.max_num_seqs
        for w0 in range(0, len(seqs), wave):
            chunk = seqs[w0:w0 + wave]
            for batch in self._prefill_batches(chunk):
                logits = self.run_prefill(batch, params)
                temps = torch.tensor([s.sampling_params.temperature for s in batch], device=self.device)
                first = self.sampler(logits, temps).tolist()
                for s, t in zip(batch, first):
                    s.status = SequenceStatus.RUNNING
                    self._accept(s, t)
            self.run_decode_loop(chunk)
            self.kv.free_sequences([s.seq_id for s in chunk])
