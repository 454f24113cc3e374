"""`LLM` — the engine shell's public entry (reference: core/llm_engine.py:34-294).

Same call shape as the reference's `LLMEngine.generate`: prompts (token-id lists, or strings when a
tokenizer is available locally), `SamplingParams`, `BatchCompressionParams`, optional per-sequence
`SequenceCompressionParams`; returns the completions (decoded strings with a tokenizer, token-id lists
without) and optionally the `Sequence` objects.  The model runs with synthetic weights unless the
caller loads a state dict into `llm.model_runner.model` — there are no checkpoints on the GPU box.
"""
from __future__ import annotations

from typing import Any, List, Optional, Union

from compactor_vllm.api_types import BatchCompressionParams, SamplingParams, SequenceCompressionParams
from compactor_vllm.config.engine_config import LLMConfig
from compactor_vllm.core.model_runner import ModelRunner
from compactor_vllm.utils.sequence import Sequence

PromptLike = Union[str, List[int]]


class LLMEngine:
    def __init__(self, config: LLMConfig, *, tokenizer: Any = None, **runner_kwargs):
        self.config = config
        self.tokenizer = tokenizer
        if tokenizer is None and config.path is not None:
            try:  # local files only: the GPU box has no network
                from transformers import AutoTokenizer
                self.tokenizer = AutoTokenizer.from_pretrained(config.path, local_files_only=True)
            except Exception:
                self.tokenizer = None
        if self.tokenizer is not None and config.eos == -1:
            eos = getattr(self.tokenizer, "eos_token_id", None)
            config.eos = -1 if eos is None else int(eos)  # token id 0 is a legitimate EOS
        self.model_runner = ModelRunner(config, **runner_kwargs)
        self.master_model_runner = self.model_runner  # the reference's attribute name

    def tokenize_prompt(self, prompt: PromptLike, **kw) -> List[int]:
        if isinstance(prompt, str):
            if self.tokenizer is None:
                raise ValueError("string prompts need a tokenizer; pass token-id lists instead")
            return list(self.tokenizer(prompt, **kw)["input_ids"])
        return [int(t) for t in prompt]

    def _build_sequences(self, prompts, sampling_params, per_sequence_compression_params=None,
                         tokenizer_kwargs=None) -> List[Sequence]:
        """core/llm_engine.py:99-153, including the rule that a prompt no longer than its protected ends is
        not compressed (:144-145)."""
        if not isinstance(prompts, list) or (prompts and isinstance(prompts[0], int)):
            prompts = [prompts]
        n = len(prompts)
        sp = [sampling_params] * n if isinstance(sampling_params, SamplingParams) else list(sampling_params)
        assert len(sp) == n, "sampling_params list must match prompts length"
        if per_sequence_compression_params is None:
            cp = [SequenceCompressionParams(1.0) for _ in range(n)]
        elif isinstance(per_sequence_compression_params, SequenceCompressionParams):
            cp = [per_sequence_compression_params] * n
        else:
            cp = list(per_sequence_compression_params)
            assert len(cp) == n, "per_sequence_compression_params list must match prompts length"
        seqs = []
        for prompt, s, c in zip(prompts, sp, cp):
            ids = self.tokenize_prompt(prompt, **(tokenizer_kwargs or {}))
            if c.protected_first_tokens + c.protected_last_tokens >= len(ids):
                c = SequenceCompressionParams(1.0, c.protected_first_tokens, c.protected_last_tokens)
            seqs.append(Sequence(prompt_token_ids=ids, sampling_params=s, compression_params=c))
        return seqs

    def generate(self, prompts, sampling_params, batch_compression_params: Optional[BatchCompressionParams] = None,
                 *, per_sequence_compression_params=None, tokenizer_kwargs=None, detokenizer_kwargs=None,
                 return_sequences: bool = False):
        batch_compression_params = batch_compression_params or BatchCompressionParams()
        seqs = self._build_sequences(prompts, sampling_params, per_sequence_compression_params, tokenizer_kwargs)
        self.model_runner.generate(seqs, batch_compression_params)
        if self.tokenizer is not None:
            outs = self.tokenizer.batch_decode([s.completion_token_ids for s in seqs], **(detokenizer_kwargs or {}))
        else:
            outs = [list(s.completion_token_ids) for s in seqs]
        return (outs, seqs) if return_sequences else outs


LLM = LLMEngine
