"""compactor_vllm — B200-native attention + KV-compression hot path.

Import paths mirror vnchari/compactor-vllm (src/compactor_vllm/__init__.py:1-17)
for the hot-path surface.  The generation engine (`LLM`: tokenizer, scheduler,
model runner, HF weight loading) is out of scope for this build — SURVEY.md
section 8(f) lists it as "next" — so `LLM` raises on construction instead of
pretending to work.
"""
from compactor_vllm.compression import CompressionMethod
from compactor_vllm.config.engine_config import AttentionBackend, LLMConfig
from compactor_vllm.config.sampling_params import SamplingParams


class LLM:
    def __init__(self, *args, **kwargs):
        raise NotImplementedError(
            "The generation engine is outside this build's scope (attention + KV-compression "
            "hot path only); drive compactor_vllm.layers.attention.Attention from your runner."
        )


__all__ = ["LLMConfig", "LLM", "SamplingParams", "AttentionBackend", "CompressionMethod"]
