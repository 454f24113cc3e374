"""compactor_vllm — B200-native attention + KV-compression hot path.

Import paths mirror vnchari/compactor-vllm (src/compactor_vllm/__init__.py:7-17).  `LLM` is the
engine shell of SURVEY 8(f)-3 (core/llm_engine.py): single rank, synthetic weights, CUDA-graph
decode — it exists to drive the hot path through the reference's top-level call; tokenizer
download, HF checkpoint loading and multi-process tensor parallelism are outside this build.
"""
from compactor_vllm.compression import CompressionMethod
from compactor_vllm.config.engine_config import AttentionBackend, LLMConfig
from compactor_vllm.config.sampling_params import SamplingParams


def __getattr__(name):  # lazy: importing the package must not need torch.cuda or the native library
    if name == "LLM":
        from compactor_vllm.core.llm_engine import LLM
        return LLM
    raise AttributeError(name)


__all__ = ["LLMConfig", "LLM", "SamplingParams", "AttentionBackend", "CompressionMethod"]
