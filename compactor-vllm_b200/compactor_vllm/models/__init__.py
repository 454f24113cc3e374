"""Model registry of the engine shell (reference: models/__init__.py:1-9, core/model_runner.py:44-52)."""
from compactor_vllm.models.decoder import LlamaForCausalLM, Qwen3ForCausalLM

MODEL_REGISTRY = {"llama": LlamaForCausalLM, "qwen3": Qwen3ForCausalLM}

__all__ = ["LlamaForCausalLM", "Qwen3ForCausalLM", "MODEL_REGISTRY"]
