"""Re-export: `SamplingParams` lives in compactor_vllm.api_types."""
from compactor_vllm.api_types import SamplingParams  # noqa: F401

__all__ = ["SamplingParams"]
