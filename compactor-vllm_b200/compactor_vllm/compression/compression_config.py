"""Re-export: the compression value types live in compactor_vllm.api_types."""
from compactor_vllm.api_types import (  # noqa: F401
    BatchCompressionParams,
    CompressionMethod,
    SequenceCompressionParams,
)

__all__ = ["CompressionMethod", "SequenceCompressionParams", "BatchCompressionParams"]
