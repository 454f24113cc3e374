"""Scorer plugins keyed by CompressionMethod, and the two hooks the model code calls around RoPE.

Boundary mirrored: the reference's compression/__init__.py:13-29 (`COMPRESSION_REGISTRY`,
`apply_prerope_compression`, `apply_postrope_compression`).
"""
from compactor_vllm.api_types import BatchCompressionParams, CompressionMethod, SequenceCompressionParams
from compactor_vllm.compression import common as _common
from compactor_vllm.compression import compactor as _compactor
from compactor_vllm.compression import snapkv as _snapkv

COMPRESSION_REGISTRY = {
    CompressionMethod.NONE: _common.NoCompression,
    CompressionMethod.SNAPKV: _snapkv.SnapKVCompression,
    CompressionMethod.COMPACTOR: _compactor.CompactorCompression,
}


def _plugin(context):
    return COMPRESSION_REGISTRY[context.compression_context.compression_method]


def apply_prerope_compression(q, k, v, context):
    """Scores that only need pre-RoPE tensors (COMPACTOR: leverage sketch of K); None otherwise."""
    return _plugin(context).pre_rope_scoring(q, k, v, context=context)


def apply_postrope_compression(q, k, v, prerope_scores, context):
    """Final [N, HKV] fp32 scores (or None) from post-RoPE q / k and the pre-RoPE partial scores."""
    return _plugin(context).post_rope_scoring(q, k, v, prerope_scores, context=context)


__all__ = [
    "COMPRESSION_REGISTRY",
    "apply_prerope_compression",
    "apply_postrope_compression",
    "BatchCompressionParams",
    "CompressionMethod",
    "SequenceCompressionParams",
]
