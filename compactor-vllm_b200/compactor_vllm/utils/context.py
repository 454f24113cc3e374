"""Per-step state shared between the runner and the attention / compression code.

The reference keeps one module-global `Context` (utils/context.py:10-83) that the runner replaces
before every forward pass; models and `Attention.forward` read it through `get_context()`.  Same
surface here (`Context`, `CompressionContext`, `get_context`, `set_context`, `reset_context`, same
field names), built from one field table so the two stay in step.
"""
from __future__ import annotations

import dataclasses
from typing import Any

from compactor_vllm.api_types import CompressionMethod
from compactor_vllm.config.engine_config import AttentionBackend

_COMPRESSION_FIELDS: dict[str, Any] = {
    "compression_method": CompressionMethod.COMPACTOR,
    "compression_chunk_size": -1,        # leverage-score chunk (tokens); -1 = whole sequence
    "batch_tokens_to_retain": None,      # int32 [B] device tensor: (token, head) pairs kept per sequence
    "max_tokens_to_retain": 0,           # max_len * HKV_local
    "context_lens": None,                # host list [B]
    "PHI": None,                         # [D, sketch] Gaussian sketch, model dtype
    "protected_first_tokens": None,      # host list [B]
    "protected_last_tokens": None,       # host list [B]
}

_STEP_FIELDS: dict[str, Any] = {
    "is_prefill": False,
    "do_compression": False,
    "cu_seqlens_q": None,
    "cu_seqlens_k": None,
    "max_seqlen_q": 0,
    "max_seqlen_k": 0,
    "batch_mapping": None,               # int32 [B]: batch row -> cache slot (slot 0 is reserved)
    "max_bh_len": 0,
    "compression_context": None,
    "STORE_STREAM": None,                # side stream for scoring + store during prefill
    "key_split": None,                   # decode split count fixed at CUDA-graph capture time
    "attention_backend": AttentionBackend.COMPACTOR_TRITON,
}


def _record_type(name: str, table: dict[str, Any], doc: str):
    """A real dataclass (so `dataclasses.replace / asdict / fields` work as on the reference's classes,
    utils/context.py:10-40) generated from a field table: every field optional, positional order = table order."""
    cls = dataclasses.make_dataclass(name, [(key, Any, dataclasses.field(default=default))
                                            for key, default in table.items()])
    cls.__doc__ = doc
    cls.__module__ = __name__
    return cls


CompressionContext = _record_type("CompressionContext", _COMPRESSION_FIELDS,
                                  "Per-step compression arguments (reference utils/context.py:10-22).")
Context = _record_type("Context", _STEP_FIELDS, "Per-step attention state (reference utils/context.py:25-40).")


_current = Context()


def get_context() -> Context:
    return _current


def set_context(*, is_prefill, **step) -> None:
    """Replace the global context; unspecified fields take their defaults."""
    global _current
    _current = Context(is_prefill=is_prefill, **step)


def reset_context() -> None:
    global _current
    _current = Context()
