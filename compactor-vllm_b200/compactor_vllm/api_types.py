"""User-facing value types of the hot path, gathered in one place.

These mirror the reference's public names and defaults so callers switch over unchanged
(field-by-field: compression/compression_config.py:8-44 and config/sampling_params.py:5-11 of the
reference); the modules of the same name re-export them.
"""
from __future__ import annotations

import enum
import logging
from dataclasses import dataclass, field

_log = logging.getLogger("compactor_vllm.compression")


class CompressionMethod(enum.Enum):
    """Which scorer ranks the (token, head) pairs of a prompt."""

    COMPACTOR = 1  # leverage-score sketch (pre-RoPE) + chunk-local attention mass (post-RoPE)
    SNAPKV = 2     # attention mass of the last-w query window
    NONE = 3       # keep everything


@dataclass
class SequenceCompressionParams:
    """Per-request knobs: fraction of (token, head) pairs kept and the always-kept prompt ends."""

    compression_ratio: float = 1.0
    protected_first_tokens: int = 16
    protected_last_tokens: int = 64


@dataclass
class BatchCompressionParams:
    """Per-batch knobs.  SnapKV scores a whole prompt at once, so chunked compression is switched
    off for it (the reference does the same and warns)."""

    compression_method: CompressionMethod = CompressionMethod.COMPACTOR
    do_chunked_compression: bool = True
    chunk_size: int = 512

    def __post_init__(self) -> None:
        if self.compression_method is CompressionMethod.SNAPKV and self.do_chunked_compression:
            _log.warning("SNAPKV scores the whole prompt: chunked compression disabled")
            self.do_chunked_compression = False


@dataclass
class SamplingParams:
    """Decoding knobs of the engine shell (not used by the attention / compression path)."""

    temperature: float = 1.0
    max_new_tokens: int = 256
    _checked: bool = field(default=False, init=False, repr=False, compare=False)

    def __post_init__(self) -> None:
        if not (self.temperature >= 0):
            raise ValueError(f"temperature must be >= 0, got {self.temperature}")
        self._checked = True
