"""One request of the engine shell (reference: utils/sequence.py:10-83, same fields and helpers)."""
from __future__ import annotations

import enum
import itertools
from dataclasses import dataclass, field
from typing import List

from compactor_vllm.api_types import SamplingParams, SequenceCompressionParams


class SequenceStatus(enum.Enum):
    WAITING = enum.auto()
    RUNNING = enum.auto()
    FINISHED = enum.auto()


_ids = itertools.count()


@dataclass
class Sequence:
    prompt_token_ids: List[int]
    completion_token_ids: List[int] = field(default_factory=list)
    sampling_params: SamplingParams = field(default_factory=SamplingParams)
    compression_params: SequenceCompressionParams = field(default_factory=SequenceCompressionParams)
    status: SequenceStatus = SequenceStatus.WAITING
    seq_id: int = field(default_factory=lambda: next(_ids), init=False)
    num_tokens_processed: int = 0

    @property
    def num_prompt_tokens(self) -> int:
        return len(self.prompt_token_ids)

    @property
    def num_generated_tokens(self) -> int:
        return len(self.completion_token_ids)

    def add_new_token(self, token_id: int) -> None:
        if not self.completion_token_ids:
            self.num_tokens_processed += self.num_prompt_tokens
        self.completion_token_ids.append(int(token_id))
        self.num_tokens_processed += 1
