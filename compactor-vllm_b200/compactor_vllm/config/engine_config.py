"""Engine configuration surface (reference: config/engine_config.py:9-94).

Same fields, defaults and validation.  `hf_config` may be any object exposing
the HF attribute names the KV-cache manager reads (num_hidden_layers,
num_key_value_heads, head_dim, torch_dtype, max_position_embeddings); when it
is None it is resolved with transformers.AutoConfig exactly as the reference
does (needs the model files locally — there is no network on the GPU box).
"""
import os
from dataclasses import dataclass
from enum import Enum, auto
from typing import Any, Optional


class AttentionBackend(Enum):
    FLASH_ATTENTION = auto()
    COMPACTOR_TRITON = auto()  # name kept for drop-in; served by the sm_100a CUDA kernels here


@dataclass
class LLMConfig:
    model: str
    path: Optional[str] = None
    nccl_port: Optional[int] = 1218
    max_num_seqs: int = 256
    max_model_len: int = 40960
    gpu_memory_utilization: float = 0.9
    tensor_parallel_size: int = 1
    enforce_eager: bool = False
    hf_config: Any = None
    eos: int = -1
    kvcache_page_size: int = 128
    leverage_sketch_size: int = 48
    attention_backend: AttentionBackend = AttentionBackend.COMPACTOR_TRITON
    show_progress_bar: bool = True

    def __post_init__(self):
        if self.path is not None and not os.path.isdir(self.path):
            raise NotADirectoryError(f"Engine config dir {self.path} does not exist")
        if self.tensor_parallel_size <= 0 or self.tensor_parallel_size > 8:
            raise ValueError("tensor_parallel_size must be >= 1 and <= 8")
        if self.hf_config is None:
            from transformers import AutoConfig

            self.hf_config = AutoConfig.from_pretrained(self.model)
        self.max_model_len = min(self.max_model_len, self.hf_config.max_position_embeddings)
