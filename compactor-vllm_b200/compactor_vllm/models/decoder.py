"""Decoder-only transformer around the hot path (reference: models/llama3.py:20-281, models/qwen3.py:21-272).

Engine shell of SURVEY 8(f)-3: the model exists so that `LLM.generate` drives the attention +
KV-compression path the way the reference's models do.  The attention block is the B200 one — fused
q/k norm + sketch + RoPE (`cvllm_rope_qk`), scoring on the store stream, `Attention.forward` — and
the norms are `cvllm_rmsnorm`; the projections / MLP / embedding are plain library GEMMs
(`torch.nn.functional.linear` = cuBLAS), which is what they are in the reference too
(layers/linear.py).  Single rank (tensor_parallel_size == 1): the TP all-reduce after o_proj is
SURVEY 8(f)-4.  Weights: random init from a seed (`init_synthetic_weights`) or a state dict with the
reference's packed names (`qkv_proj`, `gate_up_proj`) via `load_state_dict`; there are no checkpoint
files on the GPU box.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from compactor_vllm.api_types import CompressionMethod
from compactor_vllm.compression import apply_postrope_compression, apply_prerope_compression
from compactor_vllm.compression.compactor import CompactorCompression
from compactor_vllm.layers.attention import Attention
from compactor_vllm.layers.layernorm import RMSNorm
from compactor_vllm.layers.rotary_embedding import get_rope
from compactor_vllm.utils.context import get_context


def _rope_scaling_tuple(cfg):
    rs = getattr(cfg, "rope_scaling", None)
    if not rs:
        return None
    kind = rs.get("rope_type", rs.get("type"))
    if kind != "llama3":
        return None
    return ("llama3", float(rs["factor"]), float(rs["low_freq_factor"]), float(rs["high_freq_factor"]),
            int(rs["original_max_position_embeddings"]))


class DecoderAttention(nn.Module):
    """models/llama3.py:20-112 / models/qwen3.py:21-104 (qk_norm=True)."""

    def __init__(self, cfg, qk_norm: bool):
        super().__init__()
        self.num_heads = cfg.num_attention_heads
        self.num_kv_heads = cfg.num_key_value_heads
        self.head_dim = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
        self.q_size = self.num_heads * self.head_dim
        self.kv_size = self.num_kv_heads * self.head_dim
        bias = bool(getattr(cfg, "attention_bias", False))
        self.qkv_proj = nn.Linear(cfg.hidden_size, self.q_size + 2 * self.kv_size, bias=bias)
        self.o_proj = nn.Linear(self.q_size, cfg.hidden_size, bias=False)
        self.rotary_emb = get_rope(self.head_dim, self.head_dim, cfg.max_position_embeddings,
                                   float(getattr(cfg, "rope_theta", 10000.0)), _rope_scaling_tuple(cfg))
        self.attn = Attention(self.num_heads, self.head_dim, self.head_dim ** -0.5, self.num_kv_heads)
        eps = getattr(cfg, "rms_norm_eps", 1e-6)
        self.q_norm = RMSNorm(self.head_dim, eps=eps) if qk_norm else None
        self.k_norm = RMSNorm(self.head_dim, eps=eps) if qk_norm else None

    def forward(self, positions: torch.Tensor, hidden_states: torch.Tensor) -> torch.Tensor:
        ctx = get_context()
        qkv = self.qkv_proj(hidden_states)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        q = q.view(-1, self.num_heads, self.head_dim)
        k = k.view(-1, self.num_kv_heads, self.head_dim)
        v = v.view(-1, self.num_kv_heads, self.head_dim)
        compress = bool(ctx.is_prefill and ctx.do_compression)
        cc = ctx.compression_context
        compactor = compress and cc.compression_method is CompressionMethod.COMPACTOR
        # one pass: q/k norm (Qwen3) -> sketch of the pre-RoPE keys (COMPACTOR) -> rotation
        q, k, sketch = self.rotary_emb.forward_fused(
            positions, q, k, q_norm=self.q_norm, k_norm=self.k_norm, PHI=cc.PHI if compactor else None)
        scores = None
        if compactor:
            scores = CompactorCompression.pre_rope_scoring_from_sketch(sketch, ctx)
            scores = apply_postrope_compression(q, k, v, scores, ctx)
        elif compress:  # SNAPKV / NONE: nothing to score before the rotation
            scores = apply_postrope_compression(q, k, v, None, ctx)
        o = self.attn(q, k, v, scores)
        return self.o_proj(o.flatten(1, -1))


class DecoderMLP(nn.Module):
    """models/llama3.py:115-140: down(silu(gate) * up), gate and up packed."""

    def __init__(self, cfg):
        super().__init__()
        self.gate_up_proj = nn.Linear(cfg.hidden_size, 2 * cfg.intermediate_size, bias=False)
        self.down_proj = nn.Linear(cfg.intermediate_size, cfg.hidden_size, bias=False)

    def forward(self, x):
        gate, up = self.gate_up_proj(x).chunk(2, dim=-1)
        return self.down_proj(F.silu(gate) * up)


class DecoderLayer(nn.Module):
    def __init__(self, cfg, qk_norm: bool):
        super().__init__()
        eps = getattr(cfg, "rms_norm_eps", 1e-6)
        self.self_attn = DecoderAttention(cfg, qk_norm)
        self.mlp = DecoderMLP(cfg)
        self.input_layernorm = RMSNorm(cfg.hidden_size, eps=eps)
        self.post_attention_layernorm = RMSNorm(cfg.hidden_size, eps=eps)

    def forward(self, positions, hidden_states, residual):
        # models/llama3.py:165-177: the residual stream is carried beside the normed activations
        if residual is None:
            residual = hidden_states  # updated in place by the later add-norms; nobody else holds it
            hidden_states = self.input_layernorm(hidden_states)
        else:
            hidden_states, residual = self.input_layernorm(hidden_states, residual)
        hidden_states = self.self_attn(positions, hidden_states)
        hidden_states, residual = self.post_attention_layernorm(hidden_states, residual)
        return self.mlp(hidden_states), residual


class DecoderModel(nn.Module):
    def __init__(self, cfg, qk_norm: bool):
        super().__init__()
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.layers = nn.ModuleList([DecoderLayer(cfg, qk_norm) for _ in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, eps=getattr(cfg, "rms_norm_eps", 1e-6))

    def forward(self, input_ids, positions):
        h = self.embed_tokens(input_ids)
        residual = None
        for layer in self.layers:
            h, residual = layer(positions, h, residual)
        h, _ = self.norm(h, residual)
        return h


class DecoderForCausalLM(nn.Module):
    qk_norm = False

    def __init__(self, cfg):
        super().__init__()
        self.config = cfg
        self.model = DecoderModel(cfg, self.qk_norm)
        self.lm_head = nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False)
        if getattr(cfg, "tie_word_embeddings", False):
            self.lm_head.weight = self.model.embed_tokens.weight

    def forward(self, input_ids, positions):
        return self.model(input_ids, positions)

    def compute_logits(self, hidden_states):
        """layers/embed_head.py:55-69: prefill keeps the last token of every sequence."""
        ctx = get_context()
        if ctx.is_prefill:
            hidden_states = hidden_states[ctx.cu_seqlens_q[1:].long() - 1].contiguous()
        return self.lm_head(hidden_states)

    @torch.no_grad()
    def init_synthetic_weights(self, seed: int = 0) -> None:
        """Random weights of the architecture (no checkpoints without a network): N(0, 0.02) matrices —
        the HF initializer_range — and norm gains near 1, drawn on the parameters' own device (an 8B model is
        16 GB: not something to generate on the host)."""
        for i, (name, p) in enumerate(self.named_parameters()):
            gen = torch.Generator(device=p.device).manual_seed(seed * 100003 + i)
            val = torch.randn(p.shape, generator=gen, device=p.device, dtype=torch.float32)
            if name.endswith("norm.weight") or name.endswith("layernorm.weight"):
                val = 1.0 + 0.1 * val
            else:
                val = 0.02 * val
            p.copy_(val.to(p.dtype))


class LlamaForCausalLM(DecoderForCausalLM):
    qk_norm = False


class Qwen3ForCausalLM(DecoderForCausalLM):
    qk_norm = True
