"""Rotary embedding on the native kernel `cvllm_rope_qk` (reference: layers/rotary_embedding.py:8-94).

`RotaryEmbedding(head_size, rotary_dim, max_position_embeddings, base, rope_scaling)` and
`get_rope(...)` keep the reference's names, arguments and the fp32 `cos_sin_cache`
`[max_position, 1, rotary_dim]` buffer (cos | sin, llama3 frequency scaling :31-60).
`forward(positions, query, key)` returns the rotated `(query, key)`.

`fused_qk_norm_rope` is the B200 extension: Qwen3's per-head q_norm / k_norm, the
Compactor sketch `X = K_pre . PHI` and the rotation in ONE pass over q and k — the reference runs
four kernels, each reading and writing q / k in full (models/qwen3.py:90-97,
compression/compactor.py:153-158).  There is no CPU / eager fallback.
"""
from __future__ import annotations

import math
from functools import lru_cache
from typing import Optional

import torch
from torch import nn

from compactor_vllm import _native as N


def _inverse_frequencies(rotary_dim: int, base: float, rope_scaling: tuple | None) -> torch.Tensor:
    exponents = torch.arange(0, rotary_dim, 2, dtype=torch.float, device="cpu") / rotary_dim
    inv_freq = 1.0 / (base ** exponents)
    if rope_scaling is None:
        return inv_freq
    rope_type, factor, low_freq_factor, high_freq_factor, original_max_pos = rope_scaling
    assert rope_type == "llama3"
    # llama3 scaling (reference :31-60): long wavelengths are slowed by `factor`, short ones kept,
    # the band in between interpolated
    wavelen = 2 * math.pi / inv_freq
    low_wavelen = original_max_pos / low_freq_factor
    high_wavelen = original_max_pos / high_freq_factor
    scaled = torch.where(wavelen > low_wavelen, inv_freq / factor, inv_freq)
    smooth = (original_max_pos / wavelen - low_freq_factor) / (high_freq_factor - low_freq_factor)
    blended = (1 - smooth) * scaled / factor + smooth * scaled
    in_band = ~(wavelen < high_wavelen) * ~(wavelen > low_wavelen)
    return torch.where(in_band, blended, scaled)


def build_cos_sin_cache(rotary_dim: int, max_position: int, base: float,
                        rope_scaling: tuple | None = None) -> torch.Tensor:
    """fp32 [max_position, 1, rotary_dim]: cos(t * inv_freq) | sin(t * inv_freq) (reference :62-67)."""
    inv_freq = _inverse_frequencies(rotary_dim, base, rope_scaling)
    # always on the CPU, whatever the default device: cos / sin differ by an ulp between CPU and CUDA libm and
    # the table is pinned to the reference modules' CPU-built golden vectors
    t = torch.arange(max_position, dtype=torch.float, device="cpu")
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).unsqueeze_(1)


def fused_qk_norm_rope(
    positions: torch.Tensor,      # [N] int64
    query: torch.Tensor,          # [N, HQ, D] (or [N, HQ*D]), last dim contiguous
    key: torch.Tensor,            # [N, HKV, D] (or [N, HKV*D])
    cos_sin_cache: torch.Tensor,  # fp32 [max_position, 1, D]
    *,
    head_size: int = 128,
    q_norm_weight: Optional[torch.Tensor] = None,
    k_norm_weight: Optional[torch.Tensor] = None,
    eps: float = 1e-6,
    PHI: Optional[torch.Tensor] = None,  # [D, 48]: also return the sketch of the pre-RoPE keys
    inplace: bool = False,
):
    """-> (query_rot, key_rot, sketch_rows | None); outputs keep the input shapes.

    sketch_rows [N, HKV, 48] in K's dtype feeds
    `approximate_leverage_scores(None, ..., sketch_rows=...)` /
    `CompactorCompression.pre_rope_scoring_from_sketch`."""
    D = head_size
    def heads(t):
        return t.shape[1] if t.dim() == 3 else t.shape[1] // D
    q3 = query.view(query.shape[0], heads(query), D)
    k3 = key.view(key.shape[0], heads(key), D)
    Ntok, HQ, _ = q3.shape
    HKV = k3.shape[1]
    assert k3.shape[0] == Ntok and positions.shape[0] == Ntok
    assert q3.stride(-1) == 1 and k3.stride(-1) == 1 and q3.dtype == k3.dtype
    dev = q3.device
    pos = positions if positions.dtype == torch.int64 else positions.to(torch.int64)
    pos = pos.contiguous()
    cache = cos_sin_cache.view(cos_sin_cache.shape[0], -1)
    assert cache.dtype == torch.float32 and cache.shape[1] == D and cache.is_contiguous()
    if inplace:
        qo, ko = q3, k3
    else:
        qo = torch.empty((Ntok, HQ, D), dtype=q3.dtype, device=dev)
        ko = torch.empty((Ntok, HKV, D), dtype=k3.dtype, device=dev)
    sk = phi = None
    if PHI is not None:
        phi = PHI.contiguous()
        assert phi.shape == (D, 48) and phi.dtype == k3.dtype
        sk = torch.empty((Ntok, HKV, 48), dtype=k3.dtype, device=dev)
    qw = None if q_norm_weight is None else q_norm_weight.to(q3.dtype).contiguous()
    kw = None if k_norm_weight is None else k_norm_weight.to(k3.dtype).contiguous()
    rc = N.lib().cvllm_rope_qk(
        N.ptr(q3), N.ptr(k3), N.ptr(qo), N.ptr(ko), q3.stride(0), q3.stride(1), k3.stride(0), k3.stride(1),
        qo.stride(0), qo.stride(1), ko.stride(0), ko.stride(1), N.ptr(pos), N.ptr(cache), cache.shape[0],
        N.ptr(qw), N.ptr(kw), float(eps), N.ptr(phi), N.ptr(sk), Ntok, HQ, HKV, D, 48,
        N.dtype_code(q3.dtype), N.stream(dev),
    )
    N.check(rc, "cvllm_rope_qk")
    if inplace:
        return query, key, sk
    return qo.view(query.shape), ko.view(key.shape), sk


class RotaryEmbedding(nn.Module):
    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float,
                 rope_scaling: tuple | None) -> None:
        super().__init__()
        self.head_size = head_size
        assert rotary_dim == head_size
        cache = build_cos_sin_cache(rotary_dim, max_position_embeddings, base, rope_scaling)
        self.register_buffer("cos_sin_cache", cache, persistent=False)

    def _apply(self, fn, *args, **kwargs):
        # `model.to(bfloat16)` must not round the angle table: follow device moves, stay fp32 (the reference
        # builds its model under the target dtype and computes this buffer in float explicitly, :62-67)
        cache = self.cos_sin_cache
        super()._apply(fn, *args, **kwargs)
        self.cos_sin_cache = cache.to(device=fn(cache).device)
        return self

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
        q, k, _ = fused_qk_norm_rope(positions, query, key, self.cos_sin_cache, head_size=self.head_size)
        return q, k

    def forward_fused(self, positions, query, key, *, q_norm=None, k_norm=None, PHI=None, inplace=False):
        """q_norm / k_norm: `RMSNorm` modules (Qwen3) or None; PHI: the Compactor sketch matrix."""
        return fused_qk_norm_rope(
            positions, query, key, self.cos_sin_cache, head_size=self.head_size,
            q_norm_weight=None if q_norm is None else q_norm.weight,
            k_norm_weight=None if k_norm is None else k_norm.weight,
            eps=(q_norm or k_norm).eps if (q_norm is not None or k_norm is not None) else 1e-6,
            PHI=PHI, inplace=inplace,
        )


@lru_cache(1)
def get_rope(head_size: int, rotary_dim: int, max_position: int, base: float,
             rope_scaling: tuple | None = None):
    return RotaryEmbedding(head_size, rotary_dim, max_position, base, rope_scaling)
