"""Oracle for the step in front of the hot path: per-head RMSNorm, the Compactor
sketch of the pre-RoPE keys, and the rotary embedding.  TEST INFRASTRUCTURE ONLY (see __init__).

Plain torch on CPU in the reference's eager arithmetic; pinned by tests/golden/rope_*.npz and
rmsnorm_*.npz, which oracle/gen_golden.py produced by running the reference's own modules
(layers/rotary_embedding.py, layers/layernorm.py) with TorchDynamo disabled.
"""
from __future__ import annotations

import math

import torch


def llama3_inv_freq(rotary_dim: int, base: float, rope_scaling):
    """layers/rotary_embedding.py:31-60."""
    inv = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    if rope_scaling is None:
        return inv
    _, factor, lo, hi, old_len = rope_scaling
    wl = 2 * math.pi / inv
    out = []
    for f, w in zip(inv.tolist(), wl.tolist()):
        # per-frequency restatement of the reference's three torch.where branches, in fp32
        f32 = torch.tensor(f, dtype=torch.float)
        w32 = torch.tensor(w, dtype=torch.float)
        scaled = f32 / factor if w32 > old_len / lo else f32
        if not (w32 < old_len / hi) and not (w32 > old_len / lo):
            s = (old_len / w32 - lo) / (hi - lo)
            scaled = (1 - s) * scaled / factor + s * scaled
        out.append(scaled)
    return torch.stack(out).float()


def cos_sin_cache(rotary_dim: int, max_pos: int, base: float, rope_scaling=None) -> torch.Tensor:
    """layers/rotary_embedding.py:62-67 -> fp32 [max_pos, rotary_dim] (cos | sin)."""
    inv = llama3_inv_freq(rotary_dim, base, rope_scaling)
    t = torch.arange(max_pos, dtype=torch.float)
    fr = t[:, None] * inv[None, :]
    return torch.cat((fr.cos(), fr.sin()), dim=-1)


def rope(x: torch.Tensor, positions: torch.Tensor, cache: torch.Tensor) -> torch.Tensor:
    """layers/rotary_embedding.py:8-16,76-79.  x [N, H, D] 16-bit; cache fp32 [max_pos, D]."""
    cs = cache[positions.long()][:, None, :]
    half = x.shape[-1] // 2
    c, s = cs[..., :half], cs[..., half:]
    xf = x.float()
    x1, x2 = xf[..., :half], xf[..., half:]
    return torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), dim=-1).to(x.dtype)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """layers/layernorm.py:16-40.  Returns y, or (y, new_residual)."""
    xf = x.float()
    new_res = None
    if residual is not None:
        xf = xf + residual.float()
        new_res = xf.to(x.dtype)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    y = (xf * torch.rsqrt(var + eps)).to(x.dtype) * weight.to(x.dtype)
    return y if residual is None else (y, new_res)


def sketch(k_pre: torch.Tensor, phi: torch.Tensor) -> torch.Tensor:
    """compression/compactor.py:153-158: X = K.PHI in K's dtype, fp32 accumulation -> [N, H, k]."""
    return (k_pre.float() @ phi.float()).to(k_pre.dtype)


def fused_qk_norm_rope(positions, q, k, cache, q_w=None, k_w=None, eps=1e-6, phi=None):
    """The fused kernel's contract, composed from the pieces above (models/qwen3.py:90-97)."""
    if q_w is not None:
        q = rms_norm(q, q_w, eps)
    if k_w is not None:
        k = rms_norm(k, k_w, eps)
    sk = sketch(k, phi) if phi is not None else None
    return rope(q, positions, cache), rope(k, positions, cache), sk
