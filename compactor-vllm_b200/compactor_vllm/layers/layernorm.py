"""RMSNorm (reference: layers/layernorm.py:5-49) on the native kernel `cvllm_rmsnorm`.

Same module surface: `RMSNorm(hidden_size, eps).forward(x, residual=None)` returns the normed
tensor, or `(normed, new_residual)` when a residual is given.  The reference's two torch.compile'd
functions become one CTA-per-row launch with the reference's rounding points
(x.float() * rsqrt(mean + eps) -> model dtype -> * weight).  No CPU / eager fallback.
"""
from __future__ import annotations

import torch
from torch import nn

from compactor_vllm import _native as N


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """x [..., width] (last dim contiguous, rows uniformly strided); residual is updated IN PLACE
    like the reference's add_rms_forward result (it returns the new residual)."""
    width = x.shape[-1]
    x2 = x.reshape(-1, width)
    out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    o2 = out.view(-1, width)
    r2 = None
    if residual is not None:
        assert residual.shape == x.shape and residual.dtype == x.dtype
        r2 = residual.view(-1, width)
    w = weight if weight.dtype == x.dtype else weight.to(x.dtype)
    assert x2.stride(-1) == 1 and w.is_contiguous()
    rc = N.lib().cvllm_rmsnorm(
        N.ptr(x2), N.ptr(r2), N.ptr(w), N.ptr(o2), x2.stride(0), r2.stride(0) if r2 is not None else 0,
        o2.stride(0), x2.shape[0], width, float(eps), N.dtype_code(x.dtype), N.stream(x.device),
    )
    N.check(rc, "cvllm_rmsnorm")
    return out if residual is None else (out, residual)


class RMSNorm(nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def rms_forward(self, x: torch.Tensor) -> torch.Tensor:
        return rms_norm(x, self.weight, self.eps)

    def add_rms_forward(self, x: torch.Tensor, residual: torch.Tensor):
        return rms_norm(x, self.weight, self.eps, residual)

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None):
        if residual is None:
            return self.rms_forward(x)
        return self.add_rms_forward(x, residual)
