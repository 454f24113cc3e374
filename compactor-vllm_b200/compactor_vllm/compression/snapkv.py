"""SnapKV scoring (reference: compression/snapkv.py:12-448)."""
from __future__ import annotations

import math
from typing import Optional

import torch
from compactor_vllm import _native as N
from compactor_vllm.compression.common import BaseCompressionMethod
from compactor_vllm.utils.helpers import maybe_execute_in_stream

# The reference pools inside the BLOCK_K tile its autotuner happened to pick
# (32 / 64 / 128, snapkv.py:160-168,253-262), so pooled values at tile borders
# depend on the autotune outcome.  Here the segment is an explicit constant.
POOL_SEGMENT = 64
KPOOL = 5


class SnapKVCompression(BaseCompressionMethod):
    @staticmethod
    def pre_rope_scoring(q, k, v, context) -> Optional[torch.Tensor]:
        return None

    @staticmethod
    def post_rope_scoring(q, k, v, pre_rope_scores, context) -> Optional[torch.Tensor]:
        return maybe_execute_in_stream(
            query_aware_key_scores, q, k, context.cu_seqlens_q, context.cu_seqlens_k, w=32,
            STORE_STREAM=context.STORE_STREAM,
        )


def query_aware_key_scores(
    q: torch.Tensor,  # [N_q, Hq, D]
    k: torch.Tensor,  # [N_k, Hk, D]
    cu_seqlens_q: torch.Tensor,  # [B+1], int32
    cu_seqlens_k: torch.Tensor,  # [B+1], int32
    w: int,
    sm_scale: float = None,
    *,
    accum_scores: torch.Tensor = None,
    accum_blending: float = None,
    normalize: bool = False,
    pool_segment: int = POOL_SEGMENT,
) -> Optional[torch.Tensor]:
    """Attention mass from the last-w queries (x G heads) onto keys [0, L-w).

    Reference: compression/snapkv.py:332-448 (K12 :39-157, K13 :160-276):
    softmax over the non-window keys only, column sums over the w*G rows,
    causal KPOOL-tap mean pool inside pool_segment-key segments, window keys
    <- +inf.  The reference's [N, HKV, w*G] fp32 logits scratch does not exist
    here (two passes over K instead).  Sequences with L <= w keep score 0 for
    every key (uninitialised memory in the reference, :94-95,204-205).
    """
    assert q.stride(-1) == 1 and k.stride(-1) == 1, "last dim must be contiguous"
    assert isinstance(w, int), "per-sequence window tensors are not supported"
    assert not normalize, "normalize=True is not used by the engine path (snapkv.py:27-35)"
    N_q, Hq, D = q.shape
    N_k, Hk, Dk = k.shape
    assert (Hq % Hk) == 0, "Hq must be a multiple of Hk"
    assert q.stride(1) == D and k.stride(1) == D
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(D)
    B = cu_seqlens_q.numel() - 1
    assert B == cu_seqlens_k.numel() - 1
    dev = q.device
    if w * (Hq // Hk) == 0:
        return torch.zeros((N_k, Hk), dtype=torch.float32, device=dev)
    out = torch.empty((N_k, Hk), dtype=torch.float32, device=dev)
    cu_q = cu_seqlens_q.to(device=dev, dtype=torch.int32).contiguous()
    cu_k = cu_seqlens_k.to(device=dev, dtype=torch.int32).contiguous()
    L = N.lib()
    ws_bytes = L.cvllm_snapkv_scores_workspace_bytes(B, N_k, Hq, Hk, w)
    ws = N.workspace(dev, ws_bytes, "snapkv")
    rc = L.cvllm_snapkv_scores(
        N.ptr(q), N.ptr(k), q.stride(0), k.stride(0), N.ptr(cu_q), N.ptr(cu_k), w, B, N_k, Hq, Hk,
        D, N_k, float(sm_scale), KPOOL, int(pool_segment), N.ptr(out), N.dtype_code(q.dtype),
        N.ptr(ws), ws.numel(), N.stream(dev),
    )
    N.check(rc, "cvllm_snapkv_scores")
    if accum_scores is not None:
        if accum_blending is not None:
            accum_scores.mul_(accum_blending)
        accum_scores.add_(out)
        return accum_scores
    return out
