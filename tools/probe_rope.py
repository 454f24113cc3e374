"""Timing probe for cvllm_rope_qk at the headline size (32768 tokens, HQ=32, HKV=8): back-to-back
launches inside one CUDA-event pair so the Python launch path is hidden.  python tools/probe_rope.py"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
from compactor_vllm.layers.rotary_embedding import build_cos_sin_cache, fused_qk_norm_rope  # noqa: E402

dev = torch.device("cuda", 0)
N, HQ, HKV, D = 32768, 32, 8, 128
qkv = torch.randn((N, (HQ + 2 * HKV) * D), device=dev).bfloat16()
q = qkv[:, :HQ * D].view(N, HQ, D)
k = qkv[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D)
cache = build_cos_sin_cache(D, 40960, 500000.0, None).to(dev)
pos = torch.arange(N, device=dev)
w = torch.ones(D, device=dev).bfloat16()
phi = (torch.randn((D, 48), device=dev) / 48 ** 0.5).bfloat16()
qo = torch.empty((N, HQ, D), device=dev, dtype=torch.bfloat16)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps


by = N * (HQ + HKV) * D * 2 * 2
for name, fn in [
    ("copy q (torch)", lambda: qo.copy_(q)),
    ("rope", lambda: fused_qk_norm_rope(pos, q, k, cache)),
    ("rope inplace", lambda: fused_qk_norm_rope(pos, q, k, cache, inplace=True)),
    ("norm+rope", lambda: fused_qk_norm_rope(pos, q, k, cache, q_norm_weight=w, k_norm_weight=w)),
    ("rope+sketch", lambda: fused_qk_norm_rope(pos, q, k, cache, PHI=phi)),
    ("norm+rope+sketch", lambda: fused_qk_norm_rope(pos, q, k, cache, q_norm_weight=w, k_norm_weight=w, PHI=phi)),
]:
    ms = timed(fn)
    b_ = N * HQ * D * 2 * 2 if name.startswith("copy") else by
    print(f"{name:20s} {ms * 1e3:8.1f} us  {b_ / ms / 1e6:8.1f} GB/s")
