"""One process driving two GPUs: every kernel with a raised dynamic-shared-memory limit must raise it per DEVICE
(`ensure_max_dyn_smem`, csrc/common.cuh), not once per process — the round-1 library launched with the default limit
on the second device and failed.  Skipped on single-GPU boxes."""
import math

import pytest
import torch

from helpers import fused_qkv, make_paged

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_large_smem_kernels_run_on_both_devices_of_one_process():
    from compactor_vllm.attention.sparse_decode_kernel import head_sparse_decode_attention
    from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
    from compactor_vllm.compression.compactor import approximate_leverage_scores, non_causal_attn_scores
    from compactor_vllm.compression.snapkv import query_aware_key_scores
    from compactor_vllm.utils.arguments import make_phi

    HQ, HKV, D, PS, N = 32, 8, 128, 128, 700
    results = []
    for index in (0, 1):
        dev = torch.device("cuda", index)
        with torch.cuda.device(dev):
            q, k, v = fused_qkv(N, HQ, HKV, D, torch.bfloat16, dev, seed=3)
            kc, vc, pt, bm = make_paged(1, HKV, D, PS, [[200] * HKV], torch.bfloat16, dev, seed=4)
            lens = torch.full((1, HKV), 200, dtype=torch.int32, device=dev)
            cu = torch.tensor([0, N], dtype=torch.int32, device=dev)
            out = causal_sparse_varlen_with_cache(q, k, v, kc, vc, lens, pt, bm, cu, N, 200, HKV, PS, 1 / math.sqrt(D))
            dec = head_sparse_decode_attention(q[:1].contiguous(), kc, vc, lens, pt, bm, HKV, PS)
            lev = approximate_leverage_scores(k.contiguous(), [N], make_phi(D, 48, torch.bfloat16, dev), normalize=True,
                                              chunk_size=512)
            chunk = non_causal_attn_scores(q.contiguous(), k.contiguous(), None, cu, N, 128, sm_scale=1.0, normalize=True)
            snap = query_aware_key_scores(q.contiguous(), k.contiguous(), cu, cu, w=32)
            torch.cuda.synchronize(dev)
            results.append([t.float().cpu() for t in (out, dec, lev, chunk, snap)])
    for a, b in zip(*results):  # same seeds, same kernels: same bits on both devices
        assert torch.equal(torch.isfinite(a), torch.isfinite(b))
        fin = torch.isfinite(a)
        assert torch.equal(a[fin], b[fin])
