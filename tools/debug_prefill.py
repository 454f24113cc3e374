"""GPU diagnostic: where does the prefill kernel disagree with fp32 torch?"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from helpers import fused_qkv, make_paged, torch_prefill_fp32

DEV = "cuda"
def case(name, appends, lens, HQ, HKV, PS=128, dtype=torch.bfloat16, seed=0):
    B, D = len(appends), 128
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens, dtype, DEV, seed=seed)
    cu = torch.tensor([0] + torch.tensor(appends).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    q, k, v = fused_qkv(int(cu[-1]), HQ, HKV, D, dtype, DEV, seed=seed + 1)
    l = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = causal_sparse_varlen_with_cache(q, k, v, kc, vc, l, pt, bm, cu, max(appends), int(l.max()), HKV, PS)
    torch.cuda.synchronize()
    ref = torch_prefill_fp32(q, k, v, kc, vc, l, pt, bm, cu, HKV, PS)
    d = (out.float() - ref).abs()
    print(f"{name}: max err {d.max().item():.4f}  nan={torch.isnan(out.float()).any().item()}")
    cul = cu.tolist()
    for b in range(B):
        for t0 in range(cul[b], cul[b + 1], 128):
            t1 = min(t0 + 128, cul[b + 1])
            e = d[t0:t1].amax(dim=(0, 2)).tolist()
            rows = d[t0:t1].amax(dim=(1, 2))
            print(f"   b{b} qtile@{t0 - cul[b]:5d}: per-head max " + " ".join(f"{x:.3f}" for x in e[:8]) +
                  f" | worst row {int(rows.argmax())} ({rows.max().item():.3f}) first-rows " +
                  " ".join(f"{x:.3f}" for x in rows[:4].tolist()))

case("1 tile, no prefix", [128], [[0]], 1, 1)
case("2 tiles", [256], [[0]], 1, 1)
case("4 tiles", [512], [[0]], 1, 1)
case("short", [37], [[0]], 1, 1)
case("prefix 70 + 37", [37], [[70]], 1, 1)
case("prefix 128 + 37", [37], [[128]], 1, 1)
case("prefix 256 + 150", [150], [[256]], 1, 1)
case("gqa", [200], [[0, 0]], 8, 2)
case("multi-seq", [37, 1, 150], [[0, 70], [129, 5], [256, 0]], 8, 2)
case("long 2048", [2048], [[0]], 4, 1)
case("f16 long 2048", [2048], [[0]], 4, 1, dtype=torch.float16)
