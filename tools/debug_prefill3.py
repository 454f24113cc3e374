import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from helpers import fused_qkv, make_paged, torch_prefill_fp32
DEV = "cuda"
def case(name, append, Lc, dtype=torch.float16, seed=0, PS=128):
    kc, vc, pt, bm = make_paged(1, 1, 128, PS, [[Lc]], dtype, DEV, seed=seed)
    cu = torch.tensor([0, append], dtype=torch.int32, device=DEV)
    q, k, v = fused_qkv(append, 1, 1, 128, dtype, DEV, seed=seed + 1)
    l = torch.tensor([[Lc]], dtype=torch.int32, device=DEV)
    out = causal_sparse_varlen_with_cache(q, k, v, kc, vc, l, pt, bm, cu, append, Lc, 1, PS)
    torch.cuda.synchronize()
    ref = torch_prefill_fp32(q, k, v, kc, vc, l, pt, bm, cu, 1, PS)
    d = (out.float() - ref).abs().amax(dim=(1, 2))
    bad = (d > 0.02).nonzero().flatten().tolist()
    print(f"{name:40s} max {d.max().item():.4f} bad rows: {bad[:12]}{'...' if len(bad) > 12 else ''} (n={len(bad)})")
for env in [("0", "8"), ("1", "8"), ("0", "0"), ("0", "1e30")]:
    os.environ["CVLLM_PREFILL_DEBUG"], os.environ["CVLLM_PREFILL_THRESHOLD"] = env
    print("=== debug", env[0], "threshold", env[1])
    for Lc in [1, 2, 3, 8, 17, 70, 127, 129]:
        case(f"append 64, Lc {Lc}", 64, Lc)
    case("append 200 Lc 1", 200, 1)
    case("append 200 Lc 0", 200, 0)
    case("append 700 Lc 0 bf16", 700, 0, dtype=torch.bfloat16)
