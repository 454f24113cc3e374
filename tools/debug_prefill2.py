import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from conftest import load_golden
DEV = "cuda"
for name in ["prefill_bf16", "prefill_f16_ps256"]:
    g = load_golden(name)
    N, HQ, D = g["q"].shape; HKV = g["k"].shape[1]
    qkv = torch.cat([g["q"].reshape(N, -1), g["k"].reshape(N, -1), g["v"].reshape(N, -1)], 1).to(DEV)
    q = qkv[:, : HQ * D].view(N, HQ, D); k = qkv[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D); v = qkv[:, (HQ + HKV) * D:].view(N, HKV, D)
    cu = g["cu_seqlens_q"].to(DEV); lens = g["seq_lens_bh"].to(DEV)
    for rep in range(3):
        out = causal_sparse_varlen_with_cache(q, k, v, g["k_cache"].to(DEV), g["v_cache"].to(DEV), lens, g["page_table"].to(DEV),
                                              g["batch_mapping"].to(DEV), cu, int((cu[1:] - cu[:-1]).max()), int(lens.max()), g["HKV"], g["PS"])
        torch.cuda.synchronize()
        d = (out.float().cpu() - g["out"].float()).abs()
        print(name, "rep", rep, "max", d.max().item(), "cu", cu.tolist(), "lens", lens.tolist())
        tok = d.amax(dim=(1, 2))
        bad = (tok > 0.02).nonzero().flatten().tolist()
        print("   bad tokens:", bad[:40])
        for t in bad[:6]:
            print("   tok", t, "per-head", [round(x, 3) for x in d[t].amax(dim=1).tolist()])
