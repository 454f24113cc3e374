"""A handful of decode launches for ncu: B=16 at the metric's shape (16432 rows per head = 32K context, 50 % keep;
default) or `cfg3` = B=16 ragged, mean 10240 rows/head (BASELINE config 3)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
from ref_compare import _paged, HQ, HKV, D, PS
from compactor_vllm.attention.sparse_decode_kernel import head_sparse_decode_attention

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
B, mean = 16, 10240
gd = torch.Generator().manual_seed(11 + B)
if len(sys.argv) > 1 and sys.argv[1] == "cfg3":
    lens_bh = torch.randint(mean // 2, mean * 3 // 2 + 1, (B, HKV), generator=gd).to(torch.int32)
else:
    lens_bh = torch.full((B, HKV), 16432, dtype=torch.int32)
kc, vc, pt, bm = _paged(torch, B, int(lens_bh.max()) + 64, 12 + B, dev)
ld = lens_bh.to(dev)
q = torch.randn((B, HQ, D), device=dev, dtype=torch.bfloat16)
for _ in range(6):
    out = head_sparse_decode_attention(q, kc, vc, ld, pt, bm, HKV, PS)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
