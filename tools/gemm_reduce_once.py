"""Three launches of csrc/gemm_reduce.cu at world 1 on the config-5 tp = 4 shard shape (ncu target)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "compactor-vllm_b200"):
    sys.path.insert(0, str(p))
from compactor_vllm.layers.linear import PeerGemmAllReduce  # noqa: E402

dev = torch.device("cuda", 0)
m, n, k = 32768, 8192, int(sys.argv[1]) if len(sys.argv) > 1 else 2048
(pg,), free = PeerGemmAllReduce.virtual_ranks(1, m, n, dev)
x = torch.randn((m, k), device=dev).to(torch.bfloat16)
w = torch.randn((n, k), device=dev).to(torch.bfloat16)
for _ in range(3):
    pg.linear(x, w)
torch.cuda.synchronize()
free()
print("ok")
