"""Inputs for tools/ubench/sketch_order.cu: the reference's X = matmul(K^T-view, PHI) for the structured bench keys."""
import math, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent.parent
sys.path[:0] = [str(ROOT / "tools")]
import torch
from leverage_selfdistance import _keys
dev = torch.device("cuda", 0)
D = 128
phi = (torch.randn((D, 48), generator=torch.Generator(device=dev).manual_seed(42), device=dev).to(torch.bfloat16) / math.sqrt(48))
q, k, v, k_pre = _keys(torch, "structured", 32768, dev)
X = torch.matmul(k_pre.transpose(0, 1), phi)  # [H, N, 48], the reference's call (compactor.py:158)
h = 3
k_pre[:, h].contiguous().cpu().view(torch.int16).numpy().tofile("/tmp/sk_k.bin")
phi.cpu().view(torch.int16).numpy().tofile("/tmp/sk_phi.bin")
X[h].contiguous().cpu().view(torch.int16).numpy().tofile("/tmp/sk_x.bin")
print("wrote", X.shape)
# which kernel does cuBLAS run for this GEMM?  (name hints at the instruction family: nvjet / sm100 = tcgen05, sm80/sm90 = mma.sync / wgmma)
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    torch.matmul(k_pre.transpose(0, 1), phi)
    ch = X.view(8, -1, 512, 48).clone()
    torch.matmul(ch.transpose(-1, -2), ch)
    torch.cuda.synchronize()
for e in prof.key_averages():
    print("kernel:", e.key[:160])
