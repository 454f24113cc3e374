"""Decode step (fused append + attention, CUDA-graph replay, cold clean L2) at B = 1 on the headline's compacted shape
(8 heads x 16344 rows = 67 MB) and at B = 16 ragged; for A/B runs (CVLLM_LIB=...) and ncu launch lists."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
import bench as B_

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
lens = torch.full((1, B_.HKV), 16344, dtype=torch.int32)
ms, nbytes = B_.decode_bench(dev, torch.bfloat16, 1, lens, iters, flush, seed=3)
out = f"B=1 {ms * 1e3:.1f} us {nbytes / ms / 1e6:.0f} GB/s"
if len(sys.argv) > 2:
    gd = torch.Generator().manual_seed(27)
    l16 = torch.randint(5120, 15361, (16, B_.HKV), generator=gd).to(torch.int32)
    ms, nbytes = B_.decode_bench(dev, torch.bfloat16, 16, l16, iters, flush, seed=4)
    out += f" | B=16 {ms * 1e3:.1f} us {nbytes / ms / 1e6:.0f} GB/s"
print(out)
