"""A handful of decode steps (B=16 ragged, mean 10240 rows/head: BASELINE config 3's shape) for ncu."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
g = torch.Generator().manual_seed(3)
lens16 = ((torch.rand((16, bench.HKV), generator=g) + 0.5) * 10240).long()
ms, nbytes = bench.decode_bench(dev, torch.bfloat16, 16, lens16, 5, flush, 6)
print(f"decode B=16: {ms * 1e3:.1f} us, {nbytes / ms / 1e6:.1f} GB/s algorithmic")
