"""Prefill attention alone at 32K (and 8K): ms and TFLOP/s; CVLLM_LIB selects a library variant."""
import math, os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
from ref_compare import Timer, _inputs, _paged, HQ, HKV, D, PS
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
T = Timer(torch); T.flush_mode = "none"
for N in (32768, 8192):
    q, k, v, _ = _inputs(torch, N, 0, dev)
    kc, vc, pt, bm = _paged(torch, 1, N + PS, 1, dev, fill=False)
    zl = torch.zeros((1, HKV), dtype=torch.int32, device=dev)
    cu = torch.tensor([0, N], dtype=torch.int32, device=dev)
    f = lambda: causal_sparse_varlen_with_cache(q, k, v, kc, vc, zl, pt, bm, cu, N, 0, HKV, PS, 1 / math.sqrt(D))
    ms = T(f, iters=10, warm=3)
    fl = 4.0 * D * HQ * (N * (N + 1) / 2)
    print(f"{os.environ.get('CVLLM_LIB', 'default').split('/')[-1]} N={N}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
