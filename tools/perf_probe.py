"""Quick device timing of the two attention kernels (CUDA events, L2 flushed between iterations)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
from compactor_vllm.attention.sparse_decode_kernel import head_sparse_decode_attention
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
DEV = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)

def timeit(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]

def prefill(N, HQ=32, HKV=8, D=128, PS=128, dtype=torch.bfloat16):
    q = torch.randn(N, HQ, D, device=DEV, dtype=dtype); k = torch.randn(N, HKV, D, device=DEV, dtype=dtype); v = torch.randn_like(k)
    kc = torch.zeros(PS * 8, D, device=DEV, dtype=dtype); vc = torch.zeros_like(kc)
    pt = torch.zeros(2, HKV, 1, dtype=torch.int32, device=DEV); bm = torch.ones(1, dtype=torch.int32, device=DEV)
    lens = torch.zeros(1, HKV, dtype=torch.int32, device=DEV); cu = torch.tensor([0, N], dtype=torch.int32, device=DEV)
    ms = timeit(lambda: causal_sparse_varlen_with_cache(q, k, v, kc, vc, lens, pt, bm, cu, N, 0, HKV, PS), iters=5)
    flops = 4 * D * HQ * N * (N + 1) / 2
    print(f"prefill N={N}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s")

def decode(B, L, HQ=32, HKV=8, D=128, PS=128, dtype=torch.bfloat16, ragged=False):
    g = torch.Generator().manual_seed(0)
    lens = torch.full((B, HKV), L, dtype=torch.int64)
    if ragged:
        lens = (torch.rand((B, HKV), generator=g) * 1.5 * L + 0.25 * L).long()
    NLP = int((lens.max() + PS - 1) // PS)
    n_pages = (B + 1) * HKV * NLP
    pt = torch.randperm(n_pages, generator=g).to(torch.int32).view(B + 1, HKV, NLP).to(DEV)
    kc = torch.randn(n_pages * PS, D, device=DEV, dtype=dtype); vc = torch.randn_like(kc)
    bm = torch.arange(1, B + 1, dtype=torch.int32, device=DEV)
    q = torch.randn(B, HQ, D, device=DEV, dtype=dtype)
    l = lens.to(torch.int32).to(DEV)
    ms = timeit(lambda: head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS), iters=20)
    nbytes = int(lens.sum()) * D * 2 * 2
    print(f"decode B={B} L={L} ragged={ragged}: {ms * 1e3:.1f} us  {nbytes / ms / 1e6:.1f} GB/s  ({nbytes / 1e6:.1f} MB)")

for N in (4096, 16384, 32768):
    prefill(N)
decode(1, 8192); decode(1, 16384); decode(16, 10240); decode(16, 10240, ragged=True); decode(64, 4096, ragged=True)
