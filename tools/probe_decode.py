"""Decode attention alone under CUDA-graph replay (L2 flushed), B=16 ragged and B=1 at 131K pairs."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
from ref_compare import Timer, _paged, HQ, HKV, D, PS
from compactor_vllm.attention.sparse_decode_kernel import head_sparse_decode_attention

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
T = Timer(torch)
for name, B, mean in (("b16", 16, 10240), ("b1", 1, 16400), ("b64", 64, 4096), ("b4", 4, 32768)):
    gd = torch.Generator().manual_seed(11 + B)
    lens_bh = (torch.randint(mean // 2, mean * 3 // 2 + 1, (B, HKV), generator=gd) if B > 1
               else torch.full((1, HKV), mean)).to(torch.int32)
    kc, vc, pt, bm = _paged(torch, B, int(lens_bh.max()) + 64, 12 + B, dev)
    ld = lens_bh.to(dev)
    q = torch.randn((B, HQ, D), device=dev, dtype=torch.bfloat16)
    f = lambda: head_sparse_decode_attention(q, kc, vc, ld, pt, bm, HKV, PS)
    f()
    nbytes = int(lens_bh.sum()) * D * 2 * 2 + B * HQ * D * 2 * 2
    for fm in ("write", "write+read", "none"):
        T.flush_mode = fm
        us = T.graph(f, iters=30) * 1e3
        print(f"decode {name} [flush={fm}]: {us:.1f} us  {nbytes / us / 1e3:.0f} GB/s  ({nbytes / 1e6:.0f} MB)", flush=True)
    T.flush_mode = "write"
    del kc, vc

# reference points: plain streaming read and copy of 1 GiB through torch
x = torch.empty(1 << 29, dtype=torch.bfloat16, device=dev).normal_()
y = torch.empty_like(x)
for nm, fn, nb in (("torch sum (read 1 GiB)", lambda: x.sum(), x.numel() * 2),
                   ("torch copy (r+w 2 GiB)", lambda: y.copy_(x), x.numel() * 4)):
    us = T(fn, iters=10) * 1e3
    print(f"{nm}: {us:.1f} us  {nb / us / 1e3:.0f} GB/s", flush=True)
