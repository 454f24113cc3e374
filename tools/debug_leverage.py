"""Per-chunk error of the device leverage scores against the exact-factorisation oracle."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
import oracle.scoring as osc
from compactor_vllm.compression.compactor import approximate_leverage_scores
from compactor_vllm.utils.arguments import make_phi
from test_gpu_scores import _structured_keys

def q(e, qs=(0.5, 0.99, 0.999)):
    e = e.flatten()
    return [round(e.kthvalue(max(1, int(x * e.numel()))).values.item(), 4) for x in qs] + [round(e.max().item(), 3)]

for dtype, scale in ((torch.bfloat16, 1.0), (torch.float16, 0.1)):
    for lens, H, chunk in [([1300, 700], 8, 512), ([100, 49, 600], 2, 512), ([512], 8, 512)]:
        N = sum(lens)
        keys = (_structured_keys(N, H, 128, torch.float32, seed=N + H) * scale).to(dtype)
        phi = make_phi(128, 48, dtype, "cpu")
        raw = approximate_leverage_scores(keys.cuda(), lens, phi.cuda(), normalize=False, chunk_size=chunk).cpu().float()
        ex = osc.approximate_leverage_scores(keys, lens, phi, chunk_size=chunk, factorization="eigh64").float()
        rf = osc.approximate_leverage_scores(keys, lens, phi, chunk_size=chunk, factorization="svd32").float()
        _, cl = osc.split_into_chunks(lens, chunk)
        off = 0
        print(os.environ.get("CVLLM_LIB", "new"), dtype, lens)
        for L in cl:
            sl = slice(off, off + L)
            den = ex[sl].abs().clamp_min(1)
            print(f"   chunk {L:4d}: mine {q((raw[sl]-ex[sl]).abs()/den)}  ref {q((rf[sl]-ex[sl]).abs()/den)}")
            off += L
