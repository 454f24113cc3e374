"""CTA-pair prefill kernel (test option prefill_v3) against the cta_group::1 kernel: max |diff| on a ladder of shapes
(single tile, odd tile counts, ragged prefixes, several sequences), then timing at 8K / 32K."""
import math, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests"), str(ROOT / "tools")]
import torch
from compactor_vllm import _native as N
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from helpers import fused_qkv, make_paged

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
HQ, HKV, D, PS = 32, 8, 128, 128
OPT = sys.argv[1] if len(sys.argv) > 1 else "prefill_v3"


def run(v3, q, k, v, kc, vc, lens, pt, bm, cu):
    N.set_test_option(OPT, 1.0 if v3 else 0.0)
    seqlens = cu[1:] - cu[:-1]
    out = causal_sparse_varlen_with_cache(q, k, v, kc, vc, lens, pt, bm, cu, int(seqlens.max()), int(lens.max()), HKV, PS,
                                          None)
    torch.cuda.synchronize()
    return out


cases = [([100], 0), ([128], 0), ([129], 0), ([256], 0), ([300], 0), ([700], 0), ([300], 70), ([300], 128), ([513], 200),
         ([40, 700, 129], 333), ([2048], 1000), ([1, 1, 5], 5)]
ok = True
for lens_q, cache_len in cases:
    B = len(lens_q)
    g = torch.Generator().manual_seed(7)
    lens_bh = [[max(0, cache_len - 13 * h) for h in range(HKV)] for _ in range(B)] if cache_len else [[0] * HKV] * B
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens_bh, torch.bfloat16, dev, seed=3, poison=True)
    cu = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32, device=dev)
    q, k, v = fused_qkv(sum(lens_q), HQ, HKV, D, torch.bfloat16, dev, seed=5)
    lens = torch.tensor(lens_bh, dtype=torch.int32, device=dev)
    a = run(False, q, k, v, kc, vc, lens, pt, bm, cu)
    b = run(True, q, k, v, kc, vc, lens, pt, bm, cu)
    d = (a.float() - b.float()).abs().max().item()
    fin = bool(torch.isfinite(b.float()).all())
    print(f"lens {lens_q} prefix {cache_len}: max |v3 - v2| = {d:.3e} finite={fin}", flush=True)
    ok &= d < 2e-2 and fin
print("parity", "OK" if ok else "FAILED", flush=True)

from ref_compare import Timer, _inputs, _paged
T = Timer(torch); T.flush_mode = "none"
for Ntok in (8192, 32768):
    q, k, v, _ = _inputs(torch, Ntok, 0, dev)
    kc, vc, pt, bm = _paged(torch, 1, Ntok + PS, 1, dev, fill=False)
    zl = torch.zeros((1, HKV), dtype=torch.int32, device=dev)
    cu = torch.tensor([0, Ntok], dtype=torch.int32, device=dev)
    f = lambda: causal_sparse_varlen_with_cache(q, k, v, kc, vc, zl, pt, bm, cu, Ntok, 0, HKV, PS, 1 / math.sqrt(D))
    fl = 4.0 * D * HQ * (Ntok * (Ntok + 1) / 2)
    for rep in range(2):
        for v3 in (0, 1):
            N.set_test_option(OPT, float(v3))
            ms = T(f, iters=10, warm=3)
            print(f"N={Ntok} {OPT if v3 else 'prefill2'}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
