import math, sys
from pathlib import Path
ROOT = Path("/root/repo")
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests"), str(ROOT / "tools")]
import torch
from compactor_vllm import _native as N
from compactor_vllm.attention.sparse_varlen_kernel import causal_sparse_varlen_with_cache
from ref_compare import Timer, _inputs, _paged
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
HQ, HKV, D, PS = 32, 8, 128, 128
T = Timer(torch); T.flush_mode = "none"
Ntok = 32768
q, k, v, _ = _inputs(torch, Ntok, 0, dev)
kc, vc, pt, bm = _paged(torch, 1, Ntok + PS, 1, dev, fill=False)
zl = torch.zeros((1, HKV), dtype=torch.int32, device=dev)
cu = torch.tensor([0, Ntok], dtype=torch.int32, device=dev)
f = lambda: causal_sparse_varlen_with_cache(q, k, v, kc, vc, zl, pt, bm, cu, Ntok, 0, HKV, PS, 1 / math.sqrt(D))
for v3 in (0, 1, 0, 1):
    N.set_test_option("prefill_v3", float(v3))
    print(sys.argv[1], "v3" if v3 else "v2", f"{T(f, iters=10, warm=3):.3f} ms", flush=True)
