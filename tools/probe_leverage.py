"""Leverage-score chain at the headline shape: timing + output dump (compare CVLLM_EIG=jacobi vs default QL)."""
import sys, os
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
from ref_compare import Timer, _inputs, HQ, HKV, D
from compactor_vllm.compression.compactor import approximate_leverage_scores
import math
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
N = 32768
q, k, v, k_pre = _inputs(torch, N, 0, dev)
phi = (torch.randn((D, 48), generator=torch.Generator(device=dev).manual_seed(42), device=dev).to(torch.bfloat16) / math.sqrt(48))
f = lambda: approximate_leverage_scores(k_pre, [N], phi, normalize=True, chunk_size=512)
T = Timer(torch)
T.flush_mode = "none"
out = f()
ms = T.graph(f, iters=20)
tag = os.environ.get("CVLLM_EIG", "ql")
print(f"leverage chain [{tag}]: {ms*1e3:.1f} us", flush=True)
torch.save(out.float().cpu(), f"/tmp/lev_{tag}.pt")
if len(sys.argv) > 1:
    a = torch.load(f"/tmp/lev_{sys.argv[1]}.pt"); b = out.float().cpu()
    d = (a - b).abs() / a.abs().clamp_min(1.0)
    print("vs", sys.argv[1], "max", float(d.max()), "median", float(d.median()), "frac<=2e-2", float((d <= 2e-2).float().mean()),
          "nan", int(torch.isnan(b).sum()), int(torch.isnan(a).sum()))
