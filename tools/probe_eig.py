"""Phase clocks of the fp64 eigensolver (cvllm_selftest_eig, clk_out): where a 48x48 problem spends its time."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests")]
import torch
from compactor_vllm import _native as N
from test_gpu_scores import _gram_like

dev = "cuda"
for mats in (1, 148, 512, 2048):
    G = _gram_like(min(mats, 64), 12, 0.05, 0)
    G = G.repeat((mats + G.shape[0] - 1) // G.shape[0], 1, 1)[:mats].contiguous().to(dev)
    lam = torch.empty((mats, 48), dtype=torch.float64, device=dev)
    vec = torch.empty((mats, 48, 48), dtype=torch.float64, device=dev)
    sv = torch.empty((mats, 48, 48), dtype=torch.float32, device=dev)
    clk = torch.zeros((mats, 6), dtype=torch.int64, device=dev)
    f = lambda c: N.check(N.lib().cvllm_selftest_eig(N.ptr(G), N.ptr(lam), N.ptr(vec), N.ptr(sv), mats, N.CVLLM_BF16,
                                                      N.ptr(c) if c is not None else None, N.stream()), "eig")
    for _ in range(3):
        f(clk)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        f(None)
    b.record()
    b.synchronize()
    c = clk.cpu().double()
    t0 = c[:, 0]
    med = lambda x: float(x.median())
    print(f"mats={mats}: kernel {a.elapsed_time(b) * 100:.1f} us | cycles (median over CTAs): tridiag {med(c[:,1]-t0):.0f}, "
          f"QL producer end {med(c[:,2]-t0):.0f}, accumulate end {med(c[:,3]-t0):.0f}, exit {med(c[:,4]-t0):.0f}, "
          f"ring entries {med(c[:,5]):.0f}", flush=True)
