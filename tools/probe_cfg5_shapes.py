"""Config-5 per-rank shapes on ONE GPU: step time with / without compression for (HQ, HKV) = (64,8) (32,4) (16,2) (8,1)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch
import bench
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
for hq, hkv in ((64, 8), (32, 4), (16, 2), (8, 1)):
    w = bench.Workload(dev, bench.SEQ, seed=5, hq=hq, hkv=hkv)
    for _ in range(3):
        w.prefill_step(); w.prefill_step(compress=False); w.compress_only()
    torch.cuda.synchronize()
    a = bench.time_batch(lambda: w.prefill_step(compress=False), 5)
    b = bench.time_batch(lambda: w.prefill_step(compress=True), 5)
    c = bench.time_loop(w.compress_only, 5)
    fl = bench.prefill_flops(bench.SEQ, hq=hq)
    print(f"HQ={hq} HKV={hkv}: plain {a:.2f} ms ({fl/a/1e9:.0f} TFLOP/s), compressed {b:.2f} ms, chain alone {c*1e3:.0f} us", flush=True)
    del w
    torch.cuda.empty_cache()
