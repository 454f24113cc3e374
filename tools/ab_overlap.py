"""Step with / without compression in alternation (32K), with the selection chain launched with and without programmatic
dependent launch: how much of the compression chain does the attention kernel hide?"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from compactor_vllm import _native as N  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
w = bench.Workload(dev, bench.SEQ, seed=1234)
for _ in range(3):
    w.prefill_step()
    w.prefill_step(compress=False)
for rnd in range(2):
    for no_pdl in (1, 0):
        N.check(N.lib().cvllm_set_test_option(b"pdl", float(1 - no_pdl)), "pdl")
        tw = to = 0.0
        n = 8
        for _ in range(n):
            tw += bench.time_loop(lambda: w.prefill_step(), 1)
            to += bench.time_loop(lambda: w.prefill_step(compress=False), 1)
        print(f"{'plain launches' if no_pdl else 'PDL           '}: with {tw / n:.3f} ms, without {to / n:.3f} ms, added {100 * (tw - to) / to:.1f} %",
              flush=True)
N.check(N.lib().cvllm_set_test_option(b"pdl", 0.0), "pdl")
