"""COMPACTOR compression chain (32K and 4K tokens) with and without programmatic dependent launch, eager and as a CUDA
graph, alternating.  `pdl` is the library's test option (off by default: launch_pdl then passes attribute value 0)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from compactor_vllm import _native as N  # noqa: E402


def opt(v):
    N.check(N.lib().cvllm_set_test_option(b"pdl", float(1 - v)), "pdl")


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    for n_tok in (bench.SEQ, 4096):
        w = bench.Workload(dev, n_tok, seed=1234)
        for rnd in range(2):
            for no_pdl in (1, 0):
                opt(no_pdl)
                eager = timed(w.compress_only, 20)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    w.compress_only()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        w.compress_only()
                torch.cuda.current_stream().wait_stream(side)
                graph = timed(g.replay, 20)
                print(f"{n_tok} tokens, {'plain launches' if no_pdl else 'PDL           '}: eager {eager:7.1f} us, graph {graph:7.1f} us",
                      flush=True)
                del g
        del w
    opt(0)


if __name__ == "__main__":
    main()
