"""Where does a short prompt's compressed step spend its time?  For 4096 / 16384 / 32768 tokens: the compression chain
alone, the step without and with compression (synchronised per step, and back to back), the per-kernel CUDA-event
split of the chain, and the traceback of a CUDA-graph capture of the whole step."""
import sys
import traceback
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    for n in (4096, 16384, 32768):
        w = bench.Workload(dev, n, seed=4321)
        for _ in range(3):
            w.prefill_step()
            w.prefill_step(compress=False)
            w.compress_only()
        torch.cuda.synchronize()
        chain = bench.time_loop(w.compress_only, 10)
        plain = bench.time_loop(lambda: w.prefill_step(compress=False), 10)
        comp = bench.time_loop(lambda: w.prefill_step(compress=True), 10)
        plain_b = bench.time_batch(lambda: w.prefill_step(compress=False), 10)
        comp_b = bench.time_batch(lambda: w.prefill_step(compress=True), 10)
        print(f"n={n}: chain alone {chain*1e3:.0f} us | step sync'd: plain {plain*1e3:.0f} us, compressed {comp*1e3:.0f} us"
              f" | back to back: plain {plain_b*1e3:.0f} us, compressed {comp_b*1e3:.0f} us", flush=True)
        if n == 16384:
            # the same step with a default-priority store stream
            w.store_stream = torch.cuda.Stream(device=dev)
            for _ in range(2):
                w.prefill_step()
            print(f"   default-priority store stream: compressed {bench.time_loop(lambda: w.prefill_step(compress=True), 10)*1e3:.0f} us",
                  flush=True)
            try:
                gs = torch.cuda.Stream(device=dev)
                gs.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(gs):
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=gs):
                        w.prefill_step(compress=True)
                print("   graph capture of the compressed step: ok", flush=True)
            except Exception:
                traceback.print_exc()
            torch.cuda.synchronize()
        del w
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
