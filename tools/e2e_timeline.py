"""Timeline of the double-buffered end-to-end loop of bench.py: per step, when its H2D, kernels and D2H start and end
(CUDA events against a common origin).  Diagnostic for the copy / compute overlap."""
import os
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import bench as B  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n_tok = B.SEQ
    w = B.Workload(dev, n_tok, seed=1234)
    for _ in range(3):
        w.prefill_step()
    h_qkv = torch.empty(w.qkv.shape, dtype=w.qkv.dtype).pin_memory()
    h_kpre = torch.empty(w.k_pre.shape, dtype=w.k_pre.dtype).pin_memory()
    h_qkv.copy_(w.qkv)
    h_kpre.copy_(w.k_pre)
    d_in = [(torch.empty_like(w.qkv), torch.empty_like(w.k_pre)) for _ in range(2)]
    h_outs = [torch.empty((n_tok, B.HQ, B.D), dtype=w.dtype).pin_memory() for _ in range(2)]
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    E = lambda: torch.cuda.Event(enable_timing=True)
    ev_in, ev_done, ev_out = [E(), E()], [E(), E()], [E(), E()]
    main_s = torch.cuda.current_stream()
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    marks = []
    for rep in range(2):
        marks.clear()
        torch.cuda.synchronize()
        origin = E()
        origin.record(main_s)
        for ev in ev_done + ev_out:
            ev.record(main_s)
        s_in.wait_stream(main_s)
        s_out.wait_stream(main_s)
        t0 = time.perf_counter()
        for i in range(n):
            b = i & 1
            dq, dk = d_in[b]
            m = {"host_issue_ms": (time.perf_counter() - t0) * 1e3}
            with torch.cuda.stream(s_in):
                s_in.wait_event(ev_done[b])
                m["h2d0"] = E(); m["h2d0"].record(s_in)
                dk.copy_(h_kpre, non_blocking=True)
                dq.copy_(h_qkv, non_blocking=True)
                ev_in[b] = E(); ev_in[b].record(s_in)
                m["h2d1"] = ev_in[b]
            main_s.wait_event(ev_in[b])
            m["c0"] = E(); m["c0"].record(main_s)
            q, k, v = B.split_qkv(dq)
            out = w.prefill_step(q, k, v, dk) if not os.environ.get('SKIP_COMPUTE') else w.q
            ev_done[b] = E(); ev_done[b].record(main_s)
            m["c1"] = ev_done[b]
            m["host_enq_ms"] = (time.perf_counter() - t0) * 1e3
            ev_out[b].synchronize()
            m["host_sync_ms"] = (time.perf_counter() - t0) * 1e3
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_done[b])
                m["d2h0"] = E(); m["d2h0"].record(s_out)
                if not os.environ.get('SKIP_D2H'):
                    h_outs[b].copy_(out, non_blocking=True)
                    out.record_stream(s_out)
                ev_out[b] = E(); ev_out[b].record(s_out)
                m["d2h1"] = ev_out[b]
            marks.append(m)
        torch.cuda.synchronize()
        total = (time.perf_counter() - t0) * 1e3
    print(f"total {total:.2f} ms for {n} steps = {total / n:.2f} ms/step")
    for i, m in enumerate(marks):
        t = {k: origin.elapsed_time(v) for k, v in m.items() if not k.startswith("host")}
        print(f"step {i}: host issue {m['host_issue_ms']:.2f} enq {m['host_enq_ms']:.2f} sync {m['host_sync_ms']:.2f} | "
              f"h2d {t['h2d0']:.2f}-{t['h2d1']:.2f}  compute {t['c0']:.2f}-{t['c1']:.2f}  d2h {t['d2h0']:.2f}-{t['d2h1']:.2f}")


if __name__ == "__main__":
    main()
