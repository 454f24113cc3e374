"""Does a pinned H2D copy make progress while a given kernel runs?  Times the copy alone and under each load."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]
import bench as B  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
w = B.Workload(dev, B.SEQ, seed=1)
for _ in range(3):
    w.prefill_step()
h = torch.empty(w.qkv.shape, dtype=w.qkv.dtype).pin_memory()
d = torch.empty_like(w.qkv)
ho = torch.empty(w.qkv.shape, dtype=w.qkv.dtype).pin_memory()
s_in = torch.cuda.Stream(device=dev)
a = torch.randn(8192, 8192, device=dev).bfloat16()
E = lambda: torch.cuda.Event(enable_timing=True)


def run(load, name, d2h=False):
    torch.cuda.synchronize()
    main = torch.cuda.current_stream()
    c0, c1, k0, k1 = E(), E(), E(), E()
    s_in.wait_stream(main)
    k0.record(main)
    if load:
        load()
    k1.record(main)
    with torch.cuda.stream(s_in):
        c0.record(s_in)
        if d2h:
            ho.copy_(d, non_blocking=True)
        else:
            d.copy_(h, non_blocking=True)
        c1.record(s_in)
    torch.cuda.synchronize()
    print(f"{name:40s} copy {c0.elapsed_time(c1):7.2f} ms   load {k0.elapsed_time(k1):7.2f} ms   copy end after load start {k0.elapsed_time(c1):7.2f}")


def mm():
    for _ in range(12):
        torch.mm(a, a)


for _ in range(2):
    run(None, "H2D alone")
    run(mm, "H2D under cuBLAS bf16 mm x12")
    run(lambda: w.prefill_step(compress=False), "H2D under prefill (no compression)")
    run(lambda: w.prefill_step(compress=True), "H2D under prefill + compression")
    run(None, "D2H alone", d2h=True)
    run(lambda: w.prefill_step(compress=False), "D2H under prefill (no compression)", d2h=True)

print("--- orderings ---")
d2 = torch.empty_like(w.qkv)
load = lambda: w.prefill_step(compress=True)


def order_probe(name, fn):
    torch.cuda.synchronize()
    main = torch.cuda.current_stream()
    s_in.wait_stream(main)
    ev = {k: E() for k in ("k0", "k1", "c0", "c1")}
    fn(main, ev)
    torch.cuda.synchronize()
    print(f"{name:50s} copyB {ev['c0'].elapsed_time(ev['c1']):7.2f} ms   load {ev['k0'].elapsed_time(ev['k1']):7.2f} ms   "
          f"copyB start-load start {ev['k0'].elapsed_time(ev['c0']):7.2f}")


def p1(main, ev):  # copy issued first, both runnable
    with torch.cuda.stream(s_in):
        ev["c0"].record(s_in); d.copy_(h, non_blocking=True); ev["c1"].record(s_in)
    ev["k0"].record(main); load(); ev["k1"].record(main)


def p3(main, ev):  # s_in: copyA, copyB; main: wait copyA, load  (the pipeline's shape)
    e_a = torch.cuda.Event()
    with torch.cuda.stream(s_in):
        d2.copy_(h, non_blocking=True); e_a.record(s_in)
        ev["c0"].record(s_in); d.copy_(h, non_blocking=True); ev["c1"].record(s_in)
    main.wait_event(e_a)
    ev["k0"].record(main); load(); ev["k1"].record(main)


def p4(main, ev):  # as p3, but copyB is issued from the host only after the load has been launched
    e_a = torch.cuda.Event()
    with torch.cuda.stream(s_in):
        d2.copy_(h, non_blocking=True); e_a.record(s_in)
    main.wait_event(e_a)
    ev["k0"].record(main); load(); ev["k1"].record(main)
    with torch.cuda.stream(s_in):
        ev["c0"].record(s_in); d.copy_(h, non_blocking=True); ev["c1"].record(s_in)


def p5(main, ev):  # as p3 with the load = cuBLAS
    e_a = torch.cuda.Event()
    with torch.cuda.stream(s_in):
        d2.copy_(h, non_blocking=True); e_a.record(s_in)
        ev["c0"].record(s_in); d.copy_(h, non_blocking=True); ev["c1"].record(s_in)
    main.wait_event(e_a)
    ev["k0"].record(main); mm(); ev["k1"].record(main)


def p6(main, ev):  # as p3 without compression (no store stream)
    e_a = torch.cuda.Event()
    with torch.cuda.stream(s_in):
        d2.copy_(h, non_blocking=True); e_a.record(s_in)
        ev["c0"].record(s_in); d.copy_(h, non_blocking=True); ev["c1"].record(s_in)
    main.wait_event(e_a)
    ev["k0"].record(main); w.prefill_step(compress=False); ev["k1"].record(main)


for _ in range(2):
    order_probe("P1 copy first, load second", p1)
    order_probe("P3 copyA;copyB | wait A; load", p3)
    order_probe("P4 copyA | wait A; load ; then issue copyB", p4)
    order_probe("P5 = P3 with cuBLAS load", p5)
    order_probe("P6 = P3 without compression", p6)
