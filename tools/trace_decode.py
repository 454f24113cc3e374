"""Per-CTA timeline of the decode step (cvllm_debug_decode_trace): where the microseconds of a B = 1 / B = 16 launch go.
Needs the trace variant of the library:
    bash tools/build_variant.sh trace decode_attn.cu -DCVLLM_DECODE_TRACE -Icsrc -I../include
    CVLLM_LIB=$PWD/compactor-vllm_b200/lib/variants/lib_trace.so python tools/trace_decode.py"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch
import bench as B_
from compactor_vllm import _native as N
from compactor_vllm.attention.sparse_decode_kernel import decode_append_and_attend

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
HKV, HQ, D, PS = B_.HKV, B_.HQ, B_.D, B_.PS
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
trace = torch.zeros(8 * n_sm, dtype=torch.int64, device=dev)


def run(B, lens_bh):
    g = torch.Generator().manual_seed(5)
    nlp = int((int(lens_bh.max()) + 64 + PS - 1) // PS)
    n_pages = (B + 1) * HKV * nlp
    pt = torch.randperm(n_pages, generator=g).to(torch.int32).view(B + 1, HKV, nlp).to(dev)
    kc = torch.randn((n_pages * PS, D), device=dev, dtype=torch.bfloat16)
    vc = torch.randn((n_pages * PS, D), device=dev, dtype=torch.bfloat16)
    bm = torch.arange(1, B + 1, dtype=torch.int32, device=dev)
    lens = torch.zeros((B + 1, HKV), dtype=torch.int32, device=dev)
    lens[1:] = lens_bh.to(dev)
    q = torch.randn((B, HQ, D), device=dev, dtype=torch.bfloat16)
    kn = torch.randn((B, HKV, D), device=dev, dtype=torch.bfloat16)
    vn = torch.randn((B, HKV, D), device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        decode_append_and_attend(q, kn, vn, kc, vc, lens, pt, bm, HKV, PS)
    rows = []
    for it in range(5):
        flush.zero_(); flush.view(torch.float32).sum()
        trace.zero_()
        assert N.lib().cvllm_debug_decode_trace(N.ptr(trace)) == 0, "library built without -DCVLLM_DECODE_TRACE"
        decode_append_and_attend(q, kn, vn, kc, vc, lens, pt, bm, HKV, PS)
        N.lib().cvllm_debug_decode_trace(None)
        torch.cuda.synchronize()
        t = trace.view(n_sm, 8).cpu().double()
        t0 = t[:, 0].min()
        rel = (t[:, :5] - t0) / 1e3
        rows.append(rel)
        role = t[:, 5]
    rel = rows[-1]
    names = ["entry", "lens read", "first unit landed", "streaming done", "merge done"]
    print(f"B={B}: {int(lens_bh.sum()) * 512 / 1e6:.0f} MB")
    for i, nme in enumerate(names):
        col = rel[:, i]
        print(f"  {nme:20s} min {col.min():7.2f}  median {col.median():7.2f}  max {col.max():7.2f} us")
    m = role == 2
    if m.any():
        full = (trace.view(n_sm, 8).cpu().double() - t0) / 1e3
        for c in torch.nonzero(m).flatten().tolist():
            print(f"    merging CTA {c}: stream done {full[c, 3]:.2f}  ticket {full[c, 6]:.2f}  slots folded {full[c, 7]:.2f}  done {full[c, 4]:.2f}")
        others = ~m & (role == 1)
        print(f"    partial CTAs: ticket - stream done median {(full[others, 6] - full[others, 3]).median():.2f} us")
        print(f"  merging CTAs: {int(m.sum())}, their merge-done median {rel[m, 4].median():.2f} max {rel[m, 4].max():.2f}; "
              f"non-merging done max {rel[~m, 4].max():.2f}")


run(1, torch.full((1, HKV), 16344, dtype=torch.int32))
gd = torch.Generator().manual_seed(27)
run(16, torch.randint(5120, 15361, (16, HKV), generator=gd).to(torch.int32))
