"""Where the time of csrc/gemm_reduce.cu goes on ONE GPU (world 1): timing probes through the `grs_probe` test option
(1 no TMA loads, 2 no UMMA, 4 no epilogue stores / reduce, 8 no reduce, 16 no L2 hints).  Probe results are wrong by construction."""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "compactor-vllm_b200"):
    sys.path.insert(0, str(p))
from compactor_vllm import _native as N  # noqa: E402
from compactor_vllm.layers.linear import PeerGemmAllReduce  # noqa: E402


def opt(name, v):
    N.check(N.lib().cvllm_set_test_option(name.encode(), float(v)), name)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


DBG = None


def report_failure():
    rec = DBG.view(-1, 8, 8)
    print("timeout records (cta: [code thread a b seen] ...):", file=sys.stderr)
    for cta in range(rec.shape[0]):
        rows = [rec[cta, c, :5].tolist() for c in range(8) if int(rec[cta, c, 0]) != 0]
        if rows:
            print("  ", cta, rows, file=sys.stderr)


def main():
    global DBG
    dev = torch.device("cuda", 0)
    DBG = torch.zeros(64 * 160, dtype=torch.int32).pin_memory()
    opt("grs_debug_host_ptr", DBG.data_ptr())
    m, n = 32768, 8192
    out = []
    ks = [int(a) for a in sys.argv[1:]] or [1024, 4096]
    for rpr in (0, 2048):
        opt("grs_rows_per_round", rpr)
        (pg,), free = PeerGemmAllReduce.virtual_ranks(1, m, n, dev)
        for k in ks:
            x = torch.randn((m, k), device=dev).to(torch.bfloat16)
            w = torch.randn((n, k), device=dev).to(torch.bfloat16)
            row = {"k": k, "rows_per_round": rpr or 256}
            for name, probe in (("full", 0), ("full_nohint", 16), ("no_reduce", 8), ("gemm_only", 4),
                                ("gemm_only_nohint", 20), ("no_tma", 5), ("no_mma", 6), ("skeleton", 7)):
                opt("grs_probe", probe)
                print("probe start", k, name, file=sys.stderr, flush=True)
                row[name + "_ms"] = round(timed(lambda: pg.linear(x, w)), 4)
                print("probe", k, name, row[name + "_ms"], file=sys.stderr, flush=True)
            opt("grs_probe", 0)
            row["gemm_only_tflops"] = round(2.0 * m * n * k / row["gemm_only_ms"] / 1e9, 1)
            out.append(row)
            print(json.dumps(row), flush=True)
            del x, w
        free()
    opt("grs_rows_per_round", 0)


if __name__ == "__main__":
    try:
        main()
    except Exception as ex:  # noqa: BLE001
        print("FAILED:", repr(ex)[:200], file=sys.stderr)
        report_failure()
        sys.exit(1)
