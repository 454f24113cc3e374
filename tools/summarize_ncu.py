"""Turn gpurun_out/{launches.csv, prof_*.ncu-rep} into the tracked summaries under profiles/.

    python tools/summarize_ncu.py r01            # prefix for the output files
"""
import collections
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "gpurun_out"
PROF = ROOT / "profiles"
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.avg.per_second",
    "smsp__inst_executed.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "sm__cycles_elapsed.avg",
]


def launches(prefix: str):
    src = OUT / "launches.csv"
    if not src.exists():
        return
    lines = [l for l in src.read_text().splitlines(True) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for x in csv.DictReader(lines):
        agg.setdefault(x["Kernel Name"], []).append(float(x["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    ours = sum(sum(v) for k, v in agg.items() if "cvllm::" in k)
    with open(PROF / f"{prefix}_launches.md", "w") as f:
        f.write(f"# {prefix}: kernel launch list of `python bench.py --steps 2 --warmup 3 --no-extras --no-cpu`\n\n"
                "ncu `--metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare SHARES).\n"
                "Includes the workload's torch setup kernels (randn, copies) launched before the timed steps.\n\n"
                f"cvllm:: kernels are {100 * ours / tot:.1f} % of all device time in the capture.\n\n"
                "| kernel | launches | avg us | share of capture | share of cvllm time |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            name = k.split("(")[0][:80]
            mine = f"{100 * sum(v) / ours:.1f} %" if "cvllm::" in k else "-"
            f.write(f"| `{name}` | {len(v)} | {sum(v) / len(v) / 1e3:.1f} | {100 * sum(v) / tot:.1f} % | {mine} |\n")
    (PROF / f"{prefix}_launches.csv").write_text("".join(lines))


def capture(prefix: str, rep: str, tag: str, extra: dict):
    src = OUT / f"{rep}.ncu-rep"
    if not src.exists():
        return
    raw = subprocess.run(["ncu", "-i", str(src), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {}
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS or h == "Kernel Name":
            d[h] = {"value": v, "unit": u}
    def num(k):
        return float(d[k]["value"].replace(",", ""))
    def to_bytes(k):
        u = d[k]["unit"].lower()
        mul = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]
        return num(k) * mul
    summary = {"kernel": d.get("Kernel Name", {}).get("value"), "metrics": d,
               "dram_bytes_per_launch": to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")}
    summary.update(extra)
    (PROF / f"{prefix}_{tag}.json").write_text(json.dumps(summary, indent=1))


if __name__ == "__main__":
    prefix = sys.argv[1] if len(sys.argv) > 1 else "r01"
    PROF.mkdir(exist_ok=True)
    launches(prefix)
    capture(prefix, "prof_prefill", "prefill_traffic",
            {"command": "ncu --set full --clock-control none -k regex:prefill_attn -s 3 -c 1 python bench.py --steps 2 --warmup 3 --no-extras --no-cpu",
             "algorithmic_bytes": 32768 * (32 + 16) * 128 * 2 + 32768 * 32 * 128 * 2})
    capture(prefix, "prof_decode", "decode_traffic",
            {"command": "ncu --set full --clock-control none -k regex:decode_attn -s 4 -c 1 python tools/prof_decode.py "
                        "(B=16 x 8 heads x 16432 rows: the metric's 32K x 50 % shape)",
             "algorithmic_bytes": 16 * 8 * 16432 * 512 + 16 * 32 * 128 * 2 * 2})
    capture(prefix, "prof_rope", "rope_traffic",
            {"command": "ncu --set full --clock-control none -k regex:rope_qk -s 8 -c 1 python tools/probe_rope.py (plain rope, 32768 tokens)",
             "algorithmic_bytes": 32768 * 40 * 128 * 2 * 2})
    capture(prefix, "prof_rope_fused", "rope_fused_traffic",
            {"command": "ncu --set full --clock-control none -k regex:rope_qk -s 100 -c 1 python tools/probe_rope.py (norm + rope + sketch)",
             "algorithmic_bytes": 32768 * 40 * 128 * 2 * 2 + 32768 * 8 * 96})
