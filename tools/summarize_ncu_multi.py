"""Per-kernel summary of one multi-kernel `ncu --set full` report -> profiles/<prefix>_<tag>.json

    python tools/summarize_ncu_multi.py gpurun_out/prof_scoring.ncu-rep r02 scoring_kernels
One entry per distinct kernel name (the LAST launch of each: warm caches for code, same data), with duration, DRAM
bytes, achieved DRAM GB/s, tensor / xu / fma pipe utilisation, issue-slot utilisation, occupancy, registers.
"""
import csv
import json
import subprocess
import sys
from pathlib import Path

KEYS = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": "tensor_pipe_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed": "xu_pipe_pct",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active": "fp64_pipe_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slot_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__shared_mem_per_block_dynamic": "dyn_smem",
    "lts__t_bytes.sum": "l2_bytes",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
}
MUL = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3,
       "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "second": 1.0}


def main():
    rep, prefix, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    out = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0]
        d = {}
        for k, short in KEYS.items():
            if k in col and r[col[k]] not in ("", "n/a"):
                try:
                    v = float(r[col[k]].replace(",", ""))
                except ValueError:
                    continue
                d[short] = v * MUL.get(units[col[k]].lower(), 1)
        if "duration" in d and "dram_read" in d:
            d["dram_bytes"] = d["dram_read"] + d.get("dram_write", 0.0)
            d["dram_gbs"] = d["dram_bytes"] / d["duration"] / 1e9
            d["duration_us"] = d.pop("duration") * 1e6
        out[name] = d  # later launches overwrite earlier ones
    dst = Path(__file__).resolve().parent.parent / "profiles" / f"{prefix}_{tag}.json"
    dst.write_text(json.dumps({"source": Path(rep).name, "how": "ncu --set full --clock-control none (cold-ish caches, "
                               "serialised replay: durations are upper bounds, compare shares)", "kernels": out}, indent=1))
    for k, d in out.items():
        print(f"{k[:60]:60s} {d.get('duration_us', 0):8.1f} us  dram {d.get('dram_gbs', 0):7.0f} GB/s  tensor {d.get('tensor_pipe_pct', 0):5.1f}%  issue {d.get('issue_slot_pct', 0):5.1f}%")


if __name__ == "__main__":
    main()
