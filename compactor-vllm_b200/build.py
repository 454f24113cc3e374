"""Build libcompactor_sm100.so in-tree with nvcc (sm_100a only).

    python compactor-vllm_b200/build.py [--force] [--verbose]

Each csrc/*.cu is compiled to build/<name>.o (rebuilt when it or a header is
newer), then linked into lib/libcompactor_sm100.so.  No CMake, no torch: the
library is a plain C-ABI .so (include/cvllm.h) loaded with ctypes.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
BUILD = HERE / "build"
LIB = HERE / "lib" / "libcompactor_sm100.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def _newest_header() -> float:
    hdrs = list(CSRC.glob("*.cuh")) + list((HERE.parent / "include").glob("*.h"))
    return max((h.stat().st_mtime for h in hdrs), default=0.0)


def _compile(src: Path, obj: Path, verbose: bool) -> str:
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    (obj.with_suffix(".ptxas.txt")).write_text(res.stderr)
    return res.stderr if verbose else ""


def build(force: bool = False, verbose: bool = False) -> Path:
    BUILD.mkdir(exist_ok=True)
    LIB.parent.mkdir(exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    hdr_time = _newest_header()
    todo = []
    for s in srcs:
        o = BUILD / (s.stem + ".o")
        if force or not o.exists() or o.stat().st_mtime < max(s.stat().st_mtime, hdr_time):
            todo.append((s, o))
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            for out in ex.map(lambda so: _compile(so[0], so[1], verbose), todo):
                if out:
                    print(out)
    objs = [BUILD / (s.stem + ".o") for s in srcs]
    if todo or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-o", str(LIB), *map(str, objs)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose))
    sys.exit(0)
