"""Fused row-parallel linear + all-reduce (csrc/gemm_reduce.cu) against cuBLAS + NCCL on the config-5 o_proj:
x[32768, 8192 / tp] . W[8192, 8192 / tp]^T summed over tp ranks.

  python tools/gemm_reduce_bench.py                                   # one GPU: the GEMM alone against torch.mm
  torchrun --nproc-per-node N ... tools/gemm_reduce_bench.py          # N ranks: fused kernel vs torch.mm + all_reduce

Prints one JSON object (rank 0).  Times are CUDA events on the launching stream, max over ranks, after warm-up."""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "compactor-vllm_b200"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def timed(fn, reps, sync):
    for _ in range(3):
        fn()
    sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def single_gpu(dev, m=32768, n=8192):
    from compactor_vllm.layers.linear import PeerGemmAllReduce
    (pg,), free = PeerGemmAllReduce.virtual_ranks(1, m, n, dev)
    res = {"mode": "one GPU, world 1 (GEMM + local slot + reducer copy)", "rows": []}
    g = torch.Generator(device=dev).manual_seed(1)
    for k in (1024, 2048, 4096, 8192):
        x = (torch.randn((m, k), generator=g, device=dev) * 0.5).to(torch.bfloat16)
        w = (torch.randn((n, k), generator=g, device=dev) * 0.02).to(torch.bfloat16)
        y = torch.empty((m, n), dtype=torch.bfloat16, device=dev)
        ms_lib = timed(lambda: torch.mm(x, w.t(), out=y), 10, torch.cuda.synchronize)
        ms_own = timed(lambda: pg.linear(x, w), 10, torch.cuda.synchronize)
        got = pg.linear(x, w)
        torch.cuda.synchronize()
        flops = 2.0 * m * n * k
        res["rows"].append({"k": k, "cublas_ms": round(ms_lib, 4), "fused_ms": round(ms_own, 4),
                            "cublas_tflops": round(flops / ms_lib / 1e9, 1), "fused_tflops": round(flops / ms_own / 1e9, 1),
                            "mismatch_vs_cublas": float((got != y).float().mean()),
                            "max_abs_diff": float((got.float() - y.float()).abs().max())})
        del x, w, y
    free()
    return res


def multi_gpu(dev, rank, world, m=32768, n=8192, hidden=8192):
    import torch.distributed as dist
    from compactor_vllm.layers.linear import PeerGemmAllReduce
    pg = PeerGemmAllReduce(m, n, device=dev)
    k = hidden // world
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x = (torch.randn((m, k), generator=g, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device=dev) * 0.02).to(torch.bfloat16)
    y = torch.empty((m, n), dtype=torch.bfloat16, device=dev)

    def sync():
        torch.cuda.synchronize()
        dist.barrier()

    def lib_path():
        torch.mm(x, w.t(), out=y)
        dist.all_reduce(y)

    mx = lambda v: float(_allmax(v, dev))  # noqa: E731
    ms_gemm = mx(timed(lambda: torch.mm(x, w.t(), out=y), 10, sync))
    ms_lib = mx(timed(lib_path, 10, sync))
    ms_fused = mx(timed(lambda: pg.linear(x, w), 10, sync))
    # correctness: against NCCL's sum (different rounding order: tolerance) and identical bits on every rank
    lib_path()
    got = pg.linear(x, w)
    torch.cuda.synchronize()
    diff = float((got.float() - y.float()).abs().max())
    scale = float(y.float().abs().max())
    digest = torch.stack([got.view(torch.int16).to(torch.int64).sum(), (got.view(torch.int16).to(torch.int64) *
                          torch.arange(got.numel(), device=dev).view_as(got) % 1000003).sum()])
    all_d = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    same = all(bool(torch.equal(d, all_d[0])) for d in all_d)
    # exact contract: fp32 rank-order sum of the ranks' own rounded partials (gathered on rank 0 for a slice of rows)
    rows = min(512, m)
    part = torch.mm(x[:rows], w.t())
    parts = [torch.zeros_like(part) for _ in range(world)]
    dist.all_gather(parts, part)
    acc = torch.zeros((rows, n), dtype=torch.float32, device=dev)
    for p_ in parts:
        acc += p_.float()
    frac_exact = float((acc.to(torch.bfloat16) == got[:rows]).float().mean())
    msg = m * n * 2
    res = {"mode": f"{world} ranks", "tp": world, "m": m, "n": n, "k_local": k,
           "cublas_gemm_ms": round(ms_gemm, 4), "cublas_plus_nccl_ms": round(ms_lib, 4), "fused_ms": round(ms_fused, 4),
           "speedup": round(ms_lib / ms_fused, 3),
           "fused_nvlink_gbs_per_rank": round(2 * (world - 1) / world * msg / (ms_fused * 1e-3) / 1e9, 1),
           "max_abs_diff_vs_nccl": diff, "max_abs": scale, "identical_on_all_ranks": same,
           "frac_equal_rank_order_sum_of_cublas_partials": frac_exact}
    dist.barrier()
    pg.close()
    return res


def _allmax(v, dev):
    import torch.distributed as dist
    t = torch.tensor([v], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world == 1:
        print(json.dumps(single_gpu(dev)))
        return
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    res = multi_gpu(dev, rank, world, m=m)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
