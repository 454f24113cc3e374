"""Multi-GPU check + timing of the one-shot peer all-reduce.  One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/tp_allreduce_check.py

Parity: against NCCL `all_reduce` of the same partials (fp32 reference = sum of the bf16 partials in fp32,
rounded once), identical bits on every rank, through `RowParallelLinear` with a sharded weight against the
unsharded GEMM, and replayed from a CUDA graph.  Timing: CUDA events, max over ranks, decode-size messages
([B, 8192] bf16, Llama-3.1-70B hidden) and the config-5 prefill message on NCCL for context.  Rank 0 prints JSON."""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]


def timed(fn, iters=50):
    for _ in range(5):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    b.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3  # us


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    res = run_checks(dev, rank, world, timing=True)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


def run_checks(dev, rank: int, world: int, timing: bool = True) -> dict:
    """All parity checks (and, with timing=True, the latency tables).  Needs an initialised NCCL process group;
    bench.py calls it with timing=False inside its --gpus N run so the driver's own launch exercises the
    multi-rank all-reduce."""
    from compactor_vllm.layers.linear import PeerAllReduce, RowParallelLinear, shard_rows

    par = PeerAllReduce(max_bytes=8 << 20, twoshot_min_bytes=1 << 62)  # one-shot only
    par2 = PeerAllReduce(max_bytes=8 << 20, twoshot_min_bytes=0)      # two-shot only
    res = {"world": world, "checks": {}, "us": {}}
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    ok = True
    for rows in (1, 16, 64, 256, 509):
        x = torch.randn((rows, 8192), generator=g, device=dev).bfloat16()
        got = par.all_reduce(x.clone())
        # fp32 reference: gather all partials, sum in rank order, round once
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        want = torch.stack([p.float() for p in parts]).sum(0)
        # the kernel accumulates in fp32 in rank order: same arithmetic
        acc = torch.zeros_like(want)
        for p in parts:
            acc += p.float()
        exact = torch.equal(got, acc.bfloat16())
        got2 = par2.all_reduce(x.clone())
        exact2 = torch.equal(got2, got)  # the two-shot form sums in the same order: same bits
        res["checks"][f"twoshot_rows{rows}_equals_oneshot"] = exact2
        ok &= exact2
        same = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(same, got)
        identical = all(torch.equal(same[0], s) for s in same)
        nccl = x.clone()
        dist.all_reduce(nccl)
        close = (got.float() - want).abs().max().item() <= 2e-2 * want.abs().max().item()
        near_nccl = (got.float() - nccl.float()).abs().max().item() <= 4e-2 * want.abs().max().item()
        res["checks"][f"rows{rows}"] = {"bit_exact_vs_fp32_rank_order": exact, "identical_on_all_ranks": identical,
                                        "close_to_fp32": close, "close_to_nccl": near_nccl}
        ok &= exact and identical and close and near_nccl
        if timing:
            xs = x.clone()
            res["us"][f"peer_{rows}x8192"] = timed(lambda: par.all_reduce(xs))
            xn = x.clone()
            res["us"][f"nccl_{rows}x8192"] = timed(lambda: dist.all_reduce(xn))

    # ragged slices (15 vectors over `world` ranks) and an empty slice (1 vector)
    for shape in ((3, 40), (1, 8)):
        xr = torch.randn(shape, generator=g, device=dev).bfloat16()
        same_bits = torch.equal(par2.all_reduce(xr.clone()), par.all_reduce(xr.clone()))
        res["checks"][f"twoshot_ragged_{shape[0]}x{shape[1]}"] = same_bits
        ok &= same_bits

    # stress: 64 back-to-back two-shot calls on fresh data (no host sync in between), every output checked against
    # the fp32 rank-order sum — a stale slice (a fence that is too weak) shows up as a bit difference
    xs_all = [torch.randn((96 + 8 * (i % 5), 8192), generator=g, device=dev).bfloat16() for i in range(64)]
    outs = [par2.all_reduce(xi.clone()) for xi in xs_all]  # all_reduce is in place (dist.all_reduce semantics)
    stress_ok = True
    for xi, oi in zip(xs_all, outs):
        parts = [torch.empty_like(xi) for _ in range(world)]
        dist.all_gather(parts, xi)
        acc = torch.zeros(xi.shape, dtype=torch.float32, device=dev)
        for pp in parts:
            acc += pp.float()
        stress_ok &= torch.equal(oi, acc.bfloat16())
    res["checks"]["twoshot_stress_64_calls_bit_exact"] = bool(stress_ok)
    ok &= bool(stress_ok)

    # in-place contract on both routes: the argument holds the sum afterwards and is the returned tensor
    xi = torch.randn((16, 8192), generator=g, device=dev).bfloat16()
    xc = xi.clone()
    ret = par.all_reduce(xc)
    nc = xi.clone()
    dist.all_reduce(nc)
    inplace_ok = ret.data_ptr() == xc.data_ptr() and (xc.float() - nc.float()).abs().max().item() <= 4e-2 * nc.float().abs().max().item()
    res["checks"]["all_reduce_in_place"] = bool(inplace_ok)
    ok &= bool(inplace_ok)

    # interleaved stress on ONE buffer: message sizes hop across the one-shot / two-shot switch point (and the
    # block count changes from call to call, so per-block epochs advance unevenly), 96 calls with no host sync
    par_mix = PeerAllReduce(max_bytes=8 << 20, twoshot_min_bytes=1 << 20)
    rows_mix = [1, 8, 40, 64, 65, 128, 300, 509, 2, 96, 63, 256]
    xs_mix = [torch.randn((rows_mix[(7 * i + i // 5) % len(rows_mix)], 8192), generator=g, device=dev).bfloat16()
              for i in range(96)]
    outs = [par_mix.all_reduce(xi.clone()) for xi in xs_mix]
    mix_ok = True
    for xi, oi in zip(xs_mix, outs):
        parts = [torch.empty_like(xi) for _ in range(world)]
        dist.all_gather(parts, xi)
        acc = torch.zeros(xi.shape, dtype=torch.float32, device=dev)
        for pp in parts:
            acc += pp.float()
        mix_ok &= torch.equal(oi, acc.bfloat16())
    res["checks"]["interleaved_oneshot_twoshot_stress_96_calls_bit_exact"] = bool(mix_ok)
    ok &= bool(mix_ok)
    par_mix.close()

    # RowParallelLinear: sharded weight + peer all-reduce == unsharded GEMM (bf16 tolerance)
    gw = torch.Generator(device=dev).manual_seed(7)  # same seed on every rank: the full weight
    W = (0.02 * torch.randn((8192, 8192), generator=gw, device=dev)).bfloat16()
    xin = torch.randn((16, 8192), generator=gw, device=dev).bfloat16()
    lin = RowParallelLinear(8192, 8192, peer_all_reduce=par).to(dev).to(torch.bfloat16)
    with torch.no_grad():
        lin.weight.copy_(shard_rows(W, rank, world))
        per = 8192 // world
        y = lin(xin[:, rank * per:(rank + 1) * per].contiguous())
        full = xin.float() @ W.float().t()
        rel = ((y.float() - full).norm() / full.norm()).item()
    res["checks"]["row_parallel_rel_err"] = rel
    ok &= rel < 2e-2

    # CUDA graph replay of GEMM -> staging -> one-shot all-reduce (the decode step's shape of use)
    with torch.no_grad():
        xs = xin[:, rank * per:(rank + 1) * per].contiguous()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            lin(xs)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                yg = lin(xs)
        torch.cuda.current_stream().wait_stream(side)
        dist.barrier()
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        res["checks"]["graph_replay_equals_eager"] = bool(torch.equal(yg, y))
        ok &= res["checks"]["graph_replay_equals_eager"]
        if not timing:
            res["ok"] = bool(ok)
            par.close()
            par2.close()
            return res
        res["us"]["graph_gemm_plus_peer_allreduce_16x8192"] = timed(graph.replay)
        lin_nccl = RowParallelLinear(8192, 8192).to(dev).to(torch.bfloat16)
        lin_nccl.weight.copy_(shard_rows(W, rank, world))
        res["us"]["eager_gemm_plus_nccl_allreduce_16x8192"] = timed(lambda: lin_nccl(xs))

    # like for like: both routes replayed from CUDA graphs (no Python launch path), 8 calls per replay
    def graphed(fn, reps=8):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=side):
                for _ in range(reps):
                    fn()
        torch.cuda.current_stream().wait_stream(side)
        dist.barrier()
        return timed(gr.replay) / reps

    with torch.no_grad():
        for rows in (1, 16, 64, 128, 256, 512):
            st = par.staging((rows, 8192), torch.bfloat16)
            st.normal_()
            res["us"][f"graph_peer_allreduce_{rows}x8192"] = graphed(lambda: par.reduce_staged(st))
            st2 = par2.staging((rows, 8192), torch.bfloat16)
            st2.normal_()
            res["us"][f"graph_twoshot_allreduce_{rows}x8192"] = graphed(lambda: par2.reduce_staged(st2))
            xn = torch.randn((rows, 8192), device=dev).bfloat16()
            try:
                res["us"][f"graph_nccl_allreduce_{rows}x8192"] = graphed(lambda: dist.all_reduce(xn))
            except Exception as ex:
                res["us"][f"graph_nccl_allreduce_{rows}x8192"] = repr(ex)[:80]
        res["us"]["graph_gemm_peer_16x8192"] = graphed(lambda: lin(xs))
        try:
            res["us"]["graph_gemm_nccl_16x8192"] = graphed(lambda: lin_nccl(xs))
        except Exception as ex:
            res["us"]["graph_gemm_nccl_16x8192"] = repr(ex)[:80]

    # config 5's prefill message on NCCL, for context ([32768, 8192] bf16 = 512 MiB)
    big = torch.randn((32768, 8192), device=dev).bfloat16()
    us = timed(lambda: dist.all_reduce(big), iters=5)
    res["us"]["nccl_32768x8192"] = us
    res["nccl_prefill_busbw_gbs"] = big.numel() * 2 * 2 * (world - 1) / world / (us * 1e-6) / 1e9
    res["ok"] = bool(ok)
    par.close()
    par2.close()
    return res


if __name__ == "__main__":
    main()
