"""GPU parity of the fused row-parallel linear + all-reduce (reference layers/linear.py:133-153:
`F.linear(x, W_shard)` then `dist.all_reduce(y)`), csrc/gemm_reduce.cu through the C ABI `cvllm_gemm_allreduce`.

On ONE GPU the whole protocol runs with several VIRTUAL ranks: `world` buffers in this process, `world` kernels on
`world` streams, each capped to 1/world of the CTA pairs so they are co-resident (the counters, STARTED / DONE flags,
rotated owner order, reducer rounds and peer stores are exactly those of `world` GPUs; only the wires differ).
Bars: (1) world == 1 against a torch fp32 GEMM within bf16 rounding; (2) world > 1 BIT-EXACT against the fp32
rank-order sum, rounded once, of each rank's own world == 1 output (the all-reduce's contract, csrc/allreduce.cu)."""
import json
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _option(name, value):
    from compactor_vllm import _native as N
    N.check(N.lib().cvllm_set_test_option(name.encode(), float(value)), name)


@pytest.fixture
def options():
    yield _option
    _option("grs_pairs", 0)
    _option("grs_rows_per_round", 0)


def _shards(m, n, k_total, world, dtype, seed):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(seed)
    x = (torch.randn((m, k_total), generator=g, device=dev) * 0.5).to(dtype)
    w = (torch.randn((n, k_total), generator=g, device=dev) * 0.05).to(dtype)
    per = k_total // world
    return x, w, [(x[:, r * per:(r + 1) * per], w[:, r * per:(r + 1) * per]) for r in range(world)]


def _run_world(world, m, n, shards, pairs_total=74, calls=1):
    """y of every virtual rank after `calls` back-to-back calls (the flags must carry over)."""
    from compactor_vllm.layers.linear import PeerGemmAllReduce
    dev = torch.device("cuda", 0)
    ranks, free = PeerGemmAllReduce.virtual_ranks(world, m, n, dev)
    _option("grs_pairs", max(1, pairs_total // world))
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    try:
        outs = None
        for _ in range(calls):
            torch.cuda.synchronize()
            outs = []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    outs.append(ranks[r].linear(*shards[r]))
            torch.cuda.synchronize()
        return [o.clone() for o in outs]
    finally:
        _option("grs_pairs", 0)
        free()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (1000, 512, 192), (4096, 8192, 1024), (300, 264, 72)])
def test_world1_matches_fp32_gemm(options, dtype, m, n, k):
    x, w, shards = _shards(m, n, k, 1, dtype, seed=m + n + k)
    (y,) = _run_world(1, m, n, shards)
    ref = x.float() @ w.float().t()
    err = (y.float() - ref).abs()
    bound = ref.abs() * 2.0 ** (-8 if dtype == torch.bfloat16 else -11) + 1e-3
    assert bool((err <= bound).all()), float((err - bound).max())
    # and the library GEMM rounds the same fp32 sums to (nearly always) the same 16-bit values
    lib = torch.mm(x, w.t())
    assert float((y != lib).float().mean()) < 2e-2


@pytest.mark.parametrize("world,m,n,k", [(2, 1024, 512, 256), (2, 777, 512, 128), (4, 2048, 1024, 512),
                                         (4, 100, 256, 256), (8, 4096, 512, 512), (3, 1500, 768, 192)])
def test_virtual_ranks_bit_exact_rank_order_sum(options, world, m, n, k):
    dtype = torch.float16 if (world, m) == (2, 777) else torch.bfloat16  # one case on the f16 instantiation
    x, w, shards = _shards(m, n, k, world, dtype, seed=world * 1000 + m)
    partial = [_run_world(1, m, n, [shards[r]])[0] for r in range(world)]  # each rank's own rounded GEMM
    acc = torch.zeros((m, n), dtype=torch.float32, device=x.device)
    for p in partial:
        acc += p.float()  # rank order, fp32
    want = acc.to(dtype)
    ys = _run_world(world, m, n, shards, calls=3)
    for r, y in enumerate(ys):
        assert torch.equal(y, want), (r, float((y.float() - want.float()).abs().max()))
    ref = x.float() @ w.float().t()
    assert float((ys[0].float() - ref).abs().max()) <= float(ref.abs().max()) * 2.0 ** -6


def test_rounds_and_changing_shapes_share_one_buffer(options):
    """Several reduce rounds per owner (rows_per_round 256 on 2048 rows per rank), then a smaller M on the same
    buffers: the counters are reset per call and the layout depends on the call's M only."""
    from compactor_vllm.layers.linear import PeerGemmAllReduce
    dev, dtype, world, n, k = torch.device("cuda", 0), torch.bfloat16, 2, 512, 256
    options("grs_rows_per_round", 256)
    ranks, free = PeerGemmAllReduce.virtual_ranks(world, 4096, n, dev)
    options("grs_pairs", 37)
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    try:
        for m in (4096, 640, 4096, 1):
            x, w, shards = _shards(m, n, k, world, dtype, seed=m)
            torch.cuda.synchronize()
            outs = []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    outs.append(ranks[r].linear(*shards[r]))
            torch.cuda.synchronize()
            ref = x.float() @ w.float().t()
            assert torch.equal(outs[0], outs[1])
            assert float((outs[0].float() - ref).abs().max()) <= float(ref.abs().max()) * 2.0 ** -6 + 1e-3
    finally:
        free()


def test_refuses_unaligned_shapes(options):
    from compactor_vllm import _native as N
    from compactor_vllm.layers.linear import PeerGemmAllReduce
    dev = torch.device("cuda", 0)
    ranks, free = PeerGemmAllReduce.virtual_ranks(1, 256, 256, dev)
    try:
        x = torch.zeros((256, 60), dtype=torch.bfloat16, device=dev)
        w = torch.zeros((256, 60), dtype=torch.bfloat16, device=dev)
        assert not ranks[0].supports(x, w)
        with pytest.raises((AssertionError, N.NativeLibraryError)):
            ranks[0].linear(x, w)
    finally:
        free()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NVLink peer memory between processes)")
@pytest.mark.parametrize("rows", [4096, 1000])
def test_two_processes_two_gpus(rows):
    """One process per GPU, buffers shared by cudaIpc handles: every rank holds the same bits, and they are the fp32
    rank-order sum of the ranks' own cuBLAS partials (tools/gemm_reduce_bench.py does the checks on the devices)."""
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), str(ROOT / "tools" / "gemm_reduce_bench.py"), str(rows)]
    run = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stderr[-2000:]
    line = [ln for ln in run.stdout.splitlines() if ln.startswith("{")][-1]
    res = json.loads(line)
    assert res["identical_on_all_ranks"] is True
    assert res["frac_equal_rank_order_sum_of_cublas_partials"] == 1.0
    assert res["max_abs_diff_vs_nccl"] <= res["max_abs"] * 2.0 ** -7
