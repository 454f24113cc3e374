"""GPU parity of the step after the hot path: the one-shot peer-memory all-reduce behind
`RowParallelLinear` (reference layers/linear.py:133-153).  On one GPU the kernel runs with world == 1 (own
buffer only: barrier words, epochs, staging and graph replay are all exercised); with >= 2 GPUs visible the
2-rank check in tools/tp_allreduce_check.py is launched as one process per GPU.
Bar: bit-exact against the fp32 rank-order sum rounded once to the 16-bit dtype."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.fixture
def single_rank_group():
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()), RANK="0", WORLD_SIZE="1")
    dist.init_process_group("gloo", rank=0, world_size=1)
    yield
    dist.destroy_process_group()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_oneshot_allreduce_world1(single_rank_group, dtype):
    from compactor_vllm.layers.linear import PeerAllReduce

    dev = torch.device("cuda", 0)
    par = PeerAllReduce(max_bytes=1 << 20, device=dev)
    try:
        g = torch.Generator(device=dev).manual_seed(5)
        for rows in (1, 3, 16, 64):
            x = torch.randn((rows, 8192), generator=g, device=dev).to(dtype)
            got = par.all_reduce(x.clone())
            assert torch.equal(got, x)  # one rank: fp32 round trip of a 16-bit value is the identity
        # repeated calls advance the epochs; a captured call replays with the same arguments
        st = par.staging((16, 8192), dtype)
        st.copy_(torch.randn((16, 8192), generator=g, device=dev).to(dtype))
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            par.reduce_staged(st)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                out = par.reduce_staged(st)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(4):
            graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, st)
        # the two-shot form (slice reduce + gather through the result area) gives the same bits
        par2 = PeerAllReduce(max_bytes=1 << 20, device=dev, twoshot_min_bytes=0)
        try:
            for shape in ((64, 8192), (3, 40), (1, 8)):
                x = torch.randn(shape, generator=g, device=dev).to(dtype)
                assert torch.equal(par2.all_reduce(x.clone()), x)
        finally:
            par2.close()
        # too large for the staging buffer -> the process-group route, same result
        big = torch.randn((128, 8192), generator=g, device=dev).to(dtype)
        assert not par.fits(big)
    finally:
        par.close()


def test_oneshot_allreduce_rejects_bad_arguments():
    import ctypes as C

    from compactor_vllm import _native as N

    lib = N.lib()
    bases = (C.c_void_p * 1)(None)
    out = torch.empty(16, dtype=torch.bfloat16, device="cuda")
    assert lib.cvllm_allreduce_oneshot(bases, 0, 9, N.ptr(out), 16, N.dtype_code(out.dtype), None) < 0  # world > 8
    assert lib.cvllm_allreduce_oneshot(bases, 1, 1, N.ptr(out), 16, N.dtype_code(out.dtype), None) < 0  # rank >= world
    assert lib.cvllm_allreduce_oneshot(bases, 0, 1, N.ptr(out), 12, N.dtype_code(out.dtype), None) < 0  # numel % 8


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one NVLink box")
def test_oneshot_allreduce_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(ROOT / "tools" / "tp_allreduce_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"ok": true' in r.stdout
