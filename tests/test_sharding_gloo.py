"""N > 1 host logic on CPU: two ranks over gloo (the path shards by KV head with no
exchange; the only collectives are the harness's barrier / max / sum)."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "compactor-vllm_b200", ROOT / "tests"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import bench
    import oracle.selection as osel
    from compactor_vllm.utils import arguments as A

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ranks = bench.Ranks(world, "cpu", init=False)
        # ---- harness reductions (what bench.py uses for max-over-ranks timing / whole-job totals)
        assert ranks.max(float(rank + 1)) == float(world)
        assert ranks.sum(float(rank + 1)) == world * (world + 1) / 2
        ranks.barrier()

        # ---- KV-head sharding: every rank ranks ITS heads with ITS budget, no exchange
        HQ, HKV, D, PS = 32, 8, 16, 16
        lens, ratios, first, last = [300, 77], [0.5, 0.25], [4, 2], [8, 3]
        kv0, kv1, q0, q1 = A.kv_head_shard(rank, world, HQ, HKV)
        H_local = kv1 - kv0
        assert (q1 - q0) == H_local * (HQ // HKV)
        g = torch.Generator().manual_seed(1234)  # same inputs on every rank
        n = sum(lens)
        scores = torch.randn((n, HKV), generator=g)
        keys = torch.randn((n, HKV, D), generator=g).to(torch.float16)
        vals = torch.randn((n, HKV, D), generator=g).to(torch.float16)
        cu = torch.tensor([0, lens[0], n], dtype=torch.int32)
        retain = A.tokens_to_retain(lens, ratios, first, last, H_local)
        n_lp = max(lens) // PS + 2
        pt = torch.arange(3 * H_local * n_lp, dtype=torch.int32).view(3, H_local, n_lp)
        kc = torch.zeros((3 * H_local * n_lp * PS, D), dtype=torch.float16)
        vc = torch.zeros_like(kc)
        bh = torch.zeros((2, H_local), dtype=torch.int32)
        bm = torch.tensor([1, 2], dtype=torch.int32)
        osel.extract_and_store_top_kv(scores[:, kv0:kv1].contiguous(), cu, max(lens), max(lens) * H_local, H_local,
                                      keys[:, kv0:kv1].contiguous(), vals[:, kv0:kv1].contiguous(), retain, pt, bm,
                                      bh, kc, vc, PS, pad=False)
        # the local budget is met exactly (no padding) and nothing else is kept
        assert bh.sum(dim=1).tolist() == [min(r, L * H_local) for r, L in zip(retain, lens)]
        # gather the shards' results: together they cover all heads exactly once
        shard = torch.tensor([kv0, kv1], dtype=torch.int64)
        shards = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(shards, shard)
        kept = [torch.zeros((2, H_local), dtype=torch.int32) for _ in range(world)]
        dist.all_gather(kept, bh)
        if rank == 0:
            covered = sorted(h for a, b in (s.tolist() for s in shards) for h in range(a, b))
            assert covered == list(range(HKV))
            # budgets are per rank (utils/arguments.py:109-121 uses the LOCAL head count): the shards'
            # kept totals add up to the per-rank budgets, which differ from one global top-k by rounding
            total = torch.stack(kept).sum(dim=(0, 2)).tolist()
            glob = A.tokens_to_retain(lens, ratios, first, last, HKV)
            for t, gl in zip(total, glob):
                assert abs(t - gl) <= world
        Path(out_dir, f"ok{rank}").write_text("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_kv_head_shard_partitions_heads():
    from compactor_vllm.utils import arguments as A
    for world, hq, hkv in [(1, 32, 8), (2, 32, 8), (4, 64, 8), (8, 64, 8)]:
        got = [A.kv_head_shard(r, world, hq, hkv) for r in range(world)]
        assert [g[0] for g in got] == [r * hkv // world for r in range(world)]
        assert got[-1][1] == hkv and got[-1][3] == hq
        with pytest.raises(AssertionError):
            A.kv_head_shard(0, 3, hq, hkv)


def _row_parallel_worker(rank: int, world: int, port: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from compactor_vllm.layers.linear import RowParallelLinear, shard_rows

        g = torch.Generator().manual_seed(3)  # the same full weight / input on every rank
        W = torch.randn((24, 16), generator=g)
        bias = torch.randn(24, generator=g)
        x = torch.randn((5, 16), generator=g)
        lin = RowParallelLinear(16, 24, bias=True)  # no peer buffer: torch.distributed carries the reduce
        assert lin.tp_size == world and lin.weight.shape == (24, 16 // world)
        with torch.no_grad():
            lin.weight.copy_(shard_rows(W, rank, world))
            lin.bias.copy_(bias)
            per = 16 // world
            y = lin(x[:, rank * per:(rank + 1) * per])
        want = x @ W.t() + bias  # the bias is added on rank 0 only (reference layers/linear.py:149)
        assert torch.allclose(y, want, atol=1e-5), (y - want).abs().max()
        Path(out_dir, f"rp{rank}").write_text("ok")
    finally:
        dist.destroy_process_group()


def test_row_parallel_linear_two_rank_gloo(tmp_path):
    world = 2
    mp.spawn(_row_parallel_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"rp{r}").exists() for r in range(world))
