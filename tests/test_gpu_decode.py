"""Decode attention through head_sparse_decode_attention vs the CPU oracle / golden vectors."""
import pytest
import torch

from compactor_vllm.attention.sparse_decode_kernel import head_sparse_decode_attention
from oracle import attention as oa

from conftest import load_golden
from helpers import make_paged

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = {torch.float16: 3e-3, torch.bfloat16: 2e-2}  # reference test atol 3e-3 (fp16); north_star 2e-2 bf16


def _check(q, kc, vc, lens, pt, bm, HKV, PS, dtype, scale=None):
    out = head_sparse_decode_attention(q, kc, vc, lens, pt, bm, HKV, PS, scale)
    torch.cuda.synchronize()
    ref = oa.decode_attention(q.cpu(), kc.cpu(), vc.cpu(), lens.cpu(), pt.cpu(), bm.cpu(), HKV, PS, scale)
    assert torch.isfinite(out).all()
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= TOL[dtype], err
    return out


@pytest.mark.parametrize("name", ["decode_bf16", "decode_f16"])
def test_decode_golden(name):
    g = load_golden(name)
    dtype = g["q"].dtype
    out = head_sparse_decode_attention(g["q"].to(DEV), g["k_cache"].to(DEV), g["v_cache"].to(DEV),
                                       g["seq_lens_bh"].to(DEV), g["page_table"].to(DEV),
                                       g["batch_mapping"].to(DEV), g["HKV"], g["PS"])
    valid = g["valid_rows"]
    err = (out.float().cpu()[valid] - g["out"].float()[valid]).abs().max().item()
    assert err <= TOL[dtype], err
    assert (out.float().cpu()[~valid] == 0).all()


# the reference's own grid: tests/test_triton_attention.py:52-68 (B x cache_len, HQ=32 HKV=8 D=128)
@pytest.mark.parametrize("B", [1, 2, 3, 8])
@pytest.mark.parametrize("cache_len", [1, 2, 70, 128, 8000])
def test_decode_reference_grid(B, cache_len):
    dtype, HQ, HKV, D, PS = torch.float16, 32, 8, 128, 256
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, [[cache_len] * HKV] * B, dtype, DEV, seed=1234)
    g = torch.Generator().manual_seed(B * 1000 + cache_len)
    q = torch.randn((B, HQ, D), generator=g).to(dtype).to(DEV)
    lens = torch.full((B, HKV), cache_len, dtype=torch.int32, device=DEV)
    _check(q, kc, vc, lens, pt, bm, HKV, PS, dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("HQ,HKV,PS", [(32, 8, 128), (8, 1, 128), (16, 2, 32), (32, 8, 64)])
def test_decode_head_sparse_poisoned(dtype, HQ, HKV, PS):
    """Ragged per-head lengths, empty heads, scrambled pages, non-identity slots, NaN in every
    cache row a kernel has no business reading."""
    B, D = 5, 128
    g = torch.Generator().manual_seed(HQ + PS)
    lens = torch.randint(0, 3000, (B, HKV), generator=g)
    lens[0, 0] = 0
    lens[1, :] = 0
    lens[2, -1] = 1
    lens[3, 0] = 2048
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=11, poison=True)
    q = torch.randn((B, HQ, D), generator=g).to(dtype).to(DEV)
    out = _check(q, kc, vc, lens.to(torch.int32).to(DEV), pt, bm, HKV, PS, dtype, scale=0.11)
    G = HQ // HKV
    empty = (lens == 0).repeat_interleave(G, dim=1)
    assert (out.cpu()[empty] == 0).all()


def test_decode_q_4d_and_key_split_independent():
    dtype, B, HQ, HKV, D, PS = torch.bfloat16, 2, 32, 8, 128, 128
    lens = [[5000] * HKV, [300] * HKV]
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens, dtype, DEV, seed=2)
    q = torch.randn((B, HQ, 1, D), generator=torch.Generator().manual_seed(0)).to(dtype).to(DEV)
    l = torch.tensor(lens, dtype=torch.int32, device=DEV)
    a = head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS, None, key_split=1)
    b = head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS, None, key_split=7)
    assert a.shape == (B, HQ, D)
    assert torch.equal(a, b)


def test_decode_large_batch_balanced():
    """Config-3-like raggedness at reduced size: many heads, many CTAs share heads."""
    dtype, B, HQ, HKV, D, PS = torch.bfloat16, 16, 32, 8, 128, 128
    g = torch.Generator().manual_seed(9)
    lens = torch.randint(1, 6000, (B, HKV), generator=g)
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=5)
    q = torch.randn((B, HQ, D), generator=g).to(dtype).to(DEV)
    _check(q, kc, vc, lens.to(torch.int32).to(DEV), pt, bm, HKV, PS, dtype)


def test_decode_cuda_graph_capturable():
    dtype, B, HQ, HKV, D, PS = torch.bfloat16, 4, 32, 8, 128, 128
    lens = torch.randint(1, 2000, (B, HKV), generator=torch.Generator().manual_seed(1))
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=8)
    q = torch.randn((B, HQ, D), generator=torch.Generator().manual_seed(2)).to(dtype).to(DEV)
    l = lens.to(torch.int32).to(DEV)
    eager = head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS)  # warm the workspace
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            out = head_sparse_decode_attention(q, kc, vc, l, pt, bm, HKV, PS)
    torch.cuda.current_stream().wait_stream(s)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)


def test_decode_step_matches_store_then_attend_with_reserved_rows():
    """cvllm_decode_step (slot-indexed lengths, what Attention.forward's decode branch calls) against the
    reference-shaped sequence index_select -> decode_store_kv -> attention -> index_copy_, bit for bit;
    rows mapped to RESERVED_BATCH (CUDA-graph padding, core/model_runner.py:484) are skipped."""
    from compactor_vllm.attention.sparse_decode_kernel import decode_append_and_attend
    from compactor_vllm.kv_cache.store_kv_cache import decode_store_kv
    dtype, B, HQ, HKV, D, PS = torch.bfloat16, 6, 32, 8, 128, 128
    g = torch.Generator().manual_seed(4)
    lens = torch.randint(0, 1500, (B, HKV), generator=g)
    lens[2, 3] = 0
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, (lens + 2).tolist(), dtype, DEV, seed=21, n_extra_slots=3)
    bm = bm.clone()
    bm[1] = 0
    bm[4] = 0  # two padding rows
    n_slots = pt.shape[0]
    table = torch.zeros((n_slots, HKV), dtype=torch.int32, device=DEV)
    table[bm.long()] = lens.to(torch.int32).to(DEV)
    table[0] = 0
    q = torch.randn((B, HQ, D), generator=g).to(dtype).to(DEV)
    kn = torch.randn((B, HKV, D), generator=g).to(dtype).to(DEV)
    vn = torch.randn((B, HKV, D), generator=g).to(dtype).to(DEV)

    kc2, vc2, table2 = kc.clone(), vc.clone(), table.clone()
    gathered = table2.index_select(0, bm.long()).contiguous()
    decode_store_kv(key=kn, value=vn, batch_mapping=bm, bh_lens=gathered, page_table=pt, k_cache=kc2,
                    v_cache=vc2, PAGE_SIZE=PS)
    ref = head_sparse_decode_attention(q, kc2, vc2, gathered, pt, bm, HKV, PS)
    live = bm != 0
    table2[bm[live].long()] = gathered[live]

    out = decode_append_and_attend(q, kn, vn, kc, vc, table, pt, bm, HKV, PS)
    torch.cuda.synchronize()
    assert torch.equal(table, table2) and (table[0] == 0).all()
    assert torch.equal(kc.view(torch.int16), kc2.view(torch.int16)) and torch.equal(vc.view(torch.int16), vc2.view(torch.int16))
    assert torch.equal(out[live], ref[live])
    assert (out[~live] == 0).all()
