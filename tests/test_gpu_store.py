"""Store kernels through the reference-shaped wrappers vs the CPU oracle and the golden vectors."""
import pytest
import torch

from compactor_vllm.kv_cache.store_kv_cache import decode_store_kv, prefill_store_all_kv
from compactor_vllm.kv_cache.write_page_table import scatter_to_page_table
from oracle import store as ost

from conftest import load_golden
from helpers import make_paged

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(t):
    return t.view(torch.int16) if t.element_size() == 2 else t


def test_decode_store_golden():
    g = load_golden("decode_store")
    kc, vc, bl = g["k_cache_in"].to(DEV), g["v_cache_in"].to(DEV), g["bh_lens_in"].to(DEV)
    decode_store_kv(key=g["key"].to(DEV), value=g["value"].to(DEV), batch_mapping=g["batch_mapping"].to(DEV),
                    bh_lens=bl, page_table=g["page_table"].to(DEV), k_cache=kc, v_cache=vc, PAGE_SIZE=g["PS"])
    assert torch.equal(bl.cpu(), g["bh_lens_out"])
    assert torch.equal(_bits(kc.cpu()), _bits(g["k_cache_out"]))
    assert torch.equal(_bits(vc.cpu()), _bits(g["v_cache_out"]))


def test_store_all_golden():
    g = load_golden("store_all")
    kc, vc, bl = g["k_cache_in"].to(DEV), g["v_cache_in"].to(DEV), g["bh_lens_in"].to(DEV)
    cu = g["cu_seqlens_k"].to(DEV)
    prefill_store_all_kv(new_keys=g["keys"].to(DEV), new_values=g["vals"].to(DEV), cu_seqlens_k=cu,
                         max_seqlen_k=int((cu[1:] - cu[:-1]).max()), k_cache=kc, v_cache=vc,
                         page_table=g["page_table"].to(DEV), bh_lens=bl,
                         batch_mapping=g["batch_mapping"].to(DEV), PAGE_SIZE=g["PS"])
    assert torch.equal(bl.cpu(), g["bh_lens_out"])
    assert torch.equal(_bits(kc.cpu()), _bits(g["k_cache_out"]))
    assert torch.equal(_bits(vc.cpu()), _bits(g["v_cache_out"]))


def test_scatter_pages_golden():
    g = load_golden("scatter_pages")
    table = torch.full_like(g["table_out"], -1).to(DEV)
    scatter_to_page_table(g["add_pages"].to(DEV), g["new_phys"].to(DEV), g["curr_pages"].to(DEV),
                          table[:, g["slot"]], g["max_pages"])
    assert torch.equal(table.cpu(), g["table_out"])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,HKV,D,PS", [(1, 8, 128, 128), (5, 8, 128, 128), (3, 2, 64, 256), (8, 4, 32, 32)])
def test_decode_store_vs_oracle(dtype, B, HKV, D, PS):
    g = torch.Generator().manual_seed(B * 131 + D)
    lens = torch.randint(0, 3 * PS, (B, HKV), generator=g)
    lens[0, 0] = PS - 1  # page-crossing on the next append
    lens[-1, -1] = PS
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, lens.tolist(), dtype, DEV, seed=7, extra_pages=2)
    bm[B // 2] = 0  # a CUDA-graph padding row: must be skipped
    kv = torch.randn((B, 2 * HKV * D), generator=g).to(dtype).to(DEV)
    key, val = kv[:, : HKV * D].view(B, HKV, D), kv[:, HKV * D:].view(B, HKV, D)  # strided
    bl = lens.to(torch.int32).to(DEV)
    kc0, vc0, bl0 = kc.cpu().clone(), vc.cpu().clone(), bl.cpu().clone()
    decode_store_kv(key=key, value=val, batch_mapping=bm, bh_lens=bl, page_table=pt, k_cache=kc,
                    v_cache=vc, PAGE_SIZE=PS)
    ost.decode_store_kv(key.cpu(), val.cpu(), bm.cpu(), bl0, pt.cpu(), kc0, vc0, PS)
    assert torch.equal(bl.cpu(), bl0)
    assert torch.equal(_bits(kc.cpu()), _bits(kc0)) and torch.equal(_bits(vc.cpu()), _bits(vc0))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens,HKV,D,PS", [([1], 8, 128, 128), ([300, 0, 129, 4000], 8, 128, 128),
                                           ([77, 513], 2, 64, 256), ([10, 20, 30], 4, 32, 32)])
def test_store_all_vs_oracle(dtype, lens, HKV, D, PS):
    B = len(lens)
    g = torch.Generator().manual_seed(sum(lens) + D)
    prefix = torch.randint(0, 2 * PS, (B, HKV), generator=g)
    tot = (prefix + torch.tensor(lens)[:, None]).tolist()
    kc, vc, pt, bm = make_paged(B, HKV, D, PS, tot, dtype, DEV, seed=3)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    Ntok = int(cu[-1])
    kv = torch.randn((Ntok, 3 * HKV * D), generator=g).to(dtype).to(DEV)
    keys, vals = kv[:, HKV * D:2 * HKV * D].view(Ntok, HKV, D), kv[:, 2 * HKV * D:].view(Ntok, HKV, D)
    bl = prefix.to(torch.int32).to(DEV)
    kc0, vc0, bl0 = kc.cpu().clone(), vc.cpu().clone(), bl.cpu().clone()
    prefill_store_all_kv(new_keys=keys, new_values=vals, cu_seqlens_k=cu.to(DEV), max_seqlen_k=max(lens),
                         k_cache=kc, v_cache=vc, page_table=pt, bh_lens=bl, batch_mapping=bm, PAGE_SIZE=PS)
    ost.prefill_store_all_kv(keys.cpu(), vals.cpu(), cu, kc0, vc0, pt.cpu(), bl0, bm.cpu(), PS)
    assert torch.equal(bl.cpu(), bl0)
    assert torch.equal(_bits(kc.cpu()), _bits(kc0)) and torch.equal(_bits(vc.cpu()), _bits(vc0))


def test_scatter_pages_vs_oracle_large():
    g = torch.Generator().manual_seed(5)
    L, H, P = 36, 8, 320
    add = torch.randint(0, 5, (L, H), generator=g).to(torch.int32)
    curr = torch.randint(0, P - 2, (L, H), generator=g).to(torch.int32)
    new_phys = torch.randperm(int(add.sum()) + 10, generator=g)[: int(add.sum())].to(torch.int32)
    table = torch.full((L, 3, H, P), -1, dtype=torch.int32)
    ref = table.clone()
    ost.scatter_to_page_table(add, new_phys, curr, ref[:, 1], P)
    dev = table.to(DEV)
    scatter_to_page_table(add.to(DEV), new_phys.to(DEV), curr.to(DEV), dev[:, 1], P)
    assert torch.equal(dev.cpu(), ref)
