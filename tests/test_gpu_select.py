"""Ranking, top-k eviction and compaction vs the CPU oracle and the reference-generated goldens.

Given identical scores the kept sets, bh_lens and (because both sides write a head's rows in
ascending token order) the cache bytes must be bit-exact."""
import collections

import numpy as np
import pytest
import torch

from compactor_vllm.compression.common import extract_and_store_top_kv, scores_to_retain_indices
from compactor_vllm.kv_cache.store_kv_cache import prefill_store_topk_kv
from oracle import selection as osel
from oracle import store as ost

from conftest import load_golden
from helpers import make_paged

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(t):
    return t.view(torch.int16) if t.element_size() == 2 else t


def _scores(N, H, seed, ties=False):
    g = torch.Generator().manual_seed(seed)
    s = torch.randn((N, H), generator=g)
    if ties:
        s = (s * 4).round() / 4  # many exact ties
        s[::7] = float("inf")
        s[3::11] = float("-inf")
        s[5::13, 0] = float("nan")
        s[1, :] = 0.0
        s[2, :] = -0.0
    return s


@pytest.mark.parametrize("lens,H,top_frac", [([1000], 8, 1.0), ([10, 257, 64], 4, 1.0), ([70, 30], 2, 0.3),
                                             ([4097], 8, 1.0), ([1, 1, 1], 8, 1.0)])
@pytest.mark.parametrize("ties", [False, True])
def test_rank_matches_oracle(lens, H, top_frac, ties):
    cu = torch.tensor([0] + np.cumsum(lens).tolist(), dtype=torch.int32)
    N = int(cu[-1])
    sc = _scores(N, H, seed=N + H, ties=ties)
    max_len = max(lens)
    top_k = max(1, int(max_len * H * top_frac))
    ref = osel.rank_scores(sc, cu, max_len, top_k, H)
    out = scores_to_retain_indices(sc.to(DEV), cu.to(DEV), max_len, top_k, H)
    assert out.dtype == torch.int64 and out.shape == ref.shape
    assert torch.equal(out.cpu(), ref)


def _setup_store(lens, H, D, PS, dtype, seed, with_prefix):
    B = len(lens)
    g = torch.Generator().manual_seed(seed)
    prefix = torch.randint(0, 2 * PS, (B, H), generator=g) if with_prefix else torch.zeros((B, H), dtype=torch.int64)
    tot = (prefix + torch.tensor(lens)[:, None]).tolist()
    kc, vc, pt, bm = make_paged(B, H, D, PS, tot, dtype, DEV, seed=seed + 1)
    cu = torch.tensor([0] + np.cumsum(lens).tolist(), dtype=torch.int32)
    N = int(cu[-1])
    kv = torch.randn((N, 2 * H * D), generator=g).to(dtype).to(DEV)
    keys, vals = kv[:, : H * D].view(N, H, D), kv[:, H * D:].view(N, H, D)
    return prefix.to(torch.int32), kc, vc, pt, bm, cu, keys, vals


def _compare_exact(dev, ref):
    (kc, vc, bl), (kc0, vc0, bl0) = dev, ref
    assert torch.equal(bl.cpu(), bl0)
    assert torch.equal(_bits(kc.cpu()), _bits(kc0))
    assert torch.equal(_bits(vc.cpu()), _bits(vc0))


@pytest.mark.parametrize("lens,H,D,PS", [([1000], 8, 128, 128), ([10, 300, 64], 4, 64, 128), ([33, 70], 2, 32, 256),
                                         ([129, 1, 500, 7], 8, 128, 128)])
@pytest.mark.parametrize("frac", [0.1, 0.5, 0.95])
@pytest.mark.parametrize("pad,with_prefix,ties", [(True, False, False), (True, True, True), (False, False, True)])
def test_store_topk_both_flavours_vs_oracle(lens, H, D, PS, frac, pad, with_prefix, ties):
    dtype = torch.float16
    prefix, kc, vc, pt, bm, cu, keys, vals = _setup_store(lens, H, D, PS, dtype, 17, with_prefix)
    N = int(cu[-1])
    sc = _scores(N, H, seed=N, ties=ties)
    for b, L in enumerate(lens):  # protected tokens, as the scorers produce them
        sc[int(cu[b]): int(cu[b]) + min(2, L)] = float("inf")
    max_len = max(lens)
    retain = torch.tensor([max(1, int(round(frac * L * H))) for L in lens], dtype=torch.int32)
    if len(lens) > 2:
        bm[1] = 0  # reserved slot: the whole sequence is skipped
    # oracle
    kc0, vc0, bl0 = kc.cpu().clone(), vc.cpu().clone(), prefix.clone()
    osel.extract_and_store_top_kv(sc, cu, max_len, max_len * H, H, keys.cpu(), vals.cpu(), retain, pt.cpu(),
                                  bm.cpu(), bl0, kc0, vc0, PS, pad=pad)
    # index flavour: ranking from the device sort, then the reference-shaped store call
    kc1, vc1, bl1 = kc.clone(), vc.clone(), prefix.clone().to(DEV)
    idx = scores_to_retain_indices(sc.to(DEV), cu.to(DEV), max_len, max_len * H, H)
    prefill_store_topk_kv(new_keys=keys, new_vals=vals, indices_topk=idx, num_tokens_to_retain=retain.to(DEV),
                          page_table=pt, batch_mapping=bm, bh_lens=bl1, k_cache=kc1, v_cache=vc1, PAGE_SIZE=PS,
                          PAD_TO_PAGE_SIZE=pad, cu_seqlens_k=cu.to(DEV))
    _compare_exact((kc1, vc1, bl1), (kc0, vc0, bl0))
    # fused flavour (what Attention.forward calls)
    kc2, vc2, bl2 = kc.clone(), vc.clone(), prefix.clone().to(DEV)
    extract_and_store_top_kv(sc.to(DEV), cu.to(DEV), max_len, max_len * H, H, keys, vals, retain.to(DEV), pt, bm,
                             bl2, kc2, vc2, PS, PAD_TO_PAGE_SIZE=pad)
    _compare_exact((kc2, vc2, bl2), (kc0, vc0, bl0))


# the reference's own test shape: tests/test_store_kv.py:27-173 (no padding, reserved batch = -1,
# ranking truncated to TOP_K, identity page layout), unordered multiset check included
@pytest.mark.parametrize("B", [1, 3, 8])
@pytest.mark.parametrize("frac", [0.1, 0.4])
@pytest.mark.parametrize("H,D", [(2, 32), (4, 64), (8, 128)])
@pytest.mark.parametrize("L", [10, 70, 1000])
@pytest.mark.parametrize("PS", [128, 256])
def test_reference_store_test_shape(B, frac, H, D, L, PS):
    dtype = torch.float16
    TOP_K = int(L * H * frac)
    cu = torch.arange(B + 1, dtype=torch.int32) * L
    N = B * L
    g = torch.Generator().manual_seed(B + H + L)
    keys = torch.randn((N, H, D), generator=g).to(dtype).to(DEV)
    vals = torch.randn((N, H, D), generator=g).to(dtype).to(DEV)
    sc = torch.randn((N, H), generator=g).to(DEV)
    top_k_eff = max(0, min(TOP_K, L * H))
    idx = scores_to_retain_indices(sc, cu.to(DEV), L, top_k_eff, H)
    LP = max(1, (top_k_eff + PS - 1) // PS)
    n_pages = B * H * LP + 32
    kc = torch.zeros((n_pages * PS, D), dtype=dtype, device=DEV)
    vc = torch.zeros_like(kc)
    pt = torch.arange(B * H * LP, dtype=torch.int32, device=DEV).view(B, H, LP).contiguous()
    bl = torch.zeros((B, H), dtype=torch.int32, device=DEV)
    bm = torch.arange(B, dtype=torch.int32, device=DEV)  # identity, slot 0 in use -> reserved = -1
    prefill_store_topk_kv(new_keys=keys, new_vals=vals, indices_topk=idx,
                          num_tokens_to_retain=torch.full((B,), top_k_eff, dtype=torch.int32, device=DEV),
                          page_table=pt, batch_mapping=bm, bh_lens=bl, PAGE_SIZE=PS, k_cache=kc, v_cache=vc,
                          PAD_TO_PAGE_SIZE=False, TRITON_RESERVED_BATCH=-1, cu_seqlens_k=cu.to(DEV))
    idx_c, bl_c, kc_c, vc_c, keys_c, vals_c = idx.cpu(), bl.cpu(), kc.cpu(), vc.cpu(), keys.cpu(), vals.cpu()
    for b in range(B):
        heads = (idx_c[b] % H).tolist()
        toks = (idx_c[b] // H).tolist()
        cnt = collections.Counter(heads)
        for h in range(H):
            Lh = int(bl_c[b, h])
            assert Lh == cnt.get(h, 0)
            if Lh == 0:
                continue
            sel = [t for t, hh in zip(toks, heads) if hh == h]
            rows = torch.from_numpy(ost.head_rows(pt.cpu(), b, h, 0, Lh, PS))
            exp = collections.Counter(zip(map(bytes, _bits(keys_c[sel, h]).numpy()), map(bytes, _bits(vals_c[sel, h]).numpy())))
            act = collections.Counter(zip(map(bytes, _bits(kc_c[rows]).numpy()), map(bytes, _bits(vc_c[rows]).numpy())))
            assert exp == act


@pytest.mark.parametrize("name,pad,reserved", [("store_topk_pad", True, None), ("store_topk_nopad", False, -1)])
def test_store_topk_golden(name, pad, reserved):
    """Against what the reference's Triton K4/K5 wrote (multiset per head: its slot order is atomic)."""
    g = load_golden(name)
    PS, H = g["PS"], g["keys"].shape[1]
    kc, vc = g["k_cache_in"].to(DEV), g["v_cache_in"].to(DEV)
    bl = (g["bh_lens_in"] if "bh_lens_in" in g else torch.zeros_like(g["bh_lens_out"])).to(DEV)
    prefill_store_topk_kv(new_keys=g["keys"].to(DEV), new_vals=g["vals"].to(DEV), indices_topk=g["indices"].to(DEV),
                          num_tokens_to_retain=g["num_retain"].to(DEV), page_table=g["page_table"].to(DEV),
                          batch_mapping=g["batch_mapping"].to(DEV), bh_lens=bl, k_cache=kc, v_cache=vc,
                          PAGE_SIZE=PS, PAD_TO_PAGE_SIZE=pad, cu_seqlens_k=g["cu_seqlens_k"].to(DEV),
                          TRITON_RESERVED_BATCH=reserved)
    assert torch.equal(bl.cpu(), g["bh_lens_out"])
    kc, vc = kc.cpu(), vc.cpu()
    for b in range(bl.shape[0]):
        bt = int(g["batch_mapping"][b])
        for h in range(H):
            rows = torch.from_numpy(ost.head_rows(g["page_table"], bt, h, 0, int(bl[b, h]), PS))
            mine = collections.Counter(zip(map(bytes, _bits(kc[rows]).numpy()), map(bytes, _bits(vc[rows]).numpy())))
            ref = collections.Counter(zip(map(bytes, _bits(g["k_cache_out"][rows]).numpy()),
                                          map(bytes, _bits(g["v_cache_out"][rows]).numpy())))
            assert mine == ref, (b, h)
    if pad:  # and the fused path from the same scores
        kc2, vc2, bl2 = g["k_cache_in"].to(DEV), g["v_cache_in"].to(DEV), g["bh_lens_in"].to(DEV)
        cu = g["cu_seqlens_k"]
        max_len = int((cu[1:] - cu[:-1]).max())
        extract_and_store_top_kv(g["scores"].to(DEV), cu.to(DEV), max_len, max_len * H, H, g["keys"].to(DEV),
                                 g["vals"].to(DEV), g["num_retain"].to(DEV), g["page_table"].to(DEV),
                                 g["batch_mapping"].to(DEV), bl2, kc2, vc2, PS)
        assert torch.equal(bl2.cpu(), g["bh_lens_out"])
        assert torch.equal(_bits(kc2.cpu()), _bits(kc)) and torch.equal(_bits(vc2.cpu()), _bits(vc))
