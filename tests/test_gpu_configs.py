"""BASELINE.json's configurations 2-5 at their FULL sizes through the reference-shaped API
(plugin scoring -> Attention.forward -> decode), checked by what stays affordable at that size:

* attention: a device-side fp32 restatement on SAMPLED query rows (the CPU oracle would take minutes);
  helpers.torch_prefill_fp32 — the same math — is pinned to the oracle at small sizes in test_gpu_prefill.py
* eviction: kept sets / bh_lens / stored rows bit-exact against the oracle's selection fed with the
  device's scores (the contract: identical scores -> identical cache)
* size-independent properties: protected tokens always kept, every head a page multiple, budget respected
"""
import math

import numpy as np
import pytest
import torch

from compactor_vllm.compression.compactor import CompactorCompression
from compactor_vllm.compression.compression_config import CompressionMethod
from compactor_vllm.compression.snapkv import SnapKVCompression
from compactor_vllm.kv_cache.page_table import KVAllocationStatus, PagedKVCache
from compactor_vllm.layers.attention import Attention
from compactor_vllm.utils import arguments as A
from compactor_vllm.utils.context import CompressionContext, get_context, set_context
from oracle import selection as osel
from oracle import store as ost

from helpers import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
D, PS = 128, 128


def _qkv(n, HQ, HKV, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    qkv = torch.randn((n, (HQ + 2 * HKV) * D), generator=g, device=DEV, dtype=torch.float32)
    pos = torch.linspace(0, 40.0, n, device=DEV)[:, None, None]
    low = 0.5 * torch.sin(pos + torch.arange(D, device=DEV) * 0.1) + 0.2 * torch.randn((1, HKV, D), generator=g, device=DEV)
    k_pre = (low + torch.randn((n, HKV, D), generator=g, device=DEV)).to(torch.bfloat16)
    qkv = qkv.to(torch.bfloat16)
    qkv[:, HQ * D:(HQ + HKV) * D].view(n, HKV, D).add_((0.3 * low).to(torch.bfloat16))
    return (qkv[:, : HQ * D].view(n, HQ, D), qkv[:, HQ * D:(HQ + HKV) * D].view(n, HKV, D),
            qkv[:, (HQ + HKV) * D:].view(n, HKV, D), k_pre)


class Engine:
    """What the reference's KVCacheManager + ModelRunner set up around one attention layer."""

    def __init__(self, HQ, HKV, lens, extra_tokens=PS):
        self.HQ, self.HKV, self.lens = HQ, HKV, list(lens)
        n_lp = (max(lens) + extra_tokens) // PS + 2
        B = len(lens)
        self.cache = PagedKVCache(1, n_lp, B * HKV * n_lp + 8, PS, HKV, D, B + 1, torch.bfloat16, device=DEV)
        self.slots = [self.cache.new_batch() for _ in lens]
        for s, L in zip(self.slots, lens):
            assert self.cache.reserve_tokens(s, L + extra_tokens) == KVAllocationStatus.SUCCESS
        self.attn = Attention(HQ, D, 1.0 / math.sqrt(D), HKV)
        self.attn.k_cache, self.attn.v_cache, self.attn.page_table, self.attn.bh_seq_lens = self.cache.layer_slices(0)
        self.attn.page_size = PS
        self.bm = torch.tensor(self.slots, dtype=torch.int32, device=DEV)
        self.cu = torch.tensor([0] + np.cumsum(lens).tolist(), dtype=torch.int32, device=DEV)
        self.side = torch.cuda.Stream(priority=-1)

    def prefill(self, q, k, v, k_pre=None, method=None, ratio=1.0, first=0, last=0):
        B = len(self.lens)
        cctx = None
        if method is not None:
            pf, pl = [first] * B, [last] * B
            retain = A.tokens_to_retain(self.lens, [ratio] * B, pf, pl, self.HKV)
            cctx = CompressionContext(method, 512, torch.tensor(retain, dtype=torch.int32, device=DEV),
                                      A.max_tokens_to_retain(self.lens, self.HKV), self.lens,
                                      A.make_phi(D, 48, torch.bfloat16, DEV), pf, pl)
            self.retain = retain
        set_context(is_prefill=True, do_compression=method is not None, cu_seqlens_q=self.cu, cu_seqlens_k=self.cu,
                    max_seqlen_q=max(self.lens), max_seqlen_k=max(self.lens), batch_mapping=self.bm, max_bh_len=0,
                    compression_context=cctx, STORE_STREAM=self.side)
        ctx = get_context()
        scores = None
        if method is not None:
            plugin = CompactorCompression if method == CompressionMethod.COMPACTOR else SnapKVCompression
            scores = plugin.pre_rope_scoring(q, k_pre, v, ctx)
            scores = plugin.post_rope_scoring(q, k, v, scores, ctx)
        out = self.attn(q, k, v, scores)
        torch.cuda.current_stream().wait_stream(self.side)
        torch.cuda.synchronize()
        return out, scores

    def decode(self, q1, k1, v1):
        set_context(is_prefill=False, batch_mapping=self.bm)
        out = self.attn(q1, k1, v1)
        torch.cuda.synchronize()
        return out

    def head_rows(self, b, h, L):
        pos = torch.arange(L, device=DEV)
        return self.attn.page_table[self.slots[b], h, pos // PS].long() * PS + pos % PS


@torch.no_grad()
def _ref_rows(q, k, v, rows, HKV, scale, prefix=None):
    """fp32 causal attention of the sampled query positions `rows` of ONE sequence (no cached prefix unless
    `prefix` = (Kc [HKV][Lc,D], Vc) lists are given): [len(rows), HQ, D]."""
    N, HQ, _ = q.shape
    G = HQ // HKV
    R = torch.as_tensor(rows, device=q.device)
    out = torch.empty((len(rows), HQ, D), dtype=torch.float32, device=q.device)
    for h in range(HKV):
        K, V = k[:, h].float(), v[:, h].float()
        Lc = 0
        if prefix is not None:
            Lc = prefix[0][h].shape[0]
            K = torch.cat([prefix[0][h].float(), K], 0)
            V = torch.cat([prefix[1][h].float(), V], 0)
        S = torch.einsum("rgd,nd->rgn", q[R, h * G:(h + 1) * G].float(), K) * scale
        mask = torch.arange(K.shape[0], device=q.device)[None, :] > (R + Lc)[:, None]
        S.masked_fill_(mask[:, None, :], float("-inf"))
        out[:, h * G:(h + 1) * G] = torch.einsum("rgn,nd->rgd", torch.softmax(S, -1), V)
    return out


@torch.no_grad()
def _ref_decode(eng, q1, b, lens_bh):
    """fp32 decode attention of sequence b over its (compacted) cache rows."""
    HQ, HKV = eng.HQ, eng.HKV
    G = HQ // HKV
    out = torch.empty((HQ, D), dtype=torch.float32, device=DEV)
    for h in range(HKV):
        rows = eng.head_rows(b, h, int(lens_bh[b, h]))
        K, V = eng.attn.k_cache[rows].float(), eng.attn.v_cache[rows].float()
        S = (q1[b, h * G:(h + 1) * G].float() @ K.T) / math.sqrt(D)
        out[h * G:(h + 1) * G] = torch.softmax(S, -1) @ V
    return out


def _check_eviction(eng, scores, k, v, first, last):
    """Cache contents after the compressed prefill == the oracle's selection fed with the device's scores."""
    HKV = eng.HKV
    sc = scores.float().cpu()
    cu = eng.cu.cpu()
    idx = osel.rank_scores(sc, cu, max(eng.lens), max(eng.lens) * HKV, HKV).numpy()
    dev_lens = eng.cache.bh_seq_lens[0].index_select(0, eng.bm).cpu()
    for b, L in enumerate(eng.lens):
        s = int(cu[b])
        kept = osel.kept_positions_from_ranking(idx[b], eng.retain[b], HKV, s, L, np.zeros(HKV, np.int64), PS, True)
        assert sum(len(x) for x in kept) >= eng.retain[b]
        for h in range(HKV):
            n = len(kept[h])
            assert int(dev_lens[b, h]) == n, (b, h, int(dev_lens[b, h]), n)
            assert n % PS == 0 or n == L, "every head is padded to a page multiple (or holds the whole sequence)"
            t = torch.tensor(kept[h], dtype=torch.long, device=DEV)
            loc = t - s
            assert bool(((loc < first).sum() == min(first, L))) and bool(((loc >= L - last).sum() == min(last, L))), \
                "protected first/last tokens must survive in every head"
            rows = eng.head_rows(b, h, n)
            assert torch.equal(eng.attn.k_cache[rows], k[t, h]) and torch.equal(eng.attn.v_cache[rows], v[t, h])
    return dev_lens


def test_config4_dense_128k_prefill():
    """Llama-3.1-8B geometry, one 131072-token sequence, no compression: store-all + attention (140.7 TFLOP)."""
    HQ, HKV, N = 32, 8, 131072
    q, k, v, _ = _qkv(N, HQ, HKV, 4)
    eng = Engine(HQ, HKV, [N])
    out, _ = eng.prefill(q, k, v)
    rows = [0, 1, 127, 128, 129, 4095, 4096, 65535, 65536, 100000, 131070, 131071]
    ref = _ref_rows(q, k, v, rows, HKV, 1.0 / math.sqrt(D))
    assert rel_err(out[rows], ref) < 2e-2
    assert (out[rows].float() - ref).abs().max() < 2e-2
    lens = eng.cache.bh_seq_lens[0, eng.slots[0]].cpu()
    assert (lens == N).all()
    for h in (0, HKV - 1):  # store-all is a bit-exact scatter through the page table
        r = eng.head_rows(0, h, N)
        assert torch.equal(eng.attn.k_cache[r], k[:, h]) and torch.equal(eng.attn.v_cache[r], v[:, h])


def _decode_steps(eng, lens_bh, n_steps, seed, check_every):
    """Append + attend n_steps times; compare sampled steps with the fp32 restatement over the cache."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    B = len(eng.lens)
    lens_bh = lens_bh.clone()
    for step in range(n_steps):
        q1 = torch.randn((B, eng.HQ, D), generator=g, device=DEV).to(torch.bfloat16)
        k1 = torch.randn((B, eng.HKV, D), generator=g, device=DEV).to(torch.bfloat16)
        v1 = torch.randn((B, eng.HKV, D), generator=g, device=DEV).to(torch.bfloat16)
        out = eng.decode(q1, k1, v1)
        lens_bh += 1
        assert torch.equal(eng.cache.bh_seq_lens[0].index_select(0, eng.bm).cpu(), lens_bh)
        if step % check_every == 0 or step == n_steps - 1:
            for b in sorted({0, B - 1}):
                ref = _ref_decode(eng, q1, b, lens_bh)
                assert rel_err(out[b], ref) < 2e-2, (step, b)
                h = eng.HKV - 1  # the appended row sits at the end of the head's page list
                last = eng.head_rows(b, h, int(lens_bh[b, h]))[-1]
                assert torch.equal(eng.attn.k_cache[last], k1[b, h]) and torch.equal(eng.attn.v_cache[last], v1[b, h])
    return lens_bh


def test_config2_llama8b_16k_compactor_prefill_then_decode():
    """One layer of config 2: 16384-token prefill, COMPACTOR 50 % (first 16 / last 64), 256 decode steps."""
    HQ, HKV, N, first, last = 32, 8, 16384, 16, 64
    q, k, v, k_pre = _qkv(N, HQ, HKV, 2)
    eng = Engine(HQ, HKV, [N], extra_tokens=3 * PS)
    out, scores = eng.prefill(q, k, v, k_pre, CompressionMethod.COMPACTOR, 0.5, first, last)
    assert eng.retain == [65216]
    rows = [0, 63, 64, 1000, 8191, 8192, 16383]
    assert rel_err(out[rows], _ref_rows(q, k, v, rows, HKV, 1.0 / math.sqrt(D))) < 2e-2
    sc = scores.float().cpu()
    assert torch.isinf(sc[:first]).all() and torch.isinf(sc[N - last:]).all() and torch.isfinite(sc[first:N - last]).all()
    lens_bh = _check_eviction(eng, scores, k, v, first, last)
    assert 65216 <= int(lens_bh.sum()) <= 65216 + HKV * PS
    _decode_steps(eng, lens_bh, 256, seed=22, check_every=64)


def test_config3_qwen8b_snapkv_25pct_40k_batch16_decode():
    """Config 3: SnapKV 25 % at 40960 tokens (window 32), prefilled one sequence at a time into 16 slots,
    then batch-16 decode over the compacted, ragged caches."""
    HQ, HKV, N, B, w = 32, 8, 40960, 16, 32
    eng = Engine(HQ, HKV, [N] * B, extra_tokens=PS)
    all_bm, all_cu, all_lens = eng.bm, eng.cu, eng.lens
    lens_bh = torch.zeros((B, HKV), dtype=torch.int32)
    for b in range(B):
        q, k, v, _ = _qkv(N, HQ, HKV, 100 + b)
        eng.bm, eng.lens = all_bm[b:b + 1], [N]
        eng.cu = torch.tensor([0, N], dtype=torch.int32, device=DEV)
        slots, eng.slots = eng.slots, [eng.slots[b]]
        # the engine's default protected counts (16 / 64) size the budget (utils/arguments.py:109-121 ->
        # SURVEY's 81760 pairs); SnapKV itself force-keeps only its 32-key observation window (snapkv.py:27-35)
        out, scores = eng.prefill(q, k, v, None, CompressionMethod.SNAPKV, 0.25, 16, 64)
        assert eng.retain == [81760]
        if b in (0, B - 1):  # full eviction check on two sequences (a few seconds of oracle each)
            rows = [0, 4096, 20479, 40959]
            assert rel_err(out[rows], _ref_rows(q, k, v, rows, HKV, 1.0 / math.sqrt(D))) < 2e-2
            lens_bh[b] = _check_eviction(eng, scores, k, v, 0, w)[0]
        else:
            lens_bh[b] = eng.cache.bh_seq_lens[0, eng.slots[0]].cpu()
        eng.slots = slots
        del q, k, v, out, scores
    eng.bm, eng.cu, eng.lens = all_bm, all_cu, all_lens
    assert (lens_bh % PS == 0).all() and int(lens_bh.sum()) >= B * 81760
    _decode_steps(eng, lens_bh, 4, seed=33, check_every=3)


@pytest.mark.parametrize("tp", [2, 4, 8])
def test_config5_llama70b_head_sharded_rank(tp):
    """Config 5, one rank of tp: Llama-3.1-70B has HQ=64, HKV=8 -> (32,4)/(16,2)/(8,1) per rank, G=8; COMPACTOR 50 %
    at 32768 with the retain budget computed from the LOCAL head count (utils/arguments.py:109-121)."""
    HQ, HKV = 64 // tp, 8 // tp
    N, first, last = 32768, 16, 64
    q, k, v, k_pre = _qkv(N, HQ, HKV, 50 + tp)
    eng = Engine(HQ, HKV, [N], extra_tokens=PS)
    out, scores = eng.prefill(q, k, v, k_pre, CompressionMethod.COMPACTOR, 0.5, first, last)
    assert eng.retain == [int(round(0.5 * (N - first - last) * HKV))]
    rows = [0, 127, 128, 16384, 32767]
    assert rel_err(out[rows], _ref_rows(q, k, v, rows, HKV, 1.0 / math.sqrt(D))) < 2e-2
    lens_bh = _check_eviction(eng, scores, k, v, first, last)
    _decode_steps(eng, lens_bh, 3, seed=tp, check_every=1)
