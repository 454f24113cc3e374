"""Generate tests/golden/*.npz by running the REFERENCE's own kernels on CPU.

Run in the build container (needs /root/reference, torch, triton):

    python oracle/gen_golden.py

The reference's Triton kernels execute under TRITON_INTERPRET=1 (set before
triton is imported); autotuned kernels are pinned to one config because the
autotuner would otherwise try to benchmark on a CUDA device.  Nothing here is
imported by the product, and nothing under tests/ needs /root/reference at run
time — only the vectors this script wrote.

Inputs are stored next to the outputs (16-bit floats as raw uint16 bit
patterns + a dtype tag), so the fixtures do not depend on RNG stability.

bfloat16: the Triton interpreter computes through numpy, which has no bf16, and
returns garbage for bf16 tensors (checked here: outputs ~1e8 / NaN).  Fixtures
named *_bf16 therefore hold bf16-representable inputs but run the reference on
their exact fp32 up-casts (`ref_in`), so they pin the reference's math, masks
and indexing at bf16 input precision; the reference's P->bf16 rounding before
P.V is not reproduced and is covered by the parity tolerance.  float16 runs
natively.
"""
from __future__ import annotations

import os
import sys

os.environ["TRITON_INTERPRET"] = "1"
REF_SRC = os.environ.get("CVLLM_REFERENCE_SRC", "/root/reference/src")
sys.path.insert(0, REF_SRC)
sys.path.insert(1, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root: oracle.gen_golden_script

from pathlib import Path  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402
import triton  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


# --------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------
def pack(t: torch.Tensor):
    """torch tensor -> (ndarray, tag) that np.savez can hold."""
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16), "bf16"
    if t.dtype == torch.float16:
        return t.view(torch.int16).numpy().view(np.uint16), "f16"
    return t.numpy(), str(t.dtype).replace("torch.", "")


def save(name: str, **tensors):
    arrays = {}
    for key, val in tensors.items():
        if isinstance(val, torch.Tensor):
            arr, tag = pack(val.detach().cpu().contiguous())
            arrays[key] = arr
            arrays[key + "__dtype"] = np.array(tag)
        else:
            arrays[key] = np.asarray(val)
    OUT.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(OUT / f"{name}.npz", **arrays)
    print(f"wrote {name}.npz ({(OUT / f'{name}.npz').stat().st_size / 1024:.1f} KiB)")


def ref_in(t: torch.Tensor) -> torch.Tensor:
    """What the reference kernel is fed: bf16 -> exact fp32 up-cast (see module docstring)."""
    return t.float() if t.dtype == torch.bfloat16 else t


def pin(kernel, **cfg_kwargs):
    """Pin a triton autotuner to a single config (meta params + launch params)."""
    launch = {k: cfg_kwargs.pop(k) for k in ("num_warps", "num_stages") if k in cfg_kwargs}
    kernel.configs = [triton.Config(cfg_kwargs, **launch)]


def build_cache(B, HKV, D, PS, NLP, lens_bh, dtype, gen, n_slots):
    """Random page permutation, non-identity slots; returns caches, page table, mapping.

    Only the rows a test can legitimately read (the first lens_bh[b][h] positions of each mapped
    head) are random; the rest stay zero so the fixture compresses.  (Garbage / NaN in the unused
    rows is exercised by the GPU tests against the oracle, not by the goldens.)"""
    n_pages = n_slots * HKV * NLP
    perm = torch.randperm(n_pages, generator=gen).to(torch.int32)
    page_table = perm.view(n_slots, HKV, NLP).contiguous()
    k_cache = torch.zeros((n_pages * PS, D), dtype=dtype)
    v_cache = torch.zeros((n_pages * PS, D), dtype=dtype)
    batch_mapping = (torch.randperm(n_slots - 1, generator=gen)[:B] + 1).to(torch.int32)
    for b in range(B):
        for h in range(HKV):
            L = int(lens_bh[b][h])
            pos = torch.arange(L)
            rows = page_table[int(batch_mapping[b]), h, pos // PS].long() * PS + pos % PS
            k_cache[rows] = torch.randn((L, D), generator=gen).to(dtype)
            v_cache[rows] = torch.randn((L, D), generator=gen).to(dtype)
    return k_cache, v_cache, page_table, batch_mapping


# --------------------------------------------------------------------------
# attention
# --------------------------------------------------------------------------
def gen_prefill():
    from compactor_vllm.attention import sparse_varlen_kernel as svk

    pin(svk._causal_head_sparse_varlen_with_cache, BLOCK_N=32, BLOCK_M=16, WARPSPEC=False,
        num_warps=4, num_stages=2)
    for name, dtype, HQ, HKV, D, PS, appends, lens in [
        ("prefill_bf16", torch.bfloat16, 8, 2, 128, 128, [37, 1, 150], [[0, 70], [129, 5], [256, 0]]),
        ("prefill_f16_ps256", torch.float16, 4, 2, 128, 256, [64, 3], [[300, 1], [0, 0]]),
    ]:
        gen = torch.Generator().manual_seed(1234)
        B = len(appends)
        NLP = 4
        k_cache, v_cache, pt, bm = build_cache(B, HKV, D, PS, NLP, lens, dtype, gen, n_slots=B + 2)
        seq_lens = torch.tensor(lens, dtype=torch.int32)
        cu = torch.tensor([0] + list(np.cumsum(appends)), dtype=torch.int32)
        N = int(cu[-1])
        # q/k/v as views into one fused qkv buffer, like the models hand them over
        qkv = torch.randn((N, (HQ + 2 * HKV) * D), generator=gen).to(dtype)
        q = qkv[:, : HQ * D].view(N, HQ, D)
        k = qkv[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D)
        v = qkv[:, (HQ + HKV) * D:].view(N, HKV, D)
        if dtype == torch.bfloat16:
            # keep the fused-buffer striding for the fp32 run as well
            qkv32 = qkv.float()
            rq = qkv32[:, : HQ * D].view(N, HQ, D)
            rk = qkv32[:, HQ * D:(HQ + HKV) * D].view(N, HKV, D)
            rv = qkv32[:, (HQ + HKV) * D:].view(N, HKV, D)
        else:
            rq, rk, rv = q, k, v
        out = svk.causal_sparse_varlen_with_cache(
            rq, rk, rv, ref_in(k_cache), ref_in(v_cache), seq_lens, pt, bm, cu, max(appends),
            int(seq_lens.max()), HKV, PS, None)
        save(name, q=q, k=k, v=v, k_cache=k_cache, v_cache=v_cache, seq_lens_bh=seq_lens,
             page_table=pt, batch_mapping=bm, cu_seqlens_q=cu, HKV=HKV, PS=PS, out=out)


def gen_decode():
    from compactor_vllm.attention import sparse_decode_kernel as sdk

    for name, dtype, HQ, HKV, D, PS, lens, key_split in [
        ("decode_bf16", torch.bfloat16, 8, 2, 128, 128, [[1, 300], [128, 0], [257, 64]], 3),
        ("decode_f16", torch.float16, 8, 2, 128, 64, [[70, 2]], 1),
    ]:
        gen = torch.Generator().manual_seed(4321)
        B = len(lens)
        NLP = 4
        k_cache, v_cache, pt, bm = build_cache(B, HKV, D, PS, NLP, lens, dtype, gen, n_slots=B + 2)
        seq_lens = torch.tensor(lens, dtype=torch.int32)
        q = torch.randn((B, HQ, D), generator=gen).to(dtype)
        scale = 1.0 / np.sqrt(D)
        triton.set_allocator(lambda size, align, _: torch.empty(size, dtype=torch.int8))
        rdtype = torch.float32 if dtype == torch.bfloat16 else dtype
        mid_o = torch.zeros((B, key_split, HQ, D), dtype=rdtype)
        mid_lse = torch.full((B, key_split, HQ), -float("inf"), dtype=torch.float32)
        G = HQ // HKV
        tl_dtype = triton.language.float32 if dtype == torch.bfloat16 else triton.language.float16
        sdk._varkv_stage1_groupM.fn[(B, HKV, key_split)](
            q=ref_in(q), k=ref_in(k_cache), v=ref_in(v_cache), mid_o=mid_o, mid_lse=mid_lse, page_table_bhl=pt,
            batch_mapping=bm, seq_lens_bh=seq_lens, SM_SCALE=scale, B=B, HKV=HKV, HQ=HQ,
            CACHE_SIZE=k_cache.shape[0], STRIDE_LBS=mid_lse.stride(0), STRIDE_LS=mid_lse.stride(1),
            STRIDE_LH=mid_lse.stride(2), N_LOGICAL_PAGES_MAX=NLP, D=D, KEY_SPLIT=key_split,
            GROUP_M=G, DTYPE=tl_dtype, BLOCK_N=32, MIN_BLOCK_KV=8, WARPSPEC=False, PAGE_SIZE=PS)
        if key_split == 1:
            out = mid_o.squeeze(1).contiguous()
        else:
            out = torch.zeros((B, HQ, D), dtype=rdtype)
            sdk._varkv_stage2_reduce[(B, HQ)](
                mid_o=mid_o, mid_lse=mid_lse, output=out, STRIDE_LBS=mid_lse.stride(0),
                STRIDE_LS=mid_lse.stride(1), STRIDE_LH=mid_lse.stride(2), STRIDE_OBS=out.stride(0),
                STRIDE_OH=out.stride(1), B=B, HQ=HQ, D=D, KEY_SPLIT=key_split, DTYPE=tl_dtype)
        valid = (seq_lens > 0).repeat_interleave(G, dim=1)  # rows with L == 0 are garbage in the ref
        save(name, q=q, k_cache=k_cache, v_cache=v_cache, seq_lens_bh=seq_lens, page_table=pt,
             batch_mapping=bm, HKV=HKV, PS=PS, out=out, valid_rows=valid)


# --------------------------------------------------------------------------
# stores + ranking
# --------------------------------------------------------------------------
def gen_stores():
    from compactor_vllm.compression.common import scores_to_retain_indices
    from compactor_vllm.kv_cache import store_kv_cache as skc
    from compactor_vllm.kv_cache.write_page_table import _scatter_pages_kernel_lh

    dtype, H, D, PS, NLP = torch.float16, 4, 32, 16, 6
    gen = torch.Generator().manual_seed(99)
    lens = [37, 5, 64]
    B = len(lens)
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    N = int(cu[-1])
    n_slots = B + 2
    prefix = torch.tensor([[0, 3, 16, 1], [2, 0, 0, 15], [0, 0, 1, 17]], dtype=torch.int32)
    k_cache0, v_cache0, pt, bm = build_cache(B, H, D, PS, NLP, prefix.tolist(), dtype, gen, n_slots)
    kv = torch.randn((N, 2 * H * D), generator=gen).to(dtype)
    keys = kv[:, : H * D].view(N, H, D)
    vals = kv[:, H * D:].view(N, H, D)

    # decode_store_kv (one reserved row)
    kc, vc, bl = k_cache0.clone(), v_cache0.clone(), prefix.clone()
    bm_res = bm.clone()
    bm_res[1] = 0
    skc.decode_store_kv(key=keys[:B], value=vals[:B], batch_mapping=bm_res, bh_lens=bl, page_table=pt,
                        k_cache=kc, v_cache=vc, PAGE_SIZE=PS)
    save("decode_store", key=keys[:B].contiguous(), value=vals[:B].contiguous(), batch_mapping=bm_res,
         bh_lens_in=prefix, page_table=pt, k_cache_in=k_cache0, v_cache_in=v_cache0, PS=PS,
         bh_lens_out=bl, k_cache_out=kc, v_cache_out=vc)

    # prefill_store_all_kv
    kc, vc, bl = k_cache0.clone(), v_cache0.clone(), prefix.clone()
    skc.prefill_store_all_kv(new_keys=keys, new_values=vals, cu_seqlens_k=cu, max_seqlen_k=max(lens),
                             k_cache=kc, v_cache=vc, page_table=pt, bh_lens=bl, batch_mapping=bm,
                             PAGE_SIZE=PS)
    save("store_all", keys=keys, vals=vals, cu_seqlens_k=cu, batch_mapping=bm, bh_lens_in=prefix,
         page_table=pt, k_cache_in=k_cache0, v_cache_in=v_cache0, PS=PS, bh_lens_out=bl,
         k_cache_out=kc, v_cache_out=vc)

    # ranking + top-k store with padding (distinct scores -> tie-free, +inf protected rows)
    scores = torch.randn((N, H), generator=gen)
    scores[:2] = float("inf")
    scores[cu[1]:cu[1] + 1] = float("inf")
    retain = torch.tensor([40, 6, 100], dtype=torch.int32)  # > the 4 tied +inf entries of seq 1: no tie at the cut
    idx = scores_to_retain_indices(scores, cu, max(lens), max(lens) * H, H)
    kc, vc, bl = k_cache0.clone(), v_cache0.clone(), prefix.clone()
    skc.prefill_store_topk_kv(new_keys=keys, new_vals=vals, indices_topk=idx,
                              num_tokens_to_retain=retain, page_table=pt, batch_mapping=bm, bh_lens=bl,
                              k_cache=kc, v_cache=vc, PAGE_SIZE=PS, PAD_TO_PAGE_SIZE=True,
                              cu_seqlens_k=cu)
    save("store_topk_pad", keys=keys, vals=vals, scores=scores, cu_seqlens_k=cu, batch_mapping=bm,
         bh_lens_in=prefix, page_table=pt, k_cache_in=k_cache0, v_cache_in=v_cache0, PS=PS,
         num_retain=retain, indices=idx, bh_lens_out=bl, k_cache_out=kc, v_cache_out=vc)

    # same without padding, truncated ranking (the shape the reference's own test uses)
    top_k = 30
    idx2 = scores_to_retain_indices(scores, cu, max(lens), top_k, H)
    retain2 = torch.tensor([30, 12, 30], dtype=torch.int32)
    kc, vc, bl = k_cache0.clone(), v_cache0.clone(), torch.zeros_like(prefix)
    skc.prefill_store_topk_kv(new_keys=keys, new_vals=vals, indices_topk=idx2,
                              num_tokens_to_retain=retain2, page_table=pt, batch_mapping=bm, bh_lens=bl,
                              k_cache=kc, v_cache=vc, PAGE_SIZE=PS, PAD_TO_PAGE_SIZE=False,
                              TRITON_RESERVED_BATCH=-1)
    save("store_topk_nopad", keys=keys, vals=vals, scores=scores, cu_seqlens_k=cu, batch_mapping=bm,
         page_table=pt, k_cache_in=k_cache0, v_cache_in=v_cache0, PS=PS, num_retain=retain2,
         indices=idx2, top_k=top_k, bh_lens_out=bl, k_cache_out=kc, v_cache_out=vc)

    # scatter_to_page_table: launch the kernel directly (the wrapper allocates on "cuda")
    L, Hh, P = 3, 4, 6
    add = torch.tensor([[2, 0, 1, 3], [0, 0, 0, 0], [1, 1, 1, 1]], dtype=torch.int32)
    curr = torch.tensor([[0, 2, 5, 1], [1, 1, 1, 1], [0, 3, 2, 4]], dtype=torch.int32)
    new_phys = torch.arange(100, 100 + int(add.sum()), dtype=torch.int32)
    table = torch.full((L, 5, Hh, P), -1, dtype=torch.int32)
    view = table[:, 2]  # non-contiguous slice, like page_table[:, batch_index]
    cum = torch.zeros(L * Hh + 1, dtype=torch.int32)
    cum[1:] = torch.cumsum(add.view(-1), 0)
    s_l, s_h, s_p = view.stride()
    _scatter_pages_kernel_lh[(L, Hh)](add.view(-1), cum, new_phys, curr.view(-1), view, s_l, s_h, s_p,
                                      L=L, H=Hh, max_pages_per_head=P)
    save("scatter_pages", add_pages=add, curr_pages=curr, new_phys=new_phys, max_pages=P,
         table_out=table, slot=2)


# --------------------------------------------------------------------------
# scorers
# --------------------------------------------------------------------------
def _structured_qk(N, HQ, HKV, D, dtype, gen):
    """q/k with a shared low-rank component so attention is not uniform."""
    base = torch.randn((N, 1, D), generator=gen)
    q = (0.6 * torch.randn((N, HQ, D), generator=gen) + 0.5 * base).to(dtype)
    k = (0.6 * torch.randn((N, HKV, D), generator=gen) + 0.5 * base).to(dtype)
    return q, k


def gen_chunk_scores():
    from compactor_vllm.compression import compactor as cp

    dtype, HQ, HKV, D = torch.bfloat16, 8, 2, 128
    gen = torch.Generator().manual_seed(7)
    lens = [200, 65, 128]
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    N = int(cu[-1])
    q, k = _structured_qk(N, HQ, HKV, D, dtype, gen)
    pin(cp._zscore_per_batch_epilogue_no_window, BLOCK_K=64)
    raw = cp.non_causal_attn_scores(ref_in(q), ref_in(k), ref_in(k), cu, max(lens), 128, sm_scale=0.25,
                                    normalize=False)
    accum = torch.randn((N, HKV), generator=gen).to(dtype)
    full = cp.non_causal_attn_scores(
        ref_in(q), ref_in(k), ref_in(k), cu, max(lens), 128, sm_scale=0.25, normalize=True, context_lens=lens,
        protected_first_tokens=[4, 0, 16], protected_last_tokens=[8, 3, 0], accum_scores=ref_in(accum),
        accum_blending=0.5)
    save("chunk_scores", q=q, k=k, cu_seqlens=cu, chunk=128, sm_scale=0.25, raw=raw, accum=accum,
         prot_first=np.array([4, 0, 16]), prot_last=np.array([8, 3, 0]), full=full)


def gen_snapkv():
    from compactor_vllm.compression import snapkv as sk

    dtype, HQ, HKV, D, w = torch.bfloat16, 8, 2, 128, 8
    gen = torch.Generator().manual_seed(11)
    lens = [150, 8, 70]  # the second sequence has L <= w: untouched (uninitialised) in the reference
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    N = int(cu[-1])
    q, k = _structured_qk(N, HQ, HKV, D, dtype, gen)
    for seg in (32, 64):
        pin(sk._lse_and_store_logits_kernel, BLOCK_Q=32, BLOCK_K=32, num_warps=4, num_stages=3)
        pin(sk._scores_from_logits_kernel, BLOCK_Q=16, BLOCK_K=seg)
        out = sk.query_aware_key_scores(ref_in(q), ref_in(k), cu, cu, w)
        untouched = torch.zeros(N, dtype=torch.bool)
        untouched[cu[1]:cu[2]] = True
        save(f"snapkv_seg{seg}", q=q, k=k, cu_seqlens=cu, w=w, pool_segment=seg, out=out,
             untouched_rows=untouched)


def gen_leverage():
    """approximate_leverage_scores on CPU: `.cuda()` made a no-op and the cuSOLVER-only
    `driver="gesvda"` argument dropped (torch.linalg.svd's default CPU driver instead)."""
    from compactor_vllm.compression import compactor as cp

    pin(cp._zscore_per_batch_epilogue_no_window, BLOCK_K=64)
    real_svd = torch.linalg.svd
    real_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **kw: self
    cp.torch.linalg.svd = lambda A, full_matrices=True, driver=None: real_svd(A, full_matrices=full_matrices)
    try:
        for name, dtype in [("leverage_bf16", torch.bfloat16), ("leverage_f32in", torch.float32)]:
            gen = torch.Generator().manual_seed(5)
            H, D, sk = 2, 128, 48
            lens = [300, 40]
            N = sum(lens)
            low = torch.randn((N, 1, 12), generator=gen) @ torch.randn((12, D), generator=gen)
            keys = (0.5 * torch.randn((N, H, D), generator=gen) + 0.4 * low).to(dtype)
            phi = (torch.randn((D, sk), generator=gen).to(dtype) * (1 / np.sqrt(sk)))
            raw = cp.approximate_leverage_scores(keys.clone(), lens, phi, normalize=False, chunk_size=128)
            # z-score epilogue (:215-220): the kernel loads K-dtype scores as fp32 and stores them
            # back in K's dtype; run it on the exact fp32 up-cast and round (bf16 interpreter issue)
            _, chunk_lens = cp.split_into_chunks(lens, 128)
            cu_k = torch.tensor([0] + chunk_lens).cumsum(0)
            tmp = raw.float().clone().contiguous()
            cp._zscore_per_batch_epilogue_no_window[(len(chunk_lens),)](tmp, cu_k, tmp.stride(0), tmp.stride(1), H)
            norm = tmp.to(dtype)
            save(name, keys=keys, phi=phi, context_lens=np.array(lens), chunk_size=128, regularizer=5e-3,
                 raw=raw, norm=norm)
    finally:
        torch.Tensor.cuda = real_cuda
        cp.torch.linalg.svd = real_svd


def gen_prepath():
    """The reference's RotaryEmbedding / RMSNorm modules, eager (TorchDynamo disabled: their
    forward methods are torch.compile'd, which would need a CUDA-free inductor build here)."""
    import torch._dynamo
    torch._dynamo.config.disable = True
    from compactor_vllm.layers.layernorm import RMSNorm
    from compactor_vllm.layers.rotary_embedding import RotaryEmbedding

    gen = torch.Generator().manual_seed(77)
    D, HQ, HKV, Ntok, max_pos = 128, 8, 2, 37, 512
    for name, dtype, scaling, base in [
        ("rope_bf16_llama3", torch.bfloat16, ("llama3", 8.0, 1.0, 4.0, 64), 500000.0),
        ("rope_f16_plain", torch.float16, None, 10000.0),
    ]:
        rope = RotaryEmbedding(D, D, max_pos, base, scaling)
        q = torch.randn((Ntok, HQ, D), generator=gen).to(dtype)
        k = torch.randn((Ntok, HKV, D), generator=gen).to(dtype)
        pos = torch.randint(0, max_pos, (Ntok,), generator=gen)
        qo, ko = rope(pos, q, k)
        save(name, q=q, k=k, positions=pos, cos_sin_cache=rope.cos_sin_cache.view(max_pos, D),
             q_out=qo, k_out=ko, base=np.array(base),
             scaling=np.array(list(scaling[1:]) if scaling else [], dtype=np.float64))
    for name, dtype, width, rows in [("rmsnorm_bf16_128", torch.bfloat16, 128, 45),
                                     ("rmsnorm_bf16_4096", torch.bfloat16, 4096, 9),
                                     ("rmsnorm_f16_1024", torch.float16, 1024, 5)]:
        norm = RMSNorm(width, eps=1e-6)
        with torch.no_grad():
            norm.weight.copy_(1.0 + 0.2 * torch.randn(width, generator=gen))
        norm = norm.to(dtype)
        x = (2.0 * torch.randn((rows, width), generator=gen)).to(dtype)
        res = torch.randn((rows, width), generator=gen).to(dtype)
        with torch.no_grad():
            y = norm(x.clone())
            y2, r2 = norm(x.clone(), res.clone())
        save(name, x=x, residual=res, weight=norm.weight.detach(), y=y, y_add=y2, residual_out=r2,
             eps=np.array(1e-6))


from oracle.gen_golden_script import ALLOC_SCRIPT, ALLOC_SHAPE, alloc_lens  # noqa: E402


def gen_allocator():
    """The reference's PagedKVCache driven on CPU: `device="cuda"` of the one torch.tensor call in
    reserve_tokens (kv_cache/page_table.py:182) redirected to CPU, the Triton page scatter interpreted."""
    from compactor_vllm.kv_cache import page_table as ptm

    real_tensor, real_empty = torch.tensor, torch.empty

    def on_cpu(fn):
        def wrapped(*a, **k):
            if k.get("device") == "cuda":
                k["device"] = "cpu"
            return fn(*a, **k)
        return wrapped

    # page_table.py:182 and write_page_table.py's cumsum scratch name device="cuda" literally
    torch.tensor, torch.empty = on_cpu(real_tensor), on_cpu(real_empty)
    try:
        cache = ptm.PagedKVCache(dtype=torch.float16, device="cpu", **ALLOC_SHAPE)
        L, H, P = ALLOC_SHAPE["num_layers"], ALLOC_SHAPE["H_kv"], ALLOC_SHAPE["max_logical_pages_per_head"]
        status_code = {ptm.KVAllocationStatus.SUCCESS: 0, ptm.KVAllocationStatus.EXCEEDS_MAX_SEQUENCE_LENGTH: 1,
                       ptm.KVAllocationStatus.EXCEEDS_CURRENTLY_AVAILABLE_PAGES: 2}
        slots, rets, pts, pages, lens, frees = {}, [], [], [], [], []
        for op, name, arg in ALLOC_SCRIPT:
            ret = -1
            if op == "new":
                s = cache.new_batch()
                slots[name] = s
                ret = -1 if s is None else s
            elif op == "reserve":
                ret = status_code[cache.reserve_tokens(slots[name], arg)]
            elif op == "lens":
                cache.bh_seq_lens[:, slots[name]] = alloc_lens(arg, L, H, cache.bh_num_pages[:, slots[name]])
            elif op == "reclaim":
                ret = cache.reclaim_pages(slots[name], arg)
            elif op == "free":
                cache.free_batch(slots[name])
            rets.append(ret)
            valid = torch.arange(P).view(1, 1, 1, P) < cache.bh_num_pages.unsqueeze(-1)
            pts.append(torch.where(valid, cache.page_table, torch.full_like(cache.page_table, -1)).clone())
            pages.append(cache.bh_num_pages.clone())
            lens.append(cache.bh_seq_lens.clone())
            fr = torch.full((L, ALLOC_SHAPE["num_pages"]), -1, dtype=torch.int32)
            for l in range(L):
                f = sorted(cache.free_pages[l])
                fr[l, :len(f)] = torch.tensor(f, dtype=torch.int32)
            frees.append(fr)
        save("allocator_trace", rets=torch.tensor(rets, dtype=torch.int64), page_table=torch.stack(pts),
             bh_num_pages=torch.stack(pages), bh_seq_lens=torch.stack(lens), free_pages=torch.stack(frees))
    finally:
        torch.tensor, torch.empty = real_tensor, real_empty


def gen_num_splits():
    """The reference's own `num_splits_heuristic` (attention/sparse_decode_kernel.py:168-192) over a grid of
    (b x kv-head blocks, context, SM count, split cap): the split count its engine fixes at graph-capture time."""
    from compactor_vllm.attention.sparse_decode_kernel import num_splits_heuristic
    rows = []
    for sms in (108, 132, 148):
        for max_splits in (1, 4, 16, 64, 256):
            for blocks in (1, 2, 7, 8, 16, 24, 64, 100, 117, 118, 119, 128, 148, 512):
                for seq in (1, 512, 1024, 1025, 2048, 4097, 8192, 16344, 32768, 40960, 131072):
                    rows.append((blocks, seq, sms, max_splits, num_splits_heuristic(blocks, seq, sms, max_splits)))
    save("num_splits", table=torch.tensor(rows, dtype=torch.int64))


def gen_memory_budget():
    """The reference's KVCacheManager sizing methods (core/memory_manager.py:92-161) evaluated on scripted allocator
    readings: torch.cuda.mem_get_info / memory_stats are replaced by the scripted values, nothing else is touched."""
    from compactor_vllm.core import memory_manager as mm

    GB = 1 << 30
    rows_pages, rows_tokens = [], []
    real_info, real_stats = torch.cuda.mem_get_info, torch.cuda.memory_stats
    try:
        for total in (80 * GB, 180 * GB + 12345):
            for used in (17 * GB + 999, 30 * GB, 150 * GB):
                for peak, current in ((20 * GB, 16 * GB + 7), (5 * GB, 5 * GB)):
                    for frac in (0.5, 0.9, 0.95):
                        for layers, heads, hd, ps, esz, nb, mlen in ((32, 8, 128, 128, 2, 64, 40960),
                                                                    (80, 1, 128, 256, 2, 256, 131072),
                                                                    (36, 4, 64, 128, 4, 8, 4096)):
                            torch.cuda.mem_get_info = lambda *a, **k: (total - used, total)
                            torch.cuda.memory_stats = lambda *a, **k: {"allocated_bytes.all.peak": peak,
                                                                       "allocated_bytes.all.current": current}
                            m = object.__new__(mm.KVCacheManager)
                            m.gpu_frac, m.page_size, m.max_num_batches, m.num_layers = frac, ps, nb, layers
                            m.model_dtype = torch.float32 if esz == 4 else torch.bfloat16
                            m.head_dim, m.num_kv_heads = hd, heads
                            nlp = (mlen + ps - 1) // ps
                            try:
                                pages = int(m.get_num_pages(frac, nlp))
                            except RuntimeError:
                                pages = -1
                            rows_pages.append((total, used, peak, current, int(frac * 100), layers, heads, hd, ps,
                                               esz, nb, nlp, pages))
                            if pages > 0:
                                m.num_pages = pages
                                for wt, before, after in ((8192, 16 * GB, 19 * GB + 5), (1000, 4 * GB, 3 * GB)):
                                    tok = int(m.estimate_max_batched_tokens(wt, before, after))
                                    rows_tokens.append((total, used, peak, current, int(frac * 100), heads, ps,
                                                        pages, wt, before, after, tok))
    finally:
        torch.cuda.mem_get_info, torch.cuda.memory_stats = real_info, real_stats
    save("memory_budget", pages=torch.tensor(rows_pages, dtype=torch.int64),
         tokens=torch.tensor(rows_tokens, dtype=torch.int64))


if __name__ == "__main__":
    which = sys.argv[1:] or ["prefill", "decode", "stores", "chunk", "snapkv", "leverage", "prepath", "allocator",
                             "splits", "memory"]
    table = {"prefill": gen_prefill, "decode": gen_decode, "stores": gen_stores,
             "chunk": gen_chunk_scores, "snapkv": gen_snapkv, "leverage": gen_leverage,
             "prepath": gen_prepath, "allocator": gen_allocator, "splits": gen_num_splits,
             "memory": gen_memory_budget}
    for w in which:
        table[w]()
