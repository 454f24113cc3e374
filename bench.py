#!/usr/bin/env python
"""Headline benchmark of the attention + KV-compression hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload "M" (the "@32K, 50 % keep" row): per rank ONE 32768-token sequence,
HQ=32 HKV=8 D=128 page=128 bf16, COMPACTOR compression keeping 50 % (first 16 / last 64 tokens
protected), one transformer layer's hot path.  A *step* is one pass of that path over the
sequence through the reference-shaped Python API:

    scores = CompactorCompression.pre_rope_scoring(q, k_pre, v, ctx)     # leverage sketch, side stream
    scores = CompactorCompression.post_rope_scoring(q, k, v, scores, ctx) # chunk attention mass + blend
    out    = Attention.forward(q, k, v, scores)   # top-k select + page compaction on the side
                                                  # stream, tcgen05 prefill attention on the main one

`value` = algorithmic prefill attention FLOPs (causal half counted once, 4*D*HQ*L(L+1)/2) of all
ranks / max-over-ranks step time, i.e. sparse-attention prefill TFLOP/s WITH compression in the
step.  The other two numbers of the metric (decode HBM GB/s on the compacted cache, KV-compress
Mtok/s) are measured in the same run and reported under "components".

Ranks are independent sequences (weak scaling; the path has no collective).

`--impl reference`: the reference has no CPU implementation; the timed CPU arm is
the oracle/ restatement of the same path on the host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200")]

import torch  # noqa: E402

METRIC = "sparse attn prefill TFLOP/s, decode HBM GB/s, KV-compress Mtok/s @32K, 50% keep"
HQ, HKV, D, PS = 32, 8, 128, 128
SEQ = 32768
RATIO, PROT_FIRST, PROT_LAST = 0.5, 16, 64
SKETCH, LEV_CHUNK = 48, 512
CPU_SAMPLE_TOKENS = 4096


def prefill_flops(L: int, Lc: int = 0, hq: int = HQ) -> float:
    """4*D*HQ * (La*Lc + La(La+1)/2): QK^T + PV, 2 flops per MAC, only unmasked pairs."""
    return 4.0 * D * hq * (L * Lc + L * (L + 1) / 2)


def peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm": d["hbm_gbs"], "tensor_burst": d["bf16_tflops"],
                "tensor_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm": 6650.0, "tensor_burst": 1590.0, "tensor_sustained": 1400.0, "src": "fallback"}


# --------------------------------------------------------------------------------------------
# synthetic inputs (seeded; K has low-rank + positional structure so scores are not uniform)
# --------------------------------------------------------------------------------------------
def make_inputs(n_tok: int, device, dtype, seed: int, hq: int = HQ, hkv: int = HKV):
    HQ, HKV = hq, hkv  # noqa: N806 (shadow the module-level 8B head counts)
    g = torch.Generator(device=device).manual_seed(seed)
    qkv = torch.randn((n_tok, (HQ + 2 * HKV) * D), generator=g, device=device, dtype=torch.float32)
    basis = torch.randn((HKV, 12, D), generator=g, device=device)
    coef = torch.randn((n_tok, HKV, 12), generator=g, device=device) * torch.linspace(2.0, 0.2, 12, device=device)
    pos = torch.linspace(0, 50.0, n_tok, device=device)[:, None, None]
    low = torch.einsum("nhr,hrd->nhd", coef, basis) + 0.5 * torch.sin(pos + torch.arange(D, device=device) * 0.1)
    # .contiguous(): einsum hands back a head-major layout; the engine's pre-RoPE keys are token-major [N, HKV, D]
    k_pre = (low + 0.3 * torch.randn((n_tok, HKV, D), generator=g, device=device)).to(dtype).contiguous()
    qkv = qkv.to(dtype)
    kview = qkv[:, HQ * D:(HQ + HKV) * D].view(n_tok, HKV, D)
    kview.add_((0.25 * low).to(dtype))
    del basis, coef, low
    return qkv, k_pre


def split_qkv(qkv, hq: int = HQ, hkv: int = HKV):
    HQ, HKV = hq, hkv  # noqa: N806
    n = qkv.shape[0]
    q = qkv[:, : HQ * D].view(n, HQ, D)
    k = qkv[:, HQ * D:(HQ + HKV) * D].view(n, HKV, D)
    v = qkv[:, (HQ + HKV) * D:].view(n, HKV, D)
    return q, k, v


# --------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# --------------------------------------------------------------------------------------------
class ClockSampler:
    """Polls NVML (SM clock, power, clock-event reasons) every 10 ms while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, device_index: int):
        self.samples, self.power, self.reasons, self.max_mhz = [], [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = "GPU-" + str(torch.cuda.get_device_properties(device_index).uuid)
            try:
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self._nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._h = None

    def _run(self):
        nv, h = self._nv, self._h
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in self.REASONS.items():
                    if bits & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self._h is None:
            return
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s),
                "power_w_max": max(self.power) if self.power else None}


# --------------------------------------------------------------------------------------------
# the GPU arm
# --------------------------------------------------------------------------------------------
class Workload:
    """One rank's sequence, cache and the reference-shaped step."""

    def __init__(self, device, n_tok: int, seed: int, hq: int = HQ, hkv: int = HKV):
        HQ, HKV = hq, hkv  # noqa: N806
        self.hq, self.hkv = hq, hkv
        from compactor_vllm import _native as N
        from compactor_vllm.compression.compactor import CompactorCompression
        from compactor_vllm.compression.compression_config import CompressionMethod
        from compactor_vllm.kv_cache.page_table import KVAllocationStatus, PagedKVCache
        from compactor_vllm.layers.attention import Attention
        from compactor_vllm.utils import arguments as A
        from compactor_vllm.utils.context import CompressionContext

        self.N = N
        self.dev = device
        self.dtype = torch.bfloat16
        self.n_tok = n_tok
        self.plugin = CompactorCompression
        self.qkv, self.k_pre = make_inputs(n_tok, device, self.dtype, seed, hq, hkv)
        self.q, self.k, self.v = split_qkv(self.qkv, hq, hkv)
        n_lp = (n_tok + 2 * PS) // PS + 1
        self.cache = PagedKVCache(num_layers=1, max_logical_pages_per_head=n_lp, num_pages=HKV * n_lp + 8,
                                  page_size=PS, H_kv=HKV, head_dim=D, max_num_batches=2,
                                  dtype=self.dtype, device=device)
        self.slot = self.cache.new_batch()
        st = self.cache.reserve_tokens(self.slot, n_tok)
        assert st == KVAllocationStatus.SUCCESS, st
        # scramble the physical pages (the allocator hands them out ascending): paging is exercised
        pt = self.cache.page_table[0, self.slot]
        n_used = int(self.cache.bh_num_pages[0, self.slot, 0])
        flat = pt[:, :n_used].reshape(-1)
        perm = torch.randperm(flat.numel(), generator=torch.Generator().manual_seed(seed)).to(device)
        pt[:, :n_used] = flat[perm].view(HKV, n_used)
        self.attn = Attention(HQ, D, 1.0 / math.sqrt(D), HKV)
        kc, vc, ptab, bh = self.cache.layer_slices(0)
        self.attn.k_cache, self.attn.v_cache, self.attn.page_table = kc, vc, ptab
        self.attn.bh_seq_lens, self.attn.page_size = bh, PS
        self.batch_mapping = torch.tensor([self.slot], dtype=torch.int32, device=device)
        self.cu = torch.tensor([0, n_tok], dtype=torch.int32, device=device)
        retain = A.tokens_to_retain([n_tok], [RATIO], [PROT_FIRST], [PROT_LAST], HKV)
        self.retain = retain[0]
        self.cctx = CompressionContext(
            compression_method=CompressionMethod.COMPACTOR, compression_chunk_size=LEV_CHUNK,
            batch_tokens_to_retain=torch.tensor(retain, dtype=torch.int32, device=device),
            max_tokens_to_retain=A.max_tokens_to_retain([n_tok], HKV), context_lens=[n_tok],
            PHI=A.make_phi(D, SKETCH, self.dtype, device), protected_first_tokens=[PROT_FIRST],
            protected_last_tokens=[PROT_LAST])
        # the engine's STORE_STREAM (core/model_runner.py); high priority: its short kernels get SMs as attention CTAs retire (CVLLM_STORE_PRIO=0 for the reference default)
        self.store_stream = torch.cuda.Stream(device=device, priority=int(os.environ.get("CVLLM_STORE_PRIO", "-1")))
        self.attn_events = []  # (start, end) around the prefill attention launch

    # ---- context helpers -------------------------------------------------------------
    def _prefill_ctx(self, compress: bool):
        from compactor_vllm.utils.context import set_context
        set_context(is_prefill=True, do_compression=compress, cu_seqlens_q=self.cu, cu_seqlens_k=self.cu,
                    max_seqlen_q=self.n_tok, max_seqlen_k=self.n_tok, batch_mapping=self.batch_mapping,
                    max_bh_len=0, compression_context=self.cctx if compress else None,
                    STORE_STREAM=self.store_stream)

    def reset_slot(self):
        self.cache.bh_seq_lens[0, self.slot].zero_()

    # ---- the step --------------------------------------------------------------------
    def prefill_step(self, q=None, k=None, v=None, k_pre=None, compress=True):
        from compactor_vllm.utils.context import get_context
        q = self.q if q is None else q
        k = self.k if k is None else k
        v = self.v if v is None else v
        k_pre = self.k_pre if k_pre is None else k_pre
        self.reset_slot()
        self._prefill_ctx(compress)
        ctx = get_context()
        scores = None
        if compress:
            scores = self.plugin.pre_rope_scoring(q, k_pre, v, ctx)
            scores = self.plugin.post_rope_scoring(q, k, v, scores, ctx)
        out = self.attn(q, k, v, scores)
        torch.cuda.current_stream().wait_stream(self.store_stream)  # join: the step owns its stores
        return out

    def compress_only(self):
        """score -> select -> store chain alone (KV-compress Mtok/s): the same plugin calls and side streams as the
        step (scoring, selection and compaction on the store stream), no attention kernel.  Capturable in a CUDA
        graph."""
        from compactor_vllm.compression.common import extract_and_store_top_kv
        from compactor_vllm.utils.context import get_context
        from compactor_vllm.utils.helpers import maybe_execute_in_stream
        self.reset_slot()
        self._prefill_ctx(True)
        ctx, c = get_context(), self.cctx
        sc = self.plugin.pre_rope_scoring(self.q, self.k_pre, self.v, ctx)
        sc = self.plugin.post_rope_scoring(self.q, self.k, self.v, sc, ctx)
        lens = self.cache.bh_seq_lens[0].index_select(0, self.batch_mapping).contiguous()
        maybe_execute_in_stream(
            extract_and_store_top_kv, scores=sc, cu_seqlens_k=self.cu, max_k_len=self.n_tok,
            top_k=c.max_tokens_to_retain, H=self.hkv, new_keys=self.k, new_vals=self.v,
            num_tokens_to_retain=c.batch_tokens_to_retain, page_table=self.attn.page_table,
            batch_mapping=self.batch_mapping, bh_lens=lens, k_cache=self.attn.k_cache, v_cache=self.attn.v_cache,
            PAGE_SIZE=PS, PAD_TO_PAGE_SIZE=True, STORE_STREAM=self.store_stream)
        maybe_execute_in_stream(self.cache.bh_seq_lens[0].index_copy_, 0, self.batch_mapping.long(), lens,
                                STORE_STREAM=self.store_stream)
        torch.cuda.current_stream().wait_stream(self.store_stream)


def install_attention_timer(w: Workload):
    """Record CUDA events on the launching stream around the prefill attention call."""
    import compactor_vllm.layers.attention as LA
    inner = LA.causal_sparse_varlen_with_cache
    if getattr(inner, "_bench_timed", False):
        return

    def timed(*a, **kw):
        if torch.cuda.is_current_stream_capturing() or a[0].shape[0] != w.n_tok or a[0].shape[1] != w.hq:
            return inner(*a, **kw)  # a graph capture, or another workload of this run (4K / 16K / config 5)
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        out = inner(*a, **kw)
        e.record()
        w.attn_events.append((s, e))
        return out

    timed._bench_timed = True
    LA.causal_sparse_varlen_with_cache = timed


def time_loop(fn, iters: int, flush=None) -> float:
    """Average ms per call with CUDA events around each call (L2 flushed before each if given)."""
    tot = 0.0
    for _ in range(iters):
        if flush is not None:
            # cold AND clean L2: a 256 MB write leaves L2 full of dirty lines whose write-back the timed
            # kernel would pay for (measured: +6..14 us on a 120 us decode), so it is followed by a 256 MB read
            flush.zero_()
            flush.view(torch.float32).sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


def time_batch(fn, reps: int = 10) -> float:
    """Average ms per call of `reps` back-to-back calls inside ONE event pair: for kernels shorter than the
    Python launch path (the launches queue up, so the device, not the host, sets the pace)."""
    for _ in range(2):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps


def decode_bench(device, dtype, B: int, lens_bh: torch.Tensor, iters: int, flush, seed: int,
                 hq: int = HQ, hkv: int = HKV, after=None):
    """One decode step = decode_store_kv + head_sparse_decode_attention on a synthetic compacted cache,
    captured in a CUDA graph and replayed (how the reference engine runs decode, core/model_runner.py:524-528);
    every replay is timed with CUDA events after a cold-L2 flush.  Returns (ms per step, algorithmic bytes).
    `after(out)`, if given, is captured behind the attention in the same graph (config 5: o_proj + all-reduce)."""
    from compactor_vllm.attention.sparse_decode_kernel import decode_append_and_attend
    HQ, HKV = hq, hkv  # noqa: N806
    g = torch.Generator().manual_seed(seed)
    warm = 8
    nlp = int((int(lens_bh.max()) + iters + warm + 8 + PS - 1) // PS)
    n_pages = (B + 1) * HKV * nlp
    pt = torch.randperm(n_pages, generator=g).to(torch.int32).view(B + 1, HKV, nlp).to(device)
    kc = torch.randn((n_pages * PS, D), device=device, dtype=dtype)
    vc = torch.randn((n_pages * PS, D), device=device, dtype=dtype)
    bm = torch.arange(1, B + 1, dtype=torch.int32, device=device)
    lens = torch.zeros((B + 1, HKV), dtype=torch.int32, device=device)  # the cache's bh_seq_lens table, slot 0 reserved
    lens[1:] = lens_bh.to(torch.int32).to(device)
    q = torch.randn((B, HQ, D), device=device, dtype=dtype)
    kn = torch.randn((B, HKV, D), device=device, dtype=dtype)
    vn = torch.randn((B, HKV, D), device=device, dtype=dtype)

    def step():  # what Attention.forward's decode branch launches: append kernel + attention kernel
        out = decode_append_and_attend(q, kn, vn, kc, vc, lens, pt, bm, HKV, PS)
        return out if after is None else after(out)

    gs = torch.cuda.Stream(device=device)
    gs.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(gs):
        for _ in range(3):
            step()  # workspaces of the capture stream
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=gs):
            step()
    torch.cuda.synchronize()
    for _ in range(warm - 4):
        graph.replay()
    torch.cuda.synchronize()
    ms = time_loop(graph.replay, iters, flush)
    rows = int(lens_bh.sum()) + B * HKV * (warm + iters // 2)
    nbytes = rows * D * 2 * 2 + B * HQ * D * 2 * 2 + B * HKV * D * 2 * 2 * 2
    return ms, nbytes


# --------------------------------------------------------------------------------------------
# host placement, copy roofline, small-prompt overhead, config 5
# --------------------------------------------------------------------------------------------
def bind_to_gpu_numa(local: int) -> dict:
    """Pin this process to the CPUs NVML reports as local to its GPU (and, through first touch, its pinned
    buffers to that NUMA node) BEFORE any pinned allocation.  Returns what was done, for the bench line."""
    info = {"bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        before = os.sched_getaffinity(0)
        cpus &= before
        info.update(gpu_local_cpus=len(cpus), host_cpus=n_cpu)
        try:
            info["numa_node"] = int(pynvml.nvmlDeviceGetNumaNodeId(h))
        except Exception:
            pass
        if cpus and cpus != before:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
        else:
            info["note"] = "NVML reports every allowed CPU as local to this GPU: nothing to bind"
    except Exception as ex:
        info["error"] = repr(ex)[:120]
    return info


def copy_roofline(h_in, d_in, d_out, h_out, ranks, reps: int = 3, kernel_fn=None) -> dict:
    """The box's own copy ceiling on the e2e leg's buffers: H2D alone, D2H alone, both at once (two streams),
    every rank copying at the same time; GB/s summed over ranks and the duplex time of one step's bytes."""
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(do_in: bool, do_out: bool) -> float:
        ranks.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s1.wait_event(a)
        s2.wait_event(a)
        for _ in range(reps):
            if do_in:
                with torch.cuda.stream(s1):
                    for h, d in zip(h_in, d_in):
                        d.copy_(h, non_blocking=True)
            if do_out:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        b.record()
        b.synchronize()
        return a.elapsed_time(b) / reps

    by_in = sum(h.numel() * h.element_size() for h in h_in)
    by_out = h_out.numel() * h_out.element_size()
    run(True, True)
    t_in, t_out, t_dup = run(True, False), run(False, True), run(True, True)
    overlap = {}
    if kernel_fn is not None:
        # does a DMA copy overlap a running kernel on this box?  (seen both ways on this pool's single-GPU boxes:
        # some run the double-buffered loop at the copy bound, some serialise H2D behind the step's kernels)
        def both(do_copy: bool, do_kernel: bool) -> float:
            ranks.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            s1.wait_event(a)
            if do_kernel:
                kernel_fn()
            if do_copy:
                with torch.cuda.stream(s1):
                    for h, d in zip(h_in, d_in):
                        d.copy_(h, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1)
            b.record()
            b.synchronize()
            return a.elapsed_time(b)
        both(True, True)
        t_k, t_c, t_b = both(False, True), both(True, False), both(True, True)
        overlap = {"kernels_ms": t_k, "h2d_ms": t_c, "kernels_and_h2d_ms": t_b,
                   "h2d_overlaps_kernels_frac": ranks.max(-(t_k + t_c - t_b) / min(t_k, t_c)) * -1.0}
    return {**overlap, "h2d_gbs_alone": ranks.sum(by_in / t_in / 1e6), "d2h_gbs_alone": ranks.sum(by_out / t_out / 1e6),
            "duplex_gbs": ranks.sum((by_in + by_out) / t_dup / 1e6), "duplex_ms_per_step": ranks.max(t_dup),
            "note": "GB/s summed over ranks, all ranks copying at once; duplex = one step's H2D and D2H together"}


def overhead_at(dev, n_tok: int, iters: int, max_over_ranks) -> dict:
    """Compression overhead on a shorter prompt.  Reported: the step (scores + select + compaction on the store
    stream, attention on the main one) with and without compression, launched back to back so the device sets the
    pace where it can (at 4K tokens the Python launch path of the ~25 chain kernels is longer than the kernels:
    `chain_graph_ms` is the chain's device time replayed from a CUDA graph, `attn_kernel_ms` the attention kernel's).
    The whole step is not graph-captured: the reference's side-stream helper forks from the DEFAULT stream
    (utils/helpers.py:6-28), which a capture on another stream rejects."""
    w = Workload(dev, n_tok, seed=4321)
    out = {}
    for _ in range(3):
        w.prefill_step(compress=True)
        w.prefill_step(compress=False)
        w.compress_only()
    torch.cuda.synchronize()
    out["step_ms_with"] = max_over_ranks(time_batch(lambda: w.prefill_step(compress=True), iters))
    out["step_ms_without"] = max_over_ranks(time_batch(lambda: w.prefill_step(compress=False), iters))
    out["added_pct"] = 100.0 * (out["step_ms_with"] - out["step_ms_without"]) / out["step_ms_without"]
    try:
        gs = torch.cuda.Stream(device=dev)
        gs.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(gs):
            w.compress_only()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=gs):
                w.compress_only()
        torch.cuda.synchronize()
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        out["chain_graph_ms"] = max_over_ranks(time_loop(graph.replay, iters))
        del graph
    except Exception as ex:
        out["chain_graph_error"] = repr(ex)[:160]
    return out


def cfg5_block(dev, rank: int, world: int, ranks, steps: int) -> dict:
    """BASELINE config 5: ONE Llama-3.1-70B-shaped layer (HQ=64, HKV=8, D=128, hidden 8192) over a 32768-token
    prompt, COMPACTOR 50 %, sharded by KV head over the `world` ranks (reference models/llama3.py:41-50,
    core/memory_manager.py:29-32; local retain budget utils/arguments.py:109-121), then the row-parallel o_proj
    and the tensor-parallel all-reduce (layers/linear.py:149-153: NCCL for the 512 MiB prefill message, the
    peer-memory kernel for the B=16 decode step).  STRONG scaling: the layer is fixed, ranks split it."""
    import torch.distributed as dist
    from compactor_vllm.layers.linear import PeerAllReduce, PeerGemmAllReduce, RowParallelLinear
    hq, hkv, hidden, n_tok = 64 // world, 8 // world, 8192, SEQ
    mx = ranks.max
    res = {"tp": world, "hq_local": hq, "hkv_local": hkv, "hidden": hidden, "tokens": n_tok, "scaling": "strong"}
    w = Workload(dev, n_tok, seed=555 + rank, hq=hq, hkv=hkv)
    res["retain_local"] = w.retain
    peer = PeerAllReduce() if world > 1 else None
    # prefill-size o_proj: ONE kernel does the GEMM, the reduce-scatter and the all-gather over NVLink peer memory
    # (csrc/gemm_reduce.cu); the cuBLAS GEMM + NCCL all-reduce route is timed beside it
    fused = PeerGemmAllReduce(n_tok, hidden) if world > 1 else None
    lin = RowParallelLinear(hidden, hidden, peer_all_reduce=peer, peer_gemm_all_reduce=fused).to(dev).to(torch.bfloat16)
    with torch.no_grad():
        lin.weight.normal_(0.0, 0.02)

        y = torch.empty((n_tok, hidden), dtype=torch.bfloat16, device=dev)  # o_proj output, reused every step

        def prefill_lib():
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            out = w.prefill_step()
            ev[1].record()
            torch.mm(out.reshape(n_tok, hq * D), lin.weight.t(), out=y)
            ev[2].record()
            if world > 1:
                dist.all_reduce(y)
            ev[3].record()
            return ev

        def prefill_fused():
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev[0].record()
            out = w.prefill_step()
            ev[1].record()
            yf = lin(out.reshape(n_tok, hq * D))  # RowParallelLinear -> cvllm_gemm_allreduce
            ev[2].record()
            return ev, yf

        # the two routes are timed in ALTERNATION (lib step, fused step, ...) so that both see the same clocks: timed
        # one after the other, the second loop ran 5-10 % slower on power-capped boxes whatever it contained
        n_rep = max(3, steps // 4)
        for _ in range(3):
            prefill_lib()
            if world > 1:
                prefill_fused()
        ranks.barrier()
        evs, fevs = [], []
        for _ in range(n_rep):
            evs.append(prefill_lib())
            if world > 1:
                fevs.append(prefill_fused())
        torch.cuda.synchronize()
        mean = lambda es, i, j: sum(e[i].elapsed_time(e[j]) for e in es) / len(es)  # noqa: E731
        res["per_rank_ms"] = [round(mean(evs, 0, 1), 3), round(mean(evs, 1, 2), 3), round(mean(evs, 2, 3), 3)]
        res["attn_compress_ms"] = mx(mean(evs, 0, 1))
        res["oproj_gemm_ms"] = mx(mean(evs, 1, 2))
        res["allreduce_ms"] = mx(mean(evs, 2, 3)) if world > 1 else 0.0
        lib_ms = mx(mean(evs, 0, 3))
        res["prefill_ms_cublas_nccl"] = lib_ms
        msg = n_tok * hidden * 2
        res["allreduce_bytes"] = msg
        if world > 1:
            res["allreduce_busbw_gbs"] = 2 * (world - 1) / world * msg / (res["allreduce_ms"] * 1e-3) / 1e9
            only = [e for e, _ in fevs]
            res["attn_compress_ms_fused_steps"] = mx(mean(only, 0, 1))
            res["oproj_fused_ms"] = mx(mean(only, 1, 2))
            res["prefill_ms"] = mx(mean(only, 0, 2))
            res["oproj_route"] = "fused tcgen05 GEMM + reduce-scatter + all-gather over peer memory (one kernel)"
            res["oproj_fused_speedup"] = (res["oproj_gemm_ms"] + res["allreduce_ms"]) / res["oproj_fused_ms"]
            # same sum as the library route: NCCL's order may differ in the last bit of a few entries
            prefill_lib()
            _, yf = prefill_fused()
            torch.cuda.synchronize()
            res["oproj_fused_max_abs_diff_vs_nccl"] = ranks.max(float((yf.float() - y.float()).abs().max()))
            res["oproj_fused_frac_equal_nccl"] = 1.0 - ranks.max(float((yf != y).float().mean()))
        else:
            res["prefill_ms"] = lib_ms
            res["oproj_route"] = "cuBLAS (tp = 1: no collective)"
        res["allreduce_share"] = res["allreduce_ms"] / lib_ms
        res["prefill_tflops"] = prefill_flops(n_tok, hq=64) / (res["prefill_ms"] * 1e-3) / 1e12
        kept = w.cache.bh_seq_lens[0, w.slot].cpu()
        res["kept_pairs_local"] = int(kept.sum())

        # decode: B = 16 on the compacted shape, o_proj + all-reduce captured behind the attention
        B = 16
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        lens = kept.view(1, hkv).expand(B, hkv).contiguous()
        ms_attn, nbytes = decode_bench(dev, torch.bfloat16, B, lens, 20, flush, 9, hq=hq, hkv=hkv)
        ms_full, _ = decode_bench(dev, torch.bfloat16, B, lens, 20, flush, 9, hq=hq, hkv=hkv,
                                  after=lambda o: lin(o.reshape(B, hq * D)))
        res["decode_b16_attn_us"] = mx(ms_attn) * 1e3
        res["decode_b16_step_us"] = mx(ms_full) * 1e3
        res["decode_b16_oproj_allreduce_us"] = res["decode_b16_step_us"] - res["decode_b16_attn_us"]
        res["decode_b16_gbs_all_ranks"] = ranks.sum(nbytes / (ms_attn * 1e-3) / 1e9)
        res["decode_allreduce_route"] = ("peer-memory one-shot" if world > 1 else "none (tp=1)")
        del flush
        if world > 1:
            sys.path.insert(0, str(ROOT / "tools"))
            from tp_allreduce_check import run_checks
            chk = run_checks(dev, rank, world, timing=False)
            res["allreduce_checks_ok"] = bool(ranks.sum(0.0 if chk["ok"] else 1.0) == 0)
            res["allreduce_checks"] = chk["checks"]
            peer.close()
            fused.close()
    del w, lin
    torch.cuda.empty_cache()
    return res


def cpu_sample(n_tok: int, threads: int):
    """The oracle's restatement of one step on `n_tok` tokens; returns (seconds, prefill flops)."""
    import oracle.attention as oa
    import oracle.scoring as osc
    import oracle.selection as osel
    torch.set_num_threads(threads)
    qkv, k_pre = make_inputs(n_tok, "cpu", torch.bfloat16, 7)
    q, k, v = split_qkv(qkv)
    from compactor_vllm.utils import arguments as A
    phi = A.make_phi(D, SKETCH, torch.bfloat16, "cpu")
    n_lp = n_tok // PS + 2
    pt = torch.arange(2 * HKV * n_lp, dtype=torch.int32).view(2, HKV, n_lp)
    kc = torch.zeros((2 * HKV * n_lp * PS, D), dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    lens = torch.zeros((1, HKV), dtype=torch.int32)
    bm = torch.tensor([1], dtype=torch.int32)
    cu = torch.tensor([0, n_tok], dtype=torch.int32)
    retain = A.tokens_to_retain([n_tok], [RATIO], [PROT_FIRST], [PROT_LAST], HKV)
    t0 = time.perf_counter()
    pre = osc.approximate_leverage_scores(k_pre, [n_tok], phi, normalize=True, chunk_size=LEV_CHUNK)
    sc = osc.non_causal_attn_scores(q, k, cu, 128, sm_scale=1.0, normalize=True, context_lens=[n_tok],
                                    protected_first_tokens=[PROT_FIRST], protected_last_tokens=[PROT_LAST],
                                    accum_scores=pre, accum_blending=0.5)
    out = oa.prefill_attention(q, k, v, kc, vc, lens, pt, bm, cu, HKV, PS)
    osel.extract_and_store_top_kv(sc, cu, n_tok, n_tok * HKV, HKV, k, v, retain, pt, bm, lens, kc, vc, PS,
                                  pad=True)
    dt = time.perf_counter() - t0
    assert torch.isfinite(out).all()
    return dt, prefill_flops(n_tok)


def run_reference_arm(args, rank: int, world: int):
    """The CPU arm: the oracle/ restatement of one step on the host cores (the reference has no CPU path).  A step
    here is a bounded SLICE of the workload (stated in config.workload and cpu_baseline.sample), FLOP-normalised;
    warm-up and step counts are the requested ones whenever the whole run fits in ~150 s."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    budget_s = 150.0
    n = CPU_SAMPLE_TOKENS
    t_begin = time.perf_counter()
    dt0, _ = cpu_sample(n, threads)  # sizing probe (not a warm-up step of the timed configuration unless kept)
    total_steps = max(0, args.warmup) + max(1, args.steps)
    if dt0 * 16 * total_steps < budget_s:  # cost grows ~quadratically: the 16K slice costs ~16x the 4K one
        n = 4 * CPU_SAMPLE_TOKENS
        dt0 = None
    n_warm = 0
    for i in range(max(0, args.warmup)):
        if dt0 is not None and i == 0:
            n_warm += 1  # the probe ran the timed configuration: it is the first warm-up step
            continue
        if time.perf_counter() - t_begin > budget_s / 3:
            break
        cpu_sample(n, threads)
        n_warm += 1
    ts = []
    for _ in range(max(1, args.steps)):
        dt, fl = cpu_sample(n, threads)
        ts.append(dt)
        if time.perf_counter() - t_begin > budget_s:  # keep the whole arm within a few minutes
            break
    dt = sum(ts) / len(ts)
    val = fl / dt / 1e12
    cfg = workload_config()
    cfg["workload"] = (f"{n}-token SLICE of M (the CPU arm's bounded sample; TFLOP/s is FLOP-normalised): "
                       f"1 x {n}-token sequence, HQ={HQ} HKV={HKV} D={D} page={PS}, COMPACTOR keep "
                       f"{int(RATIO * 100)}% (first {PROT_FIRST}/last {PROT_LAST} protected), one layer's hot path "
                       f"per step; the GPU arm runs {SEQ} tokens")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "TFLOP/s", "n_gpus": args.gpus,
        "steps": len(ts), "warmup": n_warm, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": threads, "kind": "port",
                         "sample": f"{n}-token slice of the workload (same heads, same compression), "
                                   "oracle/ restatement: the reference has no CPU path (SURVEY 8d)"},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config() -> dict:
    return {
        "workload": f"M: 1 x {SEQ}-token sequence per rank, HQ={HQ} HKV={HKV} D={D} page={PS} bf16, "
                    f"COMPACTOR keep {int(RATIO * 100)}% (first {PROT_FIRST}/last {PROT_LAST} protected), "
                    "one layer's hot path per step",
        "l2": "prefill inputs (470 MB) exceed the 126 MB L2; decode iterations flush L2 (256 MB write, then a 256 MB read so the flushed L2 holds no dirty lines)",
        "parallelism": "independent sequences per rank, no collective in the path",
    }


class Ranks:
    """The harness's only use of torch.distributed: a barrier and max / sum reductions over the
    ranks (NCCL on GPUs, gloo in the CPU tests).  The measured path itself has no collective."""

    def __init__(self, world: int, device, backend: str | None = None, init: bool = True):
        self.world, self.device, self.dist = world, torch.device(device), None
        if world > 1:
            import torch.distributed as dist
            self.dist = dist
            if init and not dist.is_initialized():
                if backend is None:
                    backend = "nccl" if self.device.type == "cuda" else "gloo"
                kw = {"device_id": self.device} if backend == "nccl" else {}
                dist.init_process_group(backend, **kw)
            self._owns = init

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    def _reduce(self, x: float, op) -> float:
        if self.dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.MAX) if self.dist is not None else x

    def sum(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.SUM) if self.dist is not None else x

    def close(self):
        if self.dist is not None and getattr(self, "_owns", False) and self.dist.is_initialized():
            self.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--seq", type=int, default=SEQ, help=argparse.SUPPRESS)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="only the headline step (for ncu runs)")
    ap.add_argument("--only-cfg5", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    args.warmup = max(args.warmup, 3)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa(local)  # before the first pinned allocation (first touch places the pages)
    ranks = Ranks(world, dev)
    barrier, max_over_ranks, sum_over_ranks = ranks.barrier, ranks.max, ranks.sum

    from compactor_vllm import _native as N
    lib = N.lib()
    n_tok = args.seq
    w = Workload(dev, n_tok, seed=1234 + rank)
    install_attention_timer(w)
    pk = peaks()
    flops = prefill_flops(n_tok)

    # ---------------- headline: K timed steps ----------------
    for _ in range(args.warmup):
        w.prefill_step()
    def timed_block():
        barrier()
        w.attn_events.clear()
        sampler = ClockSampler(local)
        barrier()
        sampler.start()
        launches0 = lib.cvllm_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        marks = []
        for _ in range(args.steps):
            w.prefill_step()
            marks.append(torch.cuda.Event(enable_timing=True))
            marks[-1].record()
        ev1.record()
        barrier()
        n_launch = lib.cvllm_launch_count() - launches0
        ms = max_over_ranks(ev0.elapsed_time(ev1) / args.steps)
        steps_sorted = sorted(a.elapsed_time(b) for a, b in zip([ev0] + marks[:-1], marks))
        a_ms = sum(s.elapsed_time(e) for s, e in w.attn_events) / max(1, len(w.attn_events))
        return ms, steps_sorted, a_ms, n_launch, sampler.stop()

    ms_step, per_step, attn_ms, launches, clocks = timed_block()
    # The timing rules' "rejected and re-measured once": a block that saw a thermal / hw slowdown, or whose
    # mean step is far above its own median step (a stall outside the kernels: another process on the box,
    # a driver hiccup), is measured once more; the second block is the one reported, the first is kept on the line.
    remeasured = None
    bad_reasons = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(clocks.get("reasons") or [])
    stalled = max_over_ranks(1.0 if ms_step > 1.25 * per_step[len(per_step) // 2] else 0.0) > 0
    if max_over_ranks(1.0 if bad_reasons else 0.0) > 0 or stalled:  # the same decision on every rank
        remeasured = {"first_ms_per_step": ms_step, "first_clocks": clocks,
                      "why": sorted(bad_reasons) or "mean step > 1.25 x median step"}
        ms_step, per_step, attn_ms, launches, clocks = timed_block()
    value = flops * world / (ms_step * 1e-3) / 1e12
    total_launches = int(sum_over_ranks(float(launches)))

    # kept pairs after the last step (for the decode leg and the report)
    kept_bh = w.cache.bh_seq_lens[0, w.slot].cpu()
    kept_pairs = int(kept_bh.sum())

    comps = {"prefill_tflops_with_compression": value, "prefill_attn_kernel_ms": attn_ms,
             "step_ms_min_med_max": [per_step[0], per_step[len(per_step) // 2], per_step[-1]],
             **({"remeasured": remeasured} if remeasured else {}),
             "kept_pairs": kept_pairs, "retain_budget": w.retain}
    roof_dec = None
    e2e = None
    cpu_base = None
    if args.only_cfg5:
        comps["cfg5"] = cfg5_block(dev, rank, world, ranks, args.steps)
        if rank == 0:
            print(json.dumps(comps["cfg5"]), flush=True)
        ranks.close()
        return
    if not args.no_extras:
        # ---------------- attention alone (no compression in the step) ----------------
        for _ in range(2):
            w.prefill_step(compress=False)
        torch.cuda.synchronize()
        w.attn_events.clear()
        ms_plain = max_over_ranks(time_loop(lambda: w.prefill_step(compress=False), max(3, args.steps // 2)))
        # the attention kernel with the SMs to itself (in the headline step it shares them with the
        # scoring / selection / compaction kernels of the store stream)
        comps["prefill_attn_kernel_alone_ms"] = (sum(s.elapsed_time(e) for s, e in w.attn_events)
                                                 / max(1, len(w.attn_events)))
        comps["prefill_tflops_no_compression"] = flops * world / (ms_plain * 1e-3) / 1e12
        comps["compression_added_pct_sequential"] = 100.0 * (ms_step - ms_plain) / ms_plain
        # The figure on the line comes from steps with and without compression in ALTERNATION: two loops timed one after
        # the other see different SM clocks on a power-capped box (the second of two identical loops ran 5-10 % apart),
        # which used to land in this number whichever way it fell.
        n_pairs = max(5, args.steps // 2)
        t_with = t_without = 0.0
        for _ in range(n_pairs):
            t_with += time_loop(lambda: w.prefill_step(), 1)
            t_without += time_loop(lambda: w.prefill_step(compress=False), 1)
        ms_with, ms_without = max_over_ranks(t_with / n_pairs), max_over_ranks(t_without / n_pairs)
        comps["compression_added_pct"] = 100.0 * (ms_with - ms_without) / ms_without
        comps["compression_added_paired_ms"] = [ms_with, ms_without]
        # the same overhead at config 1's and config 2's prompt lengths (one 4096- / 16384-token sequence): the
        # eigensolver's cost does not shrink with the prompt, so the short-prompt figures are the honest ones
        for tag, n_small in (("4k", 4096), ("16k", 16384)):
            try:
                o = overhead_at(dev, n_small, max(5, args.steps // 2), max_over_ranks)
                comps[f"compression_added_pct_{tag}"] = o["added_pct"]
                comps[f"compression_overhead_{tag}"] = o
            except Exception as ex:
                comps[f"compression_overhead_{tag}_error"] = repr(ex)[:200]

        # ---------------- compression chain alone ----------------
        for _ in range(2):
            w.compress_only()
        torch.cuda.synchronize()
        ms_comp = max_over_ranks(time_loop(w.compress_only, max(3, args.steps // 2)))
        comps["compress_mtok_s"] = n_tok * world / (ms_comp * 1e-3) / 1e6
        comps["compress_ms"] = ms_comp
        # the same chain replayed from a CUDA graph: device time without the Python launch path
        try:
            gstream = torch.cuda.Stream(device=dev)
            gstream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(gstream):
                w.compress_only()  # workspaces for this stream
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=gstream):
                    w.compress_only()
            torch.cuda.synchronize()
            for _ in range(2):
                graph.replay()
            torch.cuda.synchronize()
            ms_graph = max_over_ranks(time_loop(graph.replay, max(3, args.steps // 2)))
            comps["compress_graph_ms"] = ms_graph
            comps["compress_graph_mtok_s"] = n_tok * world / (ms_graph * 1e-3) / 1e6
            del graph
        except Exception as ex:  # capture is a measurement aid, not part of the path
            comps["compress_graph_error"] = repr(ex)[:200]
        comp_bytes = n_tok * (2048 + 8192 + 2048 + HKV * 4 * 4) + kept_pairs * 1024
        comps["compress_implied_gbs"] = comp_bytes / (ms_comp * 1e-3) / 1e9
        if "compress_graph_ms" in comps:  # device time of the chain against its algorithmic bytes
            g_gbs = comp_bytes / (comps["compress_graph_ms"] * 1e-3) / 1e9
            comps["roofline_compress"] = {"bound": "hbm", "achieved": g_gbs, "peak": pk["hbm"], "unit": "GB/s",
                                          "frac": g_gbs / pk["hbm"], "bytes": comp_bytes,
                                          "note": "whole chain (22 launches, CUDA-graph replay): latency-bound by the "
                                                  "eigensolver, not by bandwidth"}

        # ---------------- the step in front of the path: fused q/k norm + sketch + RoPE ----------------
        try:
            from compactor_vllm.compression.compactor import approximate_leverage_scores
            from compactor_vllm.layers.rotary_embedding import build_cos_sin_cache, fused_qk_norm_rope
            cache = build_cos_sin_cache(D, 40960, 500000.0, ("llama3", 8.0, 1.0, 4.0, 8192)).to(dev)
            pos = torch.arange(n_tok, device=dev)
            qs, ks, _ = split_qkv(w.qkv)
            nw = torch.ones(D, dtype=w.dtype, device=dev)
            phi = getattr(w.cctx, "PHI", None)
            if phi is None:
                phi = (torch.randn((D, 48), device=dev) / 48 ** 0.5).to(w.dtype)
            rope_bytes = n_tok * (HQ + HKV) * D * 2 * 2
            ms_rope = max_over_ranks(time_batch(lambda: fused_qk_norm_rope(pos, qs, ks, cache), 20))
            ms_fused = max_over_ranks(time_batch(
                lambda: fused_qk_norm_rope(pos, qs, ks, cache, q_norm_weight=nw, k_norm_weight=nw, PHI=phi), 20))
            comps["rope_ms"] = ms_rope
            comps["rope_gbs"] = rope_bytes / (ms_rope * 1e-3) / 1e9
            comps["norm_rope_sketch_ms"] = ms_fused
            comps["norm_rope_sketch_gbs"] = (rope_bytes + n_tok * HKV * 96) / (ms_fused * 1e-3) / 1e9
            sk = fused_qk_norm_rope(pos, qs, ks, cache, PHI=phi)[2]
            lens = [n_tok]
            comps["leverage_from_k_ms"] = max_over_ranks(time_batch(
                lambda: approximate_leverage_scores(ks, lens, phi, normalize=True, chunk_size=512), 10))
            comps["leverage_from_sketch_ms"] = max_over_ranks(time_batch(
                lambda: approximate_leverage_scores(None, lens, phi, normalize=True, chunk_size=512,
                                                    sketch_rows=sk), 10))
            del cache, sk
        except Exception as ex:
            comps["prepath_error"] = repr(ex)[:200]

        # ---------------- page allocator: host heaps (the reference's design) vs device bitmap ----------------
        try:
            from compactor_vllm.kv_cache.page_table import DevicePagedKVCache, PagedKVCache
            shape = dict(num_layers=32, max_logical_pages_per_head=320, num_pages=4096, page_size=128, H_kv=HKV,
                         head_dim=8, max_num_batches=4, dtype=torch.bfloat16, device=dev)
            for name, cls in (("host", PagedKVCache), ("device", DevicePagedKVCache)):
                c = cls(**shape)
                s = c.new_batch()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                reps = 20
                for i in range(reps):  # a 128-token page per head per call: what a decode step's reserve does
                    c.reserve_tokens(s, 128)
                    c.bh_seq_lens[:, s] += 128
                torch.cuda.synchronize()
                comps[f"reserve_{name}_us"] = (time.perf_counter() - t0) / reps * 1e6
                del c
            c = DevicePagedKVCache(**shape)
            sl = torch.tensor([c.new_batch()], dtype=torch.int32, device=dev)
            st = torch.zeros(1, dtype=torch.int32, device=dev)
            c.reserve_tokens_device(sl, 128, st)
            comps["reserve_device_nosync_us"] = 1e3 * time_loop(
                lambda: c.reserve_tokens_device(sl, 1, st), 20)
            del c
        except Exception as ex:
            comps["allocator_error"] = repr(ex)[:200]

        # ---------------- decode on the compacted shape ----------------
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        it = 40
        ms_d1, by_d1 = decode_bench(dev, torch.bfloat16, 1, kept_bh.view(1, HKV), it, flush, 5)
        ms_d1 = max_over_ranks(ms_d1)
        comps["decode_gbs"] = by_d1 * world / (ms_d1 * 1e-3) / 1e9
        comps["decode_us"] = ms_d1 * 1e3
        g = torch.Generator().manual_seed(3)
        lens16 = (torch.rand((16, HKV), generator=g) * 1.0 + 0.5) * 10240  # ragged, mean 10240 (cfg 3: 40K x 25 %)
        ms_d16, by_d16 = decode_bench(dev, torch.bfloat16, 16, lens16.long(), it, flush, 6)
        ms_d16 = max_over_ranks(ms_d16)
        comps["decode_b16_gbs"] = by_d16 * world / (ms_d16 * 1e-3) / 1e9
        comps["decode_b16_us"] = ms_d16 * 1e3
        comps["decode_b16_cfg3_note"] = "decode_b16_*: B=16 ragged, mean 10240 rows/head (config 3: 40K x 25 %)"
        # the metric's own shape: every sequence is the headline's compacted cache (32K context, 50 % keep)
        for bsz in (8, 16):
            ms_b, by_b = decode_bench(dev, torch.bfloat16, bsz, kept_bh.view(1, HKV).expand(bsz, HKV).contiguous(),
                                      it, flush, 20 + bsz)
            ms_b = max_over_ranks(ms_b)
            comps[f"decode_b{bsz}_32k50_gbs"] = by_b * world / (ms_b * 1e-3) / 1e9
            comps[f"decode_b{bsz}_32k50_us"] = ms_b * 1e3
        gbs_h = comps["decode_b16_32k50_gbs"] / world
        roof_dec = {"bound": "hbm", "achieved": gbs_h, "peak": pk["hbm"], "unit": "GB/s",
                    "frac": gbs_h / pk["hbm"], "traffic": None,
                    "kernel": "decode step (decode_store_kernel + decode_attn_kernel) replayed from a CUDA graph, "
                              f"B=16 at the metric's shape: {int(kept_bh.sum())} cached (token, head) pairs per "
                              "sequence (32K context, 50 % keep)",
                    "peak_src": pk["src"],
                    "b1_frac": comps["decode_gbs"] / world / pk["hbm"], "b8_frac": comps["decode_b8_32k50_gbs"] / world / pk["hbm"],
                    "b16_cfg3_frac": by_d16 / (ms_d16 * 1e-3) / 1e9 / pk["hbm"]}
        del flush

        # ---------------- end to end: pinned host buffers in, host out ----------------
        h_qkv = torch.empty(w.qkv.shape, dtype=w.qkv.dtype).pin_memory()
        h_kpre = torch.empty(w.k_pre.shape, dtype=w.k_pre.dtype).pin_memory()
        h_qkv.copy_(w.qkv)
        h_kpre.copy_(w.k_pre)
        h_out = torch.empty((n_tok, HQ, D), dtype=w.dtype).pin_memory()
        d_qkv = torch.empty_like(w.qkv)
        d_kpre = torch.empty_like(w.k_pre)

        def e2e_step():
            d_kpre.copy_(h_kpre, non_blocking=True)
            d_qkv.copy_(h_qkv, non_blocking=True)
            q, k, v = split_qkv(d_qkv)
            out = w.prefill_step(q, k, v, d_kpre)
            h_out.copy_(out, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(3, args.steps // 2)
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / n_e2e)
        # The same call, software-pipelined the way a serving loop feeds it: two device input buffers and two
        # pinned output buffers; the H2D of step i+1 (copy stream) and the D2H of step i-1 (second copy stream)
        # run under step i's kernels.  Every step still moves its own inputs from pinned host memory and lands
        # its own output in pinned host memory inside the timed region; a host output buffer is reused only
        # after its previous D2H has completed (the host-side read).
        d_in = [(d_qkv, d_kpre), (torch.empty_like(w.qkv), torch.empty_like(w.k_pre))]
        h_outs = [h_out, torch.empty((n_tok, HQ, D), dtype=w.dtype).pin_memory()]
        s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_done = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]

        def e2e_pipelined(n_steps):
            main = torch.cuda.current_stream()
            for ev in ev_done + ev_out:
                ev.record(main)
            for i in range(n_steps):
                bsel = i & 1
                dq, dk = d_in[bsel]
                with torch.cuda.stream(s_in):
                    s_in.wait_event(ev_done[bsel])  # step i-2 has finished reading this input buffer
                    dk.copy_(h_kpre, non_blocking=True)
                    dq.copy_(h_qkv, non_blocking=True)
                    ev_in[bsel].record(s_in)
                main.wait_event(ev_in[bsel])
                q, k, v = split_qkv(dq)
                out = w.prefill_step(q, k, v, dk)
                ev_done[bsel].record(main)
                ev_out[bsel].synchronize()  # the host has step i-2's output: its buffer may be rewritten
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_done[bsel])
                    h_outs[bsel].copy_(out, non_blocking=True)
                    out.record_stream(s_out)
                    ev_out[bsel].record(s_out)
            for ev in ev_out:
                ev.synchronize()
            main.synchronize()

        e2e_pipelined(3)
        barrier()
        t0 = time.perf_counter()
        e2e_pipelined(2 * n_e2e)
        barrier()
        pipe_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / (2 * n_e2e))
        try:
            roofc = copy_roofline([h_kpre, h_qkv], [d_kpre, d_qkv], w.prefill_step(compress=False), h_out, ranks,
                                  kernel_fn=lambda: w.prefill_step())
        except Exception as ex:
            roofc = {"error": repr(ex)[:160]}
        e2e = {"value": flops * world / (pipe_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
               "h2d_bytes_per_step": int(h_qkv.numel() * 2 + h_kpre.numel() * 2),
               "d2h_bytes_per_step": int(h_out.numel() * 2), "ms_per_step": pipe_ms,
               "mode": "double-buffered: step i+1's H2D and step i-1's D2H overlap step i's kernels; "
                       "every step copies its own inputs from pinned host memory and its output to pinned host "
                       "memory inside the timed region (host wall clock, first H2D to last D2H)",
               "serial_value": flops * world / (e2e_ms * 1e-3) / 1e12, "serial_ms_per_step": e2e_ms,
               "copy_roofline": roofc, "numa": numa,
               "frac_of_copy_bound": (roofc["duplex_ms_per_step"] / pipe_ms) if "duplex_ms_per_step" in roofc else None,
               "kernel_ms_per_step": ms_step,
               "api": "CompactorCompression.pre/post_rope_scoring + Attention.forward, pinned host q/k/v in, out back"}
        del d_in, h_outs
        del h_qkv, h_kpre, h_out, d_qkv, d_kpre

        # ---------------- config 5: 70B layer, KV heads sharded over the ranks, o_proj + TP all-reduce ----------------
        if 8 % world == 0:
            try:
                comps["cfg5"] = cfg5_block(dev, rank, world, ranks, args.steps)
            except Exception as ex:
                if world > 1:
                    raise  # a rank that drops out would leave its peers spinning in a collective
                comps["cfg5_error"] = repr(ex)[:200]

        # ---------------- CPU baseline (rank 0, N=1 only) ----------------
        if rank == 0 and world == 1 and not args.no_cpu:
            threads = os.cpu_count() or 1
            n_cpu = CPU_SAMPLE_TOKENS
            dt, fl = cpu_sample(n_cpu, threads)
            if dt < 2.5:  # fast host: take the 16K slice so the sample is ~10-30 s of CPU work
                n_cpu = 4 * CPU_SAMPLE_TOKENS
                dt, fl = cpu_sample(n_cpu, threads)
            cpu_base = {"value": fl / dt / 1e12, "unit": "TFLOP/s", "cores": threads, "kind": "port",
                        "sample": f"one step on a {n_cpu}-token slice (same heads and compression), "
                                  f"oracle/ fp32 restatement, {dt:.1f} s"}

    if rank == 0:
        # Which measured peak is the denominator: the burst figure (cuBLAS bf16 at full clocks) unless THIS run's
        # clock sampler saw the sustained regime — median SM clock within 10 % of the clock the sustained figure
        # was taken at (MEASURED_PEAKS.json clocks_under_load.sm_mhz_median, ~1300 MHz under the 1000 W cap).
        sus_mhz = 1282.0
        try:
            sus_mhz = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
                            .get("clocks_under_load", {}).get("sm_mhz_median", sus_mhz))
        except Exception:
            pass
        sustained_regime = bool(clocks.get("sm_mhz")) and clocks["sm_mhz"] <= 1.1 * sus_mhz
        peak_t = pk["tensor_sustained"] if sustained_regime else pk["tensor_burst"]
        ach = flops / (attn_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": ach, "peak": peak_t, "unit": "TFLOP/s", "frac": ach / peak_t,
                "traffic": None, "kernel": "prefill_attn2_kernel<bf16>",
                "peak_src": pk["src"] + (" (sustained cuBLAS bf16: this run's median SM clock is in the sustained regime)"
                                         if sustained_regime else
                                         " (burst cuBLAS bf16: this run's median SM clock is far above the "
                                         "sustained-regime clock)"),
                "frac_of_burst_peak": ach / pk["tensor_burst"], "frac_of_sustained_peak": ach / pk["tensor_sustained"]}
        alone = comps.get("prefill_attn_kernel_alone_ms")
        if alone:
            roof["achieved_alone"] = flops / (alone * 1e-3) / 1e12
            roof["frac_alone_of_burst_peak"] = roof["achieved_alone"] / pk["tensor_burst"]
            roof["note"] = ("achieved/frac: the kernel timed inside the headline step, where the store stream's "
                            "compression kernels take SMs from it; achieved_alone: the same kernel in the "
                            "uncompressed step")
        # dram__bytes_read+write of one launch from the newest committed ncu --set full capture (profiles/)
        for key, r in (("prefill", roof), ("decode", roof_dec)):
            cands = sorted((ROOT / "profiles").glob(f"r*_{key}_traffic.json"))
            if cands and r is not None:
                try:
                    r["traffic"] = json.loads(cands[-1].read_text()).get("dram_bytes_per_launch")
                    r["traffic_src"] = cands[-1].name
                except Exception:
                    pass
        line = {
            "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(),
            "components": comps, "roofline": roof, "roofline_decode": roof_dec, "cpu_baseline": cpu_base,
            "e2e": e2e, "gpu_launches": total_launches, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    ranks.close()


if __name__ == "__main__":
    main()
