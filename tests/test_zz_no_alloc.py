"""`cvllm_rank_scores_static` — the allocation-free variant of `cvllm_rank_scores` (scratch in a file-scope device array
of 8 MiB, csrc/select.cu) — against the existing entry point, through the C ABI, on two calls in a row.
Bar: bit-exact (index work).  Collected LAST on purpose (`zz`).

Inputs and outputs are allocated by the test before the calls; the calls themselves allocate nothing, which the test
checks on torch's allocator counters (the library has no allocator of its own: a cudaMalloc inside would be a bug)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(dev, lens, H, seed):
    g = torch.Generator().manual_seed(seed)
    n = sum(lens)
    scores = (torch.randn((n, H), generator=g) * 8).round() / 8    # many ties: order pinned by flat index
    scores[5, 0] = float("nan")                                  # NaN ranks first (torch.topk order)
    scores[11, H - 1] = float("inf")
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    return scores.to(dev), cu.to(dev)


def _existing(L, N, scores, cu, B, H, max_k_len, top_k):
    dev = scores.device
    k_eff = min(top_k, max_k_len * H)
    out = torch.full((B, k_eff), -1, dtype=torch.int64, device=dev)
    ws_bytes = L.cvllm_rank_scores_workspace_bytes(B, H, max_k_len)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    rc = L.cvllm_rank_scores(N.ptr(scores), scores.stride(0), N.ptr(cu), B, H, max_k_len, top_k, N.ptr(out),
                             N.ptr(ws), ws.numel(), N.stream(dev))
    N.check(rc, "cvllm_rank_scores")
    return out


def test_static_scratch_ranking_equals_workspace_ranking_twice_in_a_row():
    from compactor_vllm import _native as N
    L = N.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    assert L.cvllm_rank_scores_static_max_bytes() == 8 << 20
    H = 8
    # two different problems back to back: the second call finds the first call's contents in the shared array
    cases = [([700, 300, 1], 2048), ([4096], 4096 * H), ([33, 2000], 900)]
    for rnd in range(2):
        for i, (lens, top_k) in enumerate(cases):
            B, max_k_len = len(lens), max(lens)
            scores, cu = _problem(dev, lens, H, seed=10 * rnd + i)
            want = _existing(L, N, scores, cu, B, H, max_k_len, top_k)
            out = torch.full_like(want, -1)
            torch.cuda.synchronize()
            before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
            rc = L.cvllm_rank_scores_static(N.ptr(scores), scores.stride(0), N.ptr(cu), B, H, max_k_len, top_k,
                                            N.ptr(out), N.stream(dev))
            N.check(rc, "cvllm_rank_scores_static")
            rc = L.cvllm_rank_scores_static(N.ptr(scores), scores.stride(0), N.ptr(cu), B, H, max_k_len, top_k,
                                            N.ptr(out), N.stream(dev))  # second call in a row, same stream
            N.check(rc, "cvllm_rank_scores_static")
            torch.cuda.synchronize()
            assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
            assert torch.equal(out, want), (rnd, i)


def test_static_scratch_metric_shape_and_size_limit():
    """One 32768-token sequence x 8 KV heads (the metric's shape) fits the 8 MiB array; a problem beyond it is refused
    by name, with the sizes in the error text, and nothing is launched."""
    from compactor_vllm import _native as N
    L = N.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    H, n = 8, 32768
    assert L.cvllm_rank_scores_workspace_bytes(1, H, n) <= L.cvllm_rank_scores_static_max_bytes()
    scores, cu = _problem(dev, [n], H, seed=99)
    want = _existing(L, N, scores, cu, 1, H, n, n * H // 2)
    out = torch.full_like(want, -1)
    rc = L.cvllm_rank_scores_static(N.ptr(scores), scores.stride(0), N.ptr(cu), 1, H, n, n * H // 2, N.ptr(out),
                                    N.stream(dev))
    N.check(rc, "cvllm_rank_scores_static")
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    # sortedness property at full size: scores along the ranking never increase (NaN first)
    flat = scores.flatten()[out[0]]
    finite = flat[~flat.isnan()]
    assert bool((finite[1:] <= finite[:-1]).all())

    big = 1 << 20  # 1 Mi tokens x 8 heads: 128 MiB of scratch
    launches = L.cvllm_launch_count()
    rc = L.cvllm_rank_scores_static(N.ptr(scores), scores.stride(0), N.ptr(cu), 1, H, big, 16, N.ptr(out),
                                    N.stream(dev))
    assert rc == -2  # CVLLM_E_UNSUPPORTED (include/cvllm.h)
    msg = L.cvllm_last_error().decode()
    assert "cvllm_rank_scores_static" in msg and "8388608" in msg
    assert L.cvllm_launch_count() == launches
