"""Host-side logic of the path against values produced by the REFERENCE's own functions
(oracle/gen_golden.py `splits` / `memory`): no GPU, no compute calls."""
import dataclasses

import pytest

from conftest import load_golden


def test_num_splits_heuristic_matches_reference():
    """attention/sparse_decode_kernel.py:168-192 over (b x kv-head blocks, context, SMs, split cap)."""
    from compactor_vllm.attention.sparse_decode_kernel import num_splits_heuristic

    table = load_golden("num_splits")["table"]
    assert table.shape[0] > 2000
    for blocks, seq, sms, cap, want in table.tolist():
        assert num_splits_heuristic(blocks, seq, sms, cap) == want, (blocks, seq, sms, cap)
    # the engine's capture-time call on this part: B=1, 8 kv heads, 32K context, 148 SMs
    assert num_splits_heuristic(8, 32768, 148, 64) >= 2


def test_kv_page_budget_matches_reference():
    """core/memory_manager.py:92-126 on scripted allocator readings (incl. the two RuntimeError exits)."""
    from compactor_vllm.core.memory_manager import MemorySnapshot, kv_page_budget

    g = load_golden("memory_budget")
    seen_error = seen_pages = 0
    for total, used, peak, cur, pct, layers, heads, hd, ps, esz, nb, nlp, want in g["pages"].tolist():
        snap = MemorySnapshot(total, total - used, peak, cur)
        kw = dict(num_layers=layers, heads=heads, head_dim=hd, page_size=ps, elem_size=esz, max_batches=nb,
                  logical_pages=nlp)
        if want < 0:
            with pytest.raises(RuntimeError):
                kv_page_budget(snap, pct / 100, **kw)
            seen_error += 1
        else:
            assert kv_page_budget(snap, pct / 100, **kw) == want
            seen_pages += 1
    assert seen_error > 0 and seen_pages > 0


def test_activation_token_budget_matches_reference():
    """core/memory_manager.py:128-161."""
    from compactor_vllm.core.memory_manager import MemorySnapshot, activation_token_budget

    rows = load_golden("memory_budget")["tokens"].tolist()
    assert rows
    for total, used, peak, cur, pct, heads, ps, pages, wt, before, after, want in rows:
        snap = MemorySnapshot(total, total - used, peak, cur)
        got = activation_token_budget(snap, pct / 100, warmup_tokens=wt, used_before=before, peak_after=after,
                                      num_pages=pages, page_size=ps, heads=heads)
        assert got == want


def test_context_records_are_dataclasses():
    """utils/context.py:10-40: `dataclasses.replace / asdict / fields` are legal on the reference's records."""
    from compactor_vllm.utils.context import CompressionContext, Context, get_context, reset_context, set_context

    names = [f.name for f in dataclasses.fields(Context)]
    assert names == ["is_prefill", "do_compression", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k",
                     "batch_mapping", "max_bh_len", "compression_context", "STORE_STREAM", "key_split",
                     "attention_backend"]
    assert [f.name for f in dataclasses.fields(CompressionContext)] == [
        "compression_method", "compression_chunk_size", "batch_tokens_to_retain", "max_tokens_to_retain",
        "context_lens", "PHI", "protected_first_tokens", "protected_last_tokens"]
    c = Context(True, True, max_bh_len=7)
    assert dataclasses.replace(c, max_bh_len=9).max_bh_len == 9 and c.max_bh_len == 7
    assert dataclasses.asdict(CompressionContext(context_lens=[3]))["context_lens"] == [3]
    with pytest.raises(TypeError):
        Context(no_such_field=1)
    set_context(is_prefill=True, max_seqlen_q=4)
    assert get_context().is_prefill and get_context().max_seqlen_q == 4
    reset_context()
    assert not get_context().is_prefill


def test_slot_map_reuses_and_releases():
    from compactor_vllm.core.memory_manager import SlotMap

    class FakeCache:
        def __init__(self):
            self.free, self.freed = [3, 2, 1], []

        def new_batch(self):
            return self.free.pop() if self.free else None

        def free_batch(self, slot):
            self.freed.append(slot)

    cache, slots = FakeCache(), SlotMap()
    assert [slots.acquire(cache, s) for s in (10, 11, 10, 12, 13)] == [1, 2, 1, 3, None]
    slots.release(cache, 11)
    assert cache.freed == [2] and 11 not in slots.by_seq


def test_shape_envelope_is_checked_at_construction():
    """D != 128 etc. are refused by name before any kernel is launched."""
    from types import SimpleNamespace as NS

    from compactor_vllm.core.model_runner import _check_shape_envelope

    cfg = NS(kvcache_page_size=128, leverage_sketch_size=48)
    _check_shape_envelope(NS(head_dim=128, hidden_size=4096, num_attention_heads=32, num_key_value_heads=8), cfg)
    _check_shape_envelope(NS(head_dim=None, hidden_size=8192, num_attention_heads=64, num_key_value_heads=8), cfg)
    with pytest.raises(ValueError, match="head_dim 64"):
        _check_shape_envelope(NS(head_dim=64, hidden_size=2048, num_attention_heads=32, num_key_value_heads=8), cfg)
    with pytest.raises(ValueError, match="GQA group 16"):
        _check_shape_envelope(NS(head_dim=128, hidden_size=4096, num_attention_heads=32, num_key_value_heads=2), cfg)
    with pytest.raises(ValueError, match="page_size 64"):
        _check_shape_envelope(NS(head_dim=128, hidden_size=4096, num_attention_heads=32, num_key_value_heads=8),
                              NS(kvcache_page_size=64, leverage_sketch_size=48))
