"""The scripted allocator op sequence shared by oracle/gen_golden.py (which runs it on the reference) and the
tests (which replay it on the oracle, the host allocator and the device allocator).  TEST INFRASTRUCTURE ONLY."""
import torch

ALLOC_SCRIPT = [
    # (op, slot handle, argument) — slot handles name the return values of "new"
    ("new", "a", None), ("reserve", "a", 40), ("lens", "a", 11), ("reclaim", "a", 5),
    ("new", "b", None), ("reserve", "b", 70), ("reserve", "a", 30), ("reserve", "b", 200),
    ("new", "c", None), ("reserve", "c", 90), ("free", "a", None), ("reserve", "c", 60),
    ("lens", "b", 12), ("reclaim", "b", 0), ("new", "d", None), ("reserve", "d", 17),
    ("lens", "d", 13), ("reserve", "d", 40), ("free", "c", None), ("reserve", "b", 33),
    ("new", "e", None), ("new", "f", None), ("free", "b", None), ("free", "d", None),
]
ALLOC_SHAPE = dict(num_layers=3, max_logical_pages_per_head=6, num_pages=44, page_size=16, H_kv=4,
                   head_dim=8, max_num_batches=4)


def alloc_lens(seed: int, L: int, H: int, cap_pages):
    """deterministic per-(layer, head) lengths <= the head's capacity (what a compaction leaves behind)"""
    g = torch.Generator().manual_seed(seed)
    return (torch.rand((L, H), generator=g) * cap_pages.float() * 16).floor().to(torch.int32)
