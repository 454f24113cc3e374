"""Oracle for the page allocator (kv_cache/page_table.py:72-291 of the reference).  TEST
INFRASTRUCTURE ONLY.  Pure Python ints and lists: per-layer min-heaps of free physical pages,
slot 0 reserved, all-or-nothing reservation across layers, reclaim down to ceil((len+future)/PS).

Pinned by tests/golden/allocator_trace.npz: oracle/gen_golden.py drives the reference's own
PagedKVCache on CPU (its Triton page scatter under TRITON_INTERPRET=1) through a scripted sequence of
new_batch / reserve_tokens / reclaim_pages / free_batch calls and records the state after each.
"""
from __future__ import annotations

import heapq

SUCCESS, EXCEEDS_MAX_SEQUENCE_LENGTH, EXCEEDS_CURRENTLY_AVAILABLE_PAGES = 0, 1, 2


def cdiv(a: int, b: int) -> int:
    return -(-a // b)


class Allocator:
    def __init__(self, num_layers, max_pages_per_head, num_pages, page_size, H_kv, max_num_batches):
        self.L, self.P, self.PS, self.H = num_layers, max_pages_per_head, page_size, H_kv
        self.B = max_num_batches + 1  # page_table.py:84 (+ the reserved slot)
        self.free = [list(range(num_pages)) for _ in range(num_layers)]  # :111-115 (sorted = a heap)
        self.free_batches = [b for b in reversed(range(self.B)) if b != 0]  # :117-119
        self.lens = [[[0] * H_kv for _ in range(self.B)] for _ in range(num_layers)]
        self.pages = [[[0] * H_kv for _ in range(self.B)] for _ in range(num_layers)]
        # page_table[l][b][h] = list of physical pages in logical order (only the first pages[l][b][h] are defined)
        self.table = [[[[] for _ in range(H_kv)] for _ in range(self.B)] for _ in range(num_layers)]

    def new_batch(self):  # :125-142
        if self.free_batches and all(len(f) >= self.H for f in self.free):
            return self.free_batches.pop()
        return None

    def reserve_tokens(self, b: int, add_tokens: int) -> int:  # :144-198
        add = [[0] * self.H for _ in range(self.L)]
        short = False
        for l in range(self.L):
            for h in range(self.H):
                missing = self.lens[l][b][h] + add_tokens - self.pages[l][b][h] * self.PS
                short |= missing > 0
                # the reference does not clamp (:164-166); the engine only reserves when no head has a spare
                # page, so clamping at 0 is the same on every reachable state
                add[l][h] = max(cdiv(missing, self.PS), 0)
        if not short:
            return SUCCESS
        if any(self.pages[l][b][h] + add[l][h] > self.P for l in range(self.L) for h in range(self.H)):
            return EXCEEDS_MAX_SEQUENCE_LENGTH
        if any(sum(add[l]) > len(self.free[l]) for l in range(self.L)):
            return EXCEEDS_CURRENTLY_AVAILABLE_PAGES
        for l in range(self.L):
            for h in range(self.H):  # scatter order: heads in order take consecutive pops (write_page_table.py:62-95)
                for _ in range(add[l][h]):
                    self.table[l][b][h].append(heapq.heappop(self.free[l]))
                self.pages[l][b][h] += add[l][h]
        return SUCCESS

    def reclaim_pages(self, b: int, future_reserve_tokens: int = 0) -> int:  # :200-267 -> pages freed
        freed = 0
        for l in range(self.L):
            for h in range(self.H):
                used = min(cdiv(self.lens[l][b][h] + future_reserve_tokens, self.PS), self.pages[l][b][h])
                for page in self.table[l][b][h][used:]:
                    heapq.heappush(self.free[l], page)
                    freed += 1
                del self.table[l][b][h][used:]
                self.pages[l][b][h] = used
        return freed

    def free_batch(self, b: int) -> None:  # :269-291
        for l in range(self.L):
            for h in range(self.H):
                for page in self.table[l][b][h]:
                    heapq.heappush(self.free[l], page)
                self.table[l][b][h] = []
                self.lens[l][b][h] = 0
                self.pages[l][b][h] = 0
        self.free_batches.append(b)

    # -- state in the fixture's format --
    def snapshot(self):
        import numpy as np
        pt = np.full((self.L, self.B, self.H, self.P), -1, dtype=np.int32)
        for l in range(self.L):
            for b in range(self.B):
                for h in range(self.H):
                    row = self.table[l][b][h]
                    pt[l, b, h, :len(row)] = row
        free = np.full((self.L, max(len(f) for f in self.free) if any(self.free) else 0), -1, dtype=np.int32)
        for l in range(self.L):
            free[l, :len(self.free[l])] = sorted(self.free[l])
        return pt, np.array(self.pages, dtype=np.int32), np.array(self.lens, dtype=np.int32), free
