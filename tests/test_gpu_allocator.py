"""Page allocator parity: the host allocator (`PagedKVCache`) and the
device-side allocator (`DevicePagedKVCache`, csrc/page_alloc.cu) against the trace of the REFERENCE's
PagedKVCache (tests/golden/allocator_trace.npz, oracle/gen_golden.py) — return values, page tables,
page counts and the free set after every operation, exact — plus randomised host-vs-device agreement,
batched device reservations, CUDA-graph capture and conservation at full size."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from test_oracle_golden import replay_allocator_script

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None

STATUS_CODE = {"SUCCESS": 0, "EXCEEDS_MAX_SEQUENCE_LENGTH": 1, "EXCEEDS_CURRENTLY_AVAILABLE_PAGES": 2}


def _classes():
    from compactor_vllm.kv_cache.page_table import DevicePagedKVCache, PagedKVCache
    return {"host": PagedKVCache, "device": DevicePagedKVCache}


def _state(cache):
    P = cache.page_table.shape[-1]
    valid = torch.arange(P, device=DEV).view(1, 1, 1, P) < cache.bh_num_pages.unsqueeze(-1)
    pt = torch.where(valid, cache.page_table, torch.full_like(cache.page_table, -1)).cpu().numpy()
    free = [sorted(f) for f in cache.free_pages]
    return pt, cache.bh_num_pages.cpu().numpy(), cache.bh_seq_lens.cpu().numpy(), free


@pytest.mark.parametrize("kind", ["host", "device"])
def test_allocator_replays_reference_trace(kind):
    from oracle.gen_golden_script import ALLOC_SCRIPT, ALLOC_SHAPE, alloc_lens

    g = load_golden("allocator_trace")
    L, H = ALLOC_SHAPE["num_layers"], ALLOC_SHAPE["H_kv"]
    cls = _classes()[kind]

    def make(shape):
        return cls(dtype=torch.float16, device=DEV, **shape)

    def set_lens(cache, slot, seed):
        cache.bh_seq_lens[:, slot] = alloc_lens(seed, L, H, cache.bh_num_pages[:, slot].cpu()).to(DEV)

    for step, cache, ret in replay_allocator_script(make, set_lens):
        if ALLOC_SCRIPT[step][0] == "reserve":
            ret = STATUS_CODE[ret.name]
        assert ret == int(g["rets"][step]), (step, ALLOC_SCRIPT[step], ret)
        pt, pages, lens, free = _state(cache)
        assert np.array_equal(pt, g["page_table"][step].numpy()), step
        assert np.array_equal(pages, g["bh_num_pages"][step].numpy()), step
        assert np.array_equal(lens, g["bh_seq_lens"][step].numpy()), step
        gf = g["free_pages"][step].numpy()
        for l in range(L):
            assert free[l] == [int(x) for x in gf[l] if x >= 0], (step, l)


def test_device_allocator_matches_host_on_random_ops():
    cls = _classes()
    shape = dict(num_layers=4, max_logical_pages_per_head=12, num_pages=700, page_size=32, H_kv=8, head_dim=16,
                 max_num_batches=9)
    host = cls["host"](dtype=torch.bfloat16, device=DEV, **shape)
    dev = cls["device"](dtype=torch.bfloat16, device=DEV, **shape)
    rng = np.random.default_rng(5)
    live = []
    for step in range(120):
        op = rng.choice(["new", "reserve", "shrink", "reclaim", "free"], p=[0.2, 0.35, 0.2, 0.15, 0.1])
        if op == "new":
            a, b = host.new_batch(), dev.new_batch()
            assert a == b
            if a is not None:
                live.append(a)
        elif live:
            s = int(rng.choice(live))
            if op == "reserve":
                n = int(rng.integers(1, 200))
                # the engine only reserves when no head has a spare page; skip states the reference mishandles
                spare = host.bh_num_pages[:, s] * 32 - host.bh_seq_lens[:, s]
                if bool((spare >= n).any()) and not bool((spare >= n).all()):
                    continue
                ra, rb = host.reserve_tokens(s, n), dev.reserve_tokens(s, n)
                assert ra.name == rb.name, (step, ra, rb)
                if ra.name == "SUCCESS":  # tokens arrive
                    host.bh_seq_lens[:, s] += n
                    dev.bh_seq_lens[:, s] += n
            elif op == "shrink":  # a compaction leaves ragged per-head lengths behind
                keep = torch.from_numpy(rng.random((4, 8))).to(DEV)
                new = (host.bh_seq_lens[:, s].double() * keep).floor().to(torch.int32)
                host.bh_seq_lens[:, s] = new
                dev.bh_seq_lens[:, s] = new
            elif op == "reclaim":
                fut = int(rng.integers(0, 40))
                assert host.reclaim_pages(s, fut) == dev.reclaim_pages(s, fut)
            elif op == "free":
                host.free_batch(s)
                dev.free_batch(s)
                live.remove(s)
        a, b = _state(host), _state(dev)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]), step
        assert a[3] == b[3], step
        assert dev.free_count.tolist() == [len(f) for f in b[3]]


def test_batched_device_reserve_equals_sequential_host_and_skips_reserved_slot():
    cls = _classes()
    shape = dict(num_layers=3, max_logical_pages_per_head=8, num_pages=150, page_size=16, H_kv=4, head_dim=16,
                 max_num_batches=6)
    host = cls["host"](dtype=torch.float16, device=DEV, **shape)
    dev = cls["device"](dtype=torch.float16, device=DEV, **shape)
    slots = [host.new_batch() for _ in range(5)]
    assert slots == [dev.new_batch() for _ in range(5)]
    # 5 sequences x 3 pages x 4 heads = 60 per layer of 150; the 3rd wants more than a head may hold
    want = [host.reserve_tokens(s, 40) for s in slots[:2]]
    st = dev.reserve_tokens_device(torch.tensor(slots[:2] + [0], dtype=torch.int32, device=DEV), 40)
    assert st.tolist() == [0, 0, 0] and all(w.name == "SUCCESS" for w in want)
    for c in (host, dev):
        c.bh_seq_lens[:, slots[0]] = 40
        c.bh_seq_lens[:, slots[1]] = 33
    order = [slots[2], slots[0], 0, slots[3], slots[1], slots[4]]
    want = [host.reserve_tokens(s, 100) if s else None for s in order]
    st = dev.reserve_tokens_device(torch.tensor(order, dtype=torch.int32, device=DEV), 100).tolist()
    for s, w, got in zip(order, want, st):
        assert got == (0 if w is None else STATUS_CODE[w.name]), (s, w, got)
    assert 1 in st or 2 in st  # the script must exercise a failure inside a batch
    a, b = _state(host), _state(dev)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[3] == b[3]


def test_repeated_slot_in_one_call_is_served_once():
    """A slot named twice in one batched reserve / reclaim behaves like two sequential host calls (the second is a
    no-op): free_count stays equal to the bitmap's popcount and no garbage page id reaches the page table."""
    cls = _classes()
    shape = dict(num_layers=2, max_logical_pages_per_head=8, num_pages=64, page_size=16, H_kv=4, head_dim=16,
                 max_num_batches=4)
    host = cls["host"](dtype=torch.float16, device=DEV, **shape)
    dev = cls["device"](dtype=torch.float16, device=DEV, **shape)
    a, b = host.new_batch(), host.new_batch()
    assert (a, b) == (dev.new_batch(), dev.new_batch())
    order = [a, b, a, a, b]
    for s in order:
        assert host.reserve_tokens(s, 40).name == "SUCCESS"
    st = dev.reserve_tokens_device(torch.tensor(order, dtype=torch.int32, device=DEV), 40)
    assert st.tolist() == [0] * 5
    x, y = _state(host), _state(dev)
    assert np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1]) and x[3] == y[3]
    assert dev.free_count.tolist() == [len(f) for f in y[3]]
    for c in (host, dev):
        c.bh_seq_lens[:, a] = 5
    for s in (a, a):
        host.reclaim_pages(s, 0)
    dev.reclaim_pages_device(torch.tensor([a, a], dtype=torch.int32, device=DEV), 0)
    x, y = _state(host), _state(dev)
    assert np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1]) and x[3] == y[3]
    assert dev.free_count.tolist() == [len(f) for f in y[3]]


def test_decode_reservation_in_a_cuda_graph():
    """reserve 1 token for the running slots + `bh_seq_lens += 1`, captured once and replayed: pages grow
    exactly as the host allocator's would, with no host sync inside the step."""
    cls = _classes()
    shape = dict(num_layers=2, max_logical_pages_per_head=6, num_pages=64, page_size=8, H_kv=2, head_dim=16,
                 max_num_batches=4)
    host = cls["host"](dtype=torch.float16, device=DEV, **shape)
    dev = cls["device"](dtype=torch.float16, device=DEV, **shape)
    slots = [host.new_batch() for _ in range(3)]
    assert slots == [dev.new_batch() for _ in range(3)]
    slots_dev = torch.tensor(slots + [0], dtype=torch.int32, device=DEV)  # graph padding row -> slot 0
    status = torch.zeros(4, dtype=torch.int32, device=DEV)
    one = torch.zeros_like(dev.bh_seq_lens)
    one[:, slots] = 1

    def step():
        dev.reserve_tokens_device(slots_dev, 1, status)
        dev.bh_seq_lens.add_(one)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()  # warm-up: workspace allocation happens outside the capture
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            step()
    torch.cuda.current_stream().wait_stream(side)
    n_replays = 30
    for _ in range(n_replays):
        graph.replay()
    torch.cuda.synchronize()
    for _ in range(n_replays + 1):
        for s in slots:
            assert host.reserve_tokens(s, 1).name == "SUCCESS"
            host.bh_seq_lens[:, s] += 1
    a, b = _state(host), _state(dev)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    assert a[3] == b[3] and status.tolist() == [0, 0, 0, 0]
    # out of room: 6 pages x 8 tokens = 48 per head; keep replaying until the kernel reports it
    for _ in range(30):
        graph.replay()
    assert status.tolist()[:3] == [1, 1, 1] and int(dev.bh_num_pages.max()) == 6


def test_conservation_at_full_size():
    """Llama-3.1-8B sized tables (32 layers, 8 kv heads, 40960-token sequences = 320 pages per head):
    whatever is reserved comes back — free counts and bitmaps return to the initial state."""
    cls = _classes()["device"]
    L, H, P, PS, n_pages = 32, 8, 320, 128, 12000
    c = cls(num_layers=L, max_logical_pages_per_head=P, num_pages=n_pages, page_size=PS, H_kv=H, head_dim=8,
            max_num_batches=8, dtype=torch.bfloat16, device=DEV)
    bits0 = c.free_bits.clone()
    slots = [c.new_batch() for _ in range(4)]
    st = c.reserve_tokens_device(torch.tensor(slots, dtype=torch.int32, device=DEV), 40960).tolist()
    assert st == [0, 0, 0, 0]  # 4 x 8 x 320 = 10240 pages per layer
    assert c.free_count.tolist() == [n_pages - 10240] * L
    assert c.reserve_tokens(c.new_batch(), 40960).name == "EXCEEDS_CURRENTLY_AVAILABLE_PAGES"
    # every page handed out exactly once per layer
    rows = c.page_table[:, slots].reshape(L, -1)
    assert all(len(set(r)) == 10240 for r in rows.tolist())
    # first slot's first head got the smallest ids in order (min-heap pops)
    assert c.page_table[0, slots[0], 0].tolist() == list(range(320))
    c.bh_seq_lens[:, slots] = 40960
    c.bh_seq_lens[:, slots[1]] = 20000  # after a 50 % compaction
    freed = c.reclaim_pages(slots[1], 0)
    assert freed == L * H * (320 - 157) * PS * 8 * 2 * 2
    for s in slots:
        c.free_batch(s)
    torch.cuda.synchronize()
    assert c.free_count.tolist() == [n_pages] * L and torch.equal(c.free_bits, bits0)
    assert int(c.bh_num_pages.abs().sum()) == 0 and int(c.bh_seq_lens.abs().sum()) == 0
