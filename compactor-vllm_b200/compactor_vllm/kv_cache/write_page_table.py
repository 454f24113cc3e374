"""Page-table append (reference: kv_cache/write_page_table.py:6-95)."""
import torch
from compactor_vllm import _native as N


def scatter_to_page_table(
    add_pages: torch.Tensor,  # [L, H] int32
    new_phys_pages: torch.Tensor,  # [N]
    curr_pages: torch.Tensor,  # [L, H] int32
    page_table: torch.Tensor,  # [L, H, max_pages_per_head] int32, any strides
    max_pages_per_head: int,
):
    """page_table[l, h, curr[l,h] + i] = new_phys[cum[l,h] + i] for i < add[l,h].

    The exclusive prefix sum the reference computes with torch.cumsum (:42-44)
    happens inside the kernel, so the call is a single launch.
    """
    L, H = add_pages.shape
    if L == 0 or H == 0:
        return
    assert page_table.dtype == torch.int32
    add_flat = add_pages.to(torch.int32).contiguous().view(-1)
    curr_flat = curr_pages.to(torch.int32).contiguous().view(-1)
    new_phys = new_phys_pages.to(torch.int32).contiguous()
    if new_phys.numel() == 0:
        return
    s_l, s_h, s_p = page_table.stride()
    rc = N.lib().cvllm_scatter_pages(
        N.ptr(add_flat), N.ptr(new_phys), N.ptr(curr_flat), N.ptr(page_table),
        s_l, s_h, s_p, L, H, int(max_pages_per_head), N.stream(page_table.device),
    )
    N.check(rc, "cvllm_scatter_pages")
