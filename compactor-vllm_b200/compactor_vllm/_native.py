"""ctypes binding of libcompactor_sm100.so (C ABI in include/cvllm.h).

There is deliberately NO fallback: if the library is missing or a call fails the
caller gets an exception.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_PKG_ROOT = Path(__file__).resolve().parent.parent
LIB_PATH = Path(os.environ.get("CVLLM_LIB", _PKG_ROOT / "lib" / "libcompactor_sm100.so"))

CVLLM_F16, CVLLM_BF16, CVLLM_F32_ACCUM = 0, 1, 2

_i, _i64, _f, _p, _sz, _u32 = C.c_int, C.c_int64, C.c_float, C.c_void_p, C.c_size_t, C.c_uint32

_SIGNATURES = {
    "cvllm_abi_version": (C.c_int, []),
    "cvllm_last_error": (C.c_char_p, []),
    "cvllm_launch_count": (C.c_uint64, []),
    "cvllm_set_test_option": (_i, [C.c_char_p, C.c_double]),
    "cvllm_selftest_eig": (_i, [_p, _p, _p, _p, _i, _i, _p, _p]),
    "cvllm_decode_store_kv": (_i, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p,
                                   _i, _i, _i, _i, _i, _i, _i, _p]),
    "cvllm_prefill_store_all": (_i, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p,
                                     _i, _i, _i, _i, _i, _i, _i, _p]),
    "cvllm_scatter_pages": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p]),
    "cvllm_pages_bitmap_words": (_sz, [_i]),
    "cvllm_pages_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "cvllm_pages_init": (_i, [_p, _p, _i, _i, _p]),
    "cvllm_pages_reserve": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _p, _i, _i, _i, _i, _i, _i, _i, _p, _sz, _p]),
    "cvllm_pages_release": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _p, _i, _i, _i, _i, _i, _i, _i, _p]),
    "cvllm_debug_decode_trace": (_i, [_p]),
    "cvllm_allreduce_flag_bytes": (_sz, []),
    "cvllm_ipc_alloc": (_i, [_sz, _p, _p]),
    "cvllm_ipc_open": (_i, [_p, _p]),
    "cvllm_ipc_close": (_i, [_p]),
    "cvllm_ipc_free": (_i, [_p]),
    "cvllm_allreduce_oneshot": (_i, [_p, _i, _i, _p, _i64, _i, _p]),
    "cvllm_allreduce_twoshot": (_i, [_p, _i, _i, _p, _i64, _i, _i64, _p]),
    "cvllm_gemm_allreduce_payload_bytes": (_i64, [_i64, _i64, _i]),
    "cvllm_gemm_allreduce": (_i, [_p, _i64, _p, _i64, _i64, _i64, _i64, _p, _i, _i, _i, _i64, _p, _p]),
    "cvllm_rank_scores_workspace_bytes": (_sz, [_i, _i, _i]),
    "cvllm_rank_scores": (_i, [_p, _i64, _p, _i, _i, _i, _i, _p, _p, _sz, _p]),
    "cvllm_rank_scores_static_max_bytes": (_sz, []),
    "cvllm_rank_scores_static": (_i, [_p, _i64, _p, _i, _i, _i, _i, _p, _p]),
    "cvllm_store_topk_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "cvllm_prefill_store_topk_idx": (_i, [_p, _p, _i64, _i64, _i64, _i64, _p, _i, _i, _p, _p, _p,
                                          _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _i,
                                          _p, _sz, _p]),
    "cvllm_select_store_workspace_bytes": (_sz, [_i, _i, _i]),
    "cvllm_select_store_topk": (_i, [_p, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p,
                                     _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _i, _p, _sz, _p]),
    "cvllm_decode_attention_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "cvllm_decode_attention": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _i, _i,
                                    _f, _i, _i, _p]),
    "cvllm_decode_step": (_i, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _sz,
                               _i, _i, _i, _i, _i, _i, _f, _i, _i, _i, _p]),
    "cvllm_prefill_attention": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p,
                                     _i, _i, _i, _i, _i, _i, _i, _i64, _i64, _i64, _i64, _i64, _i,
                                     _f, _i, _p]),
    "cvllm_leverage_scores_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "cvllm_leverage_scores": (_i, [_p, _i64, _i64, _p, _i, _i, _p, _p, _i, _i, _i, _i, _f, _i, _i,
                                   _p, _sz, _p]),
    "cvllm_leverage_scores_from_sketch": (_i, [_p, _p, _i, _i, _p, _p, _i, _i, _i, _f, _i, _i, _p, _sz, _p]),
    "cvllm_rope_qk": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _p, _p, _i,
                           _p, _p, _f, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "cvllm_rmsnorm": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i, _i, _f, _i, _p]),
    "cvllm_chunk_attn_scores_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "cvllm_chunk_attn_scores": (_i, [_p, _p, _i64, _i64, _p, _i, _i, _i, _i, _i, _i, _i, _f, _i,
                                     _p, _i, _i64, _f, _p, _p, _p, _i, _p, _sz, _p]),
    "cvllm_snapkv_scores_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "cvllm_snapkv_scores": (_i, [_p, _p, _i64, _i64, _p, _p, _i, _i, _i, _i, _i, _i, _i, _f, _i,
                                 _i, _p, _i, _p, _sz, _p]),
    "cvllm_selftest_umma": (_i, [_p, _p, _p, _i, _u32, _u32, _i, _p]),
}

_lib = None


class NativeLibraryError(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raise loudly if absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise NativeLibraryError(
                f"{LIB_PATH} not found: build it with `python compactor-vllm_b200/build.py` "
                "(there is no CPU or PyTorch fallback for this path)"
            )
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the ABI drifted
            fn.restype = res
            fn.argtypes = args
        if handle.cvllm_abi_version() != 1:
            raise NativeLibraryError("libcompactor_sm100.so ABI version mismatch")
        _lib = handle
    return _lib


def set_test_option(name: str, value: float) -> None:
    """Test / profiling switch of the library (include/cvllm.h: cvllm_set_test_option)."""
    check(lib().cvllm_set_test_option(name.encode(), float(value)), f"cvllm_set_test_option({name})")


def exported_symbols() -> list[str]:
    return list(_SIGNATURES)


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().cvllm_last_error().decode(errors="replace")
        raise NativeLibraryError(f"{what} failed with code {rc}: {msg}")


def ptr(t: torch.Tensor | None) -> int | None:
    if t is None:
        return None
    assert t.is_cuda, "cvllm kernels need CUDA tensors (no CPU fallback exists)"
    return t.data_ptr()


def stream(device: torch.device | None = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.bfloat16:
        return CVLLM_BF16
    if dt == torch.float16:
        return CVLLM_F16
    raise TypeError(f"cvllm kernels support float16/bfloat16, got {dt}")


_WORKSPACES: dict = {}


def workspace(device: torch.device, nbytes: int, tag: str = "", zero: bool = False) -> torch.Tensor:
    """Grow-only per-(device, stream, tag) byte scratch from torch's caching allocator.

    zero=True: the buffer is zero-filled when it is (re)allocated — for kernels whose contract is
    "control words zero on entry, restored to zero on exit" (decode attention's head tickets)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, tag)
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < nbytes:
        alloc = torch.zeros if zero else torch.empty
        buf = alloc(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = buf
    return buf
