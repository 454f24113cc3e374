"""Row-parallel linear + the tensor-parallel all-reduce behind it (reference:
layers/linear.py:133-153 `RowParallelLinear`: `F.linear` then `dist.all_reduce(y)`).

`PeerAllReduce` is the B200 route for decode-size messages: every rank keeps one cudaIpc-shared staging
buffer; the row-parallel GEMM writes its partial straight into it (`out=`), and `cvllm_allreduce_oneshot`
reads all peers' partials over NVLink / NVSwitch and sums them in rank order — one kernel, no ring, identical
bits on every rank.  Messages above `max_bytes` (prefill) go to NCCL, which is bandwidth-optimal there
(NVLS in-switch reduction).  The process group only carries the 64-byte IPC handles at start-up.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from compactor_vllm import _native as N


class _RawCudaBytes:
    """`__cuda_array_interface__` view of a raw device pointer, so torch can wrap memory it did not allocate."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerAllReduce:
    def __init__(self, group=None, max_bytes: Optional[int] = None, device: Optional[torch.device] = None,
                 twoshot_min_bytes: Optional[int] = None):
        assert dist.is_initialized(), "PeerAllReduce needs an initialised process group"
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        assert self.world <= 8
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        # One-shot reads (world-1) x message per rank, two-shot 2 x (world-1)/world x message plus one publishing
        # barrier; every barrier is a system-scope release / acquire pair (csrc/allreduce.cu).  Measured per call
        # inside a CUDA graph, [rows, 8192] bf16, on one 8 x B200 box (profiles/r02f_tp_allreduce_sys_n{2,4,8}.json;
        # the relaxed-flag build of round 1 is 4-8 us faster per call, profiles/r02f_tp_allreduce_relaxed_n8.json):
        #   8 ranks: one-shot 12.8 / 18.8 / 26.7 / 38.5 us at 16 KB / 256 KB / 1 MB / 2 MB, two-shot 16.9 / 22.0 /
        #            26.6 / 28.3 us -> switch at 1 MB; 4 MB: 32.6 us (NCCL 55.9), 8 MB: 44.4 us (NCCL 73.2);
        #            NCCL at 16 KB .. 2 MB: 27.4 / 30.7 / 31.3 / 35.2 us
        #   4 ranks: one-shot 24.9 / 35.2 us at 2 / 4 MB, two-shot 25.7 / 31.2 us -> switch at 3 MB; 8 MB 41.6 (NCCL 49.0)
        #   2 ranks: both forms move the same bytes, the one-shot has one barrier less -> never two-shot; with the
        #            formal fences the one-shot only ties NCCL there (11.0 / 15.5 / 17.7 us vs 13.6 / 15.7 / 17.7)
        if max_bytes is None:
            max_bytes = 8 << 20
        if twoshot_min_bytes is None:
            twoshot_min_bytes = ((1 << 20) if self.world >= 8 else (3 << 20) if self.world >= 3 else (1 << 62))
        self.max_bytes = int(max_bytes)
        self.twoshot_min_bytes = int(twoshot_min_bytes)
        lib = N.lib()
        self.flag_bytes = int(lib.cvllm_allreduce_flag_bytes())
        own = C.c_void_p()
        handle = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            # payload area + result area of the two-shot form
            N.check(lib.cvllm_ipc_alloc(2 * self.max_bytes, C.byref(own), handle), "cvllm_ipc_alloc")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self._own = own.value
        self._bases = (C.c_void_p * self.world)()
        for r, h in enumerate(handles):
            if r == self.rank:
                self._bases[r] = self._own
            else:
                p = C.c_void_p()
                with torch.cuda.device(self.device):
                    N.check(lib.cvllm_ipc_open(C.create_string_buffer(h, 64), C.byref(p)), "cvllm_ipc_open")
                self._bases[r] = p.value
        raw = torch.as_tensor(_RawCudaBytes(self._own + self.flag_bytes, self.max_bytes), device=self.device)
        self._payload = raw  # uint8 view of this rank's staging area
        dist.barrier(group=group)  # every rank has mapped every buffer before the first kernel

    def fits(self, t: torch.Tensor) -> bool:
        return (t.dtype in (torch.bfloat16, torch.float16) and t.numel() % 8 == 0
                and t.numel() * t.element_size() <= self.max_bytes)

    def staging(self, shape, dtype) -> torch.Tensor:
        """A tensor of `shape` living in the shared buffer: let the producer GEMM write into it."""
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        assert nbytes <= self.max_bytes
        return self._payload[:nbytes].view(dtype).view(*shape)

    def reduce_staged(self, staged: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Sum over ranks of what each rank put in its staging area (same shape everywhere), written to `out`
        (a fresh tensor when None; `out` must not alias the staging area)."""
        if out is None:
            out = torch.empty(staged.shape, dtype=staged.dtype, device=self.device)
        assert out.is_contiguous() and out.dtype == staged.dtype and out.numel() == staged.numel()
        if staged.numel() * staged.element_size() >= self.twoshot_min_bytes:
            rc = N.lib().cvllm_allreduce_twoshot(self._bases, self.rank, self.world, N.ptr(out), staged.numel(),
                                                 N.dtype_code(staged.dtype), self.max_bytes, N.stream(self.device))
            N.check(rc, "cvllm_allreduce_twoshot")
            return out
        rc = N.lib().cvllm_allreduce_oneshot(self._bases, self.rank, self.world, N.ptr(out), staged.numel(),
                                             N.dtype_code(staged.dtype), N.stream(self.device))
        N.check(rc, "cvllm_allreduce_oneshot")
        return out

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        """`dist.all_reduce(t)` semantics on both routes: `t` holds the sum over ranks afterwards and is returned
        (the reference's in-place idiom, layers/linear.py:151).  Large or odd-shaped messages go to NCCL; the rest
        are copied into the shared staging area and summed by the peer-memory kernel straight back into `t`."""
        if not self.fits(t) or not t.is_contiguous():
            dist.all_reduce(t, group=self.group)
            return t
        st = self.staging(t.shape, t.dtype)
        st.copy_(t)
        return self.reduce_staged(st, out=t)

    def close(self) -> None:
        lib = N.lib()
        torch.cuda.synchronize(self.device)
        for r in range(self.world):
            if r != self.rank and self._bases[r]:
                lib.cvllm_ipc_close(self._bases[r])
        dist.barrier(group=self.group)
        lib.cvllm_ipc_free(self._own)
        self._own = None


class PeerGemmAllReduce:
    """`F.linear(x, W)` + `dist.all_reduce` of a row-parallel layer as ONE kernel per rank (csrc/gemm_reduce.cu,
    `cvllm_gemm_allreduce`): tcgen05 GEMM on CTA pairs whose epilogue TMA-stores each tile into the owning rank's
    memory over NVLink (reduce-scatter), reducer warps that sum a round of rows as soon as it has landed from every
    rank and store the result into y on every rank (all-gather).  The route for PREFILL-size outputs, where the
    separate all-reduce is 12-37 % of the layer.

    `max_rows x out_features` sizes the shared buffer (slots + y, 2 x rows x out x 2 bytes per rank).  The returned
    tensor is a view of that buffer, valid until the next call."""

    def __init__(self, max_rows: int, out_features: int, group=None, device: Optional[torch.device] = None,
                 _virtual: Optional[tuple] = None):
        lib = N.lib()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.max_rows, self.out_features = int(max_rows), int(out_features)
        self.flag_bytes = int(lib.cvllm_allreduce_flag_bytes())
        if _virtual is not None:  # (rank, world, bases): several ranks of ONE process on one GPU (tests)
            self.group = None
            self.rank, self.world, bases = _virtual
            self.capacity = int(lib.cvllm_gemm_allreduce_payload_bytes(self.max_rows, self.out_features, self.world))
            self._own = None
            self._bases = (C.c_void_p * self.world)(*bases)
            return
        assert dist.is_initialized(), "PeerGemmAllReduce needs an initialised process group"
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        assert self.world <= 8
        self.capacity = int(lib.cvllm_gemm_allreduce_payload_bytes(self.max_rows, self.out_features, self.world))
        assert self.capacity > 0
        own = C.c_void_p()
        handle = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            N.check(lib.cvllm_ipc_alloc(self.capacity, C.byref(own), handle), "cvllm_ipc_alloc")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self._own = own.value
        self._bases = (C.c_void_p * self.world)()
        for r, h in enumerate(handles):
            if r == self.rank:
                self._bases[r] = self._own
            else:
                ptr = C.c_void_p()
                with torch.cuda.device(self.device):
                    N.check(lib.cvllm_ipc_open(C.create_string_buffer(h, 64), C.byref(ptr)), "cvllm_ipc_open")
                self._bases[r] = ptr.value
        dist.barrier(group=group)  # every rank has mapped every buffer before the first kernel

    @staticmethod
    def virtual_ranks(world: int, max_rows: int, out_features: int, device: torch.device):
        """`world` ranks inside this process on ONE device (their kernels run side by side on `world` streams, each on
        1/world of the SMs — set the `grs_pairs` test option accordingly).  Returns (ranks, free)."""
        lib = N.lib()
        cap = int(lib.cvllm_gemm_allreduce_payload_bytes(max_rows, out_features, world))
        bases = []
        with torch.cuda.device(device):
            for _ in range(world):
                own, handle = C.c_void_p(), C.create_string_buffer(64)
                N.check(lib.cvllm_ipc_alloc(cap, C.byref(own), handle), "cvllm_ipc_alloc")
                bases.append(own.value)
        ranks = [PeerGemmAllReduce(max_rows, out_features, device=device, _virtual=(r, world, bases))
                 for r in range(world)]

        def free():
            torch.cuda.synchronize(device)
            for b in bases:
                lib.cvllm_ipc_free(b)
        return ranks, free

    def supports(self, x: torch.Tensor, weight: torch.Tensor) -> bool:
        rows = x.numel() // max(1, x.shape[-1])
        return (x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and weight.dtype == x.dtype
                and 0 < rows <= self.max_rows and weight.shape[0] == self.out_features
                and weight.shape[0] % 8 == 0 and weight.shape[1] % 8 == 0 and weight.stride(1) == 1
                and weight.stride(0) % 8 == 0)

    def linear(self, x: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """sum over ranks of `x @ weight.T` ([..., K_local] x [out, K_local]) -> [..., out], on every rank."""
        rows = x.reshape(-1, x.shape[-1])
        if rows.stride(1) != 1 or rows.stride(0) % 8 != 0:
            rows = rows.contiguous()
        m, k = rows.shape
        n = weight.shape[0]
        assert weight.shape[1] == k and self.supports(x, weight)
        y_ptr = C.c_void_p()
        rc = N.lib().cvllm_gemm_allreduce(N.ptr(rows), rows.stride(0), N.ptr(weight), weight.stride(0), m, n, k,
                                          self._bases, self.rank, self.world, N.dtype_code(x.dtype), self.capacity,
                                          C.byref(y_ptr), N.stream(self.device))
        N.check(rc, "cvllm_gemm_allreduce")
        raw = torch.as_tensor(_RawCudaBytes(y_ptr.value, m * n * 2), device=self.device)
        return raw.view(x.dtype).view(*x.shape[:-1], n)

    def close(self) -> None:
        if self._own is None:
            return
        lib = N.lib()
        torch.cuda.synchronize(self.device)
        for r in range(self.world):
            if r != self.rank and self._bases[r]:
                lib.cvllm_ipc_close(self._bases[r])
        dist.barrier(group=self.group)
        lib.cvllm_ipc_free(self._own)
        self._own = None


def shard_rows(full_weight: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Input-dimension shard of a [out, in] weight for tensor-parallel rank `rank` (reference :141-146)."""
    assert full_weight.shape[1] % world == 0
    per = full_weight.shape[1] // world
    return full_weight[:, rank * per:(rank + 1) * per]


class RowParallelLinear(nn.Module):
    """y = all_reduce(x_shard @ W_shard^T (+ bias on rank 0)) — reference layers/linear.py:133-153."""

    def __init__(self, input_size: int, output_size: int, bias: bool = False,
                 peer_all_reduce: Optional[PeerAllReduce] = None,
                 peer_gemm_all_reduce: Optional[PeerGemmAllReduce] = None):
        super().__init__()
        # the tensor-parallel group is the peer all-reduce's group when one is supplied (it may be a subgroup)
        group = peer_all_reduce.group if peer_all_reduce is not None else (
            peer_gemm_all_reduce.group if peer_gemm_all_reduce is not None else None)
        self.tp_group = group
        self.tp_rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.tp_size = dist.get_world_size(group) if dist.is_initialized() else 1
        assert input_size % self.tp_size == 0
        self.weight = nn.Parameter(torch.empty(output_size, input_size // self.tp_size))
        self.bias = nn.Parameter(torch.empty(output_size)) if bias else None
        self.peer = peer_all_reduce
        self.peer_gemm = peer_gemm_all_reduce  # fused GEMM + reduce-scatter + all-gather for prefill-size outputs

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        bias = self.bias if self.tp_rank == 0 else None
        if self.tp_size == 1:
            return F.linear(x, self.weight, bias)
        rows = x.reshape(-1, x.shape[-1])
        out_shape = (*x.shape[:-1], self.weight.shape[0])
        if self.peer is not None and x.is_cuda and bias is None and self.bias is None:
            probe = torch.empty(0, dtype=x.dtype)
            if rows.shape[0] * self.weight.shape[0] * probe.element_size() <= self.peer.max_bytes \
                    and (rows.shape[0] * self.weight.shape[0]) % 8 == 0 and x.dtype in (torch.bfloat16, torch.float16):
                st = self.peer.staging((rows.shape[0], self.weight.shape[0]), x.dtype)
                torch.mm(rows, self.weight.t(), out=st)  # the partial lands in the shared buffer: no copy
                return self.peer.reduce_staged(st).view(out_shape)
        if self.peer_gemm is not None and bias is None and self.bias is None and self.peer_gemm.supports(x, self.weight):
            return self.peer_gemm.linear(x, self.weight)  # one kernel: GEMM, reduce-scatter and all-gather
        y = F.linear(x, self.weight, bias)
        dist.all_reduce(y, group=self.tp_group)
        return y
