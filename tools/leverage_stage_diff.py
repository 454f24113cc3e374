"""Stage-by-stage BIT comparison of this repo's leverage chain with the reference's torch arithmetic on the B200.

Which intermediate of `approximate_leverage_scores` (reference compression/compactor.py:153-214) first differs
from cuBLAS / torch: the sketch X = K.PHI (bf16), the chunk mean (bf16), the centred chunk, or the Gram (bf16)?
Ours come from the library's workspace (mean, Gram) and from the fused norm+RoPE kernel's sketch output (same
fragment order as the leverage kernel, tests/test_gpu_prepath.py); theirs are the reference's torch calls restated
on the same device (torch.matmul / mean / sub_, same shapes, strides and batching as the reference issues).
-> gpurun_out/r02_leverage_stage_diff.json
"""
from __future__ import annotations

import json
import math
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tools")]
import torch  # noqa: E402
from compactor_vllm import _native as N  # noqa: E402
from compactor_vllm.compression.compactor import approximate_leverage_scores  # noqa: E402
from compactor_vllm.layers.rotary_embedding import build_cos_sin_cache, fused_qk_norm_rope  # noqa: E402
from leverage_selfdistance import _keys  # noqa: E402

HQ, HKV, D, SK, CH = 32, 8, 128, 48, 512


def align(x):
    return (x + 255) & ~255


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n_tok = 32768
    phi = (torch.randn((D, SK), generator=torch.Generator(device=dev).manual_seed(42), device=dev)
           .to(torch.bfloat16) / math.sqrt(SK))
    rep = {}
    for kind in ("structured", "white"):
        q, k, v, k_pre = _keys(torch, kind, n_tok, dev)
        out = approximate_leverage_scores(k_pre, [n_tok], phi, normalize=False, chunk_size=CH)
        torch.cuda.synchronize()
        n_chunks = n_tok // CH
        mats = n_chunks * HKV
        ws = N.workspace(dev, N.lib().cvllm_leverage_scores_workspace_bytes(n_tok, HKV, SK, n_chunks), "leverage")
        off = align(n_tok * HKV * 4)
        gram = ws[off:off + mats * SK * SK * 4].view(torch.float32).view(n_chunks, HKV, SK, SK).clone()
        off += 2 * align(mats * SK * SK * 4)
        mean = ws[off:off + mats * SK * 4].view(torch.float32).view(n_chunks, HKV, SK).clone()
        # our sketch through the fused kernel (positions irrelevant for the sketch)
        cache = build_cos_sin_cache(D, 8, 10000.0, None).to(dev)
        pos = torch.zeros(n_tok, dtype=torch.int64, device=dev)
        _, _, sk = fused_qk_norm_rope(pos, q.contiguous(), k_pre, cache, PHI=phi)

        # the reference's torch arithmetic
        X = torch.matmul(k_pre.transpose(0, 1), phi)  # [H, N, 48] bf16
        Xo = sk.transpose(0, 1)
        r = {"sketch_entries_differ": int((X != Xo).sum()), "sketch_entries": X.numel()}
        ch = X.view(HKV, -1, CH, SK).clone()
        mu = ch.mean(dim=-2, keepdim=True)
        r["mean_entries_differ"] = int((mu.squeeze(2).transpose(0, 1).float() != mean).sum())
        r["mean_entries"] = mean.numel()
        ch.sub_(mu)
        g = torch.matmul(ch.transpose(-1, -2), ch)  # [H, NC, 48, 48] bf16
        G = g.to(torch.float32)
        G.diagonal(dim1=-2, dim2=-1).add_(5e-3)
        Go = gram.transpose(0, 1)
        dg = (G != Go)
        r["gram_entries_differ"] = int(dg.sum())
        r["gram_entries"] = G.numel()
        r["gram_matrices_differ"] = int(dg.flatten(2).any(dim=2).sum())
        r["gram_matrices"] = mats
        # the same Gram from OUR sketch but torch's mean/centre/GEMM: separates "sketch bits" from "Gram order"
        ch2 = Xo.contiguous().view(HKV, -1, CH, SK).clone()
        ch2.sub_(ch2.mean(dim=-2, keepdim=True))
        G2 = torch.matmul(ch2.transpose(-1, -2), ch2).to(torch.float32)
        G2.diagonal(dim1=-2, dim2=-1).add_(5e-3)
        r["gram_from_our_sketch_by_torch_vs_ours_entries_differ"] = int((G2 != Go).sum())
        r["gram_from_our_sketch_by_torch_vs_ref_entries_differ"] = int((G2 != G).sum())
        # fp64-exact Gram of the reference's centred chunk, rounded once: how often does cuBLAS itself miss it?
        ex = torch.matmul(ch.double().transpose(-1, -2), ch.double()).to(torch.bfloat16).float()
        ex.diagonal(dim1=-2, dim2=-1).add_(5e-3)
        r["ref_gram_vs_exactly_rounded_entries_differ"] = int((G != ex).sum())
        r["our_gram_vs_exactly_rounded_entries_differ"] = int((Go != ex).sum())
        rep[kind] = r
        print(kind, json.dumps(r), flush=True)
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "r02_leverage_stage_diff.json").write_text(json.dumps(rep, indent=1))


if __name__ == "__main__":
    main()
