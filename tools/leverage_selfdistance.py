"""How far is the reference's leverage scorer from ITSELF, and how far is this repo from it?  (task 1)

    python tools/leverage_selfdistance.py            # both sides as subprocesses -> gpurun_out/r02_leverage_selfdistance.json

`approximate_leverage_scores` (reference compression/compactor.py:113-221) inverts 48x48 Gram matrices that were
rounded to bf16 entrywise; with low-rank keys the small singular values of that rounded Gram are rounding noise, so
`x^T |G|^-1 x` is ill-conditioned in the reference itself.  This tool measures that on the B200 with the UNMODIFIED
reference (baseline/_ref), varying only things the reference does not pin:

  ref            as shipped: cuSOLVER gesvda (an approximate batched SVD), compactor.py:177
  ref_gesvdj     same code, torch.linalg.svd(driver="gesvdj")   (Jacobi)
  ref_gesvd      same code, driver="gesvd"                      (QR iteration)
  ref_eigh64     same code, the SVD replaced by an fp64 symmetric eigendecomposition of the same fp32 Gram
                 (S = |lambda|): the exact factorisation of what the reference factorises
  ref_ulp        as shipped, K perturbed by one bf16 ulp in 1 % of its entries
  ref_splitk     as shipped, torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction flipped (another
                 legal cuBLAS accumulation order for the bf16 GEMMs; no effect at these shapes)
  ref_gram_f32   as shipped, but the Gram X^T X is accumulated by an fp32 GEMM from the same bf16 operands and then
                 rounded to bf16 — the same sum in another fp32 order: what ANY other GEMM kernel (a different
                 cuBLAS version / heuristic, or this repo's) is expected to look like
  ref_sketch_f32 likewise for the sketch GEMM X = K.PHI
  ours / ours_sketch / ours_ulp   this repo (K route, rope.cu sketch route is covered by tests; perturbed K)

Statistics of every variant against `ref`, on the bench's structured keys and on white-noise keys, at 32768
tokens: fraction of entries within 2e-2*max(1,|s|), median / max scaled error, overlap of the top-50 % sets of the
z-scored leverage scores, and the END-TO-END kept-set overlap after blending 0.5*leverage into the reference's own
post-RoPE chunk scores (compactor.py:489-599) with the engine's protected tokens.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
HQ, HKV, D = 32, 8, 128


def _keys(torch, kind, n_tok, dev):
    from ref_compare import _inputs
    q, k, v, k_pre = _inputs(torch, n_tok, 0, dev)
    if kind == "white":
        g = torch.Generator(device=dev).manual_seed(123)
        k_pre = torch.randn((n_tok, HKV, D), generator=g, device=dev).to(torch.bfloat16)
    return q, k, v, k_pre


def _ulp_perturb(torch, x, frac, seed):
    """+-1 bf16 ulp on a random `frac` of the entries (bit pattern +-1: exactly one ulp away)."""
    g = torch.Generator(device=x.device).manual_seed(seed)
    bits = x.view(torch.int16).clone()
    pick = torch.rand(bits.shape, generator=g, device=x.device) < frac
    sign = (torch.randint(0, 2, bits.shape, generator=g, device=x.device, dtype=torch.int16) * 2 - 1)
    bits = torch.where(pick, bits + sign, bits)
    return bits.view(torch.bfloat16)


def run_side(side: str, n_tok: int):
    sys.path.insert(0, str(ROOT / ("baseline/_ref" if side == "ref" else "compactor-vllm_b200")))
    import torch
    import compactor_vllm
    assert ("baseline/_ref" in compactor_vllm.__file__) == (side == "ref"), compactor_vllm.__file__
    from compactor_vllm.compression.compactor import approximate_leverage_scores, non_causal_attn_scores
    from ref_compare import _first_call

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    phi = (torch.randn((D, 48), generator=torch.Generator(device=dev).manual_seed(42), device=dev)
           .to(torch.bfloat16) / math.sqrt(48))
    outs, notes = {}, {}
    real_svd = torch.linalg.svd

    def lev(kp):
        return approximate_leverage_scores(kp, [n_tok], phi, normalize=True, chunk_size=512).float().cpu()

    for kind in ("structured", "white"):
        q, k, v, k_pre = _keys(torch, kind, n_tok, dev)
        k_ulp = _ulp_perturb(torch, k_pre, 0.01, 7)
        if side == "ours":
            outs[f"{kind}/ours"] = lev(k_pre)
            outs[f"{kind}/ours_ulp"] = lev(k_ulp)
            cu = torch.tensor([0, n_tok], dtype=torch.int32, device=dev)
            pre = approximate_leverage_scores(k_pre, [n_tok], phi, normalize=True, chunk_size=512)
            outs[f"{kind}/ours_e2e"] = non_causal_attn_scores(
                q, k, v, cu, n_tok, chunk_size=128, sm_scale=1.0, normalize=True, accum_scores=pre,
                context_lens=[n_tok], protected_first_tokens=[16], protected_last_tokens=[64],
                accum_blending=0.5).float().cpu()
            continue
        # ---- reference side: same code, only the unpinned choices vary ----
        outs[f"{kind}/ref"] = lev(k_pre)
        outs[f"{kind}/ref_again"] = lev(k_pre)  # run-to-run determinism of the shipped path
        outs[f"{kind}/ref_ulp"] = lev(k_ulp)
        for drv in ("gesvdj", "gesvd"):
            def svd_drv(a, full_matrices=True, *, driver=None, _d=drv):
                return real_svd(a, full_matrices=full_matrices, driver=_d)
            torch.linalg.svd = svd_drv
            try:
                outs[f"{kind}/ref_{drv}"] = lev(k_pre)
            except Exception as e:  # noqa: BLE001
                notes[f"{kind}/ref_{drv}"] = f"{type(e).__name__}: {str(e)[:200]}"
            finally:
                torch.linalg.svd = real_svd

        def svd_eigh64(a, full_matrices=True, *, driver=None):
            lam, vec = torch.linalg.eigh(a.double())
            return vec.float(), lam.abs().float(), vec.transpose(-1, -2).float()
        torch.linalg.svd = svd_eigh64
        try:
            outs[f"{kind}/ref_eigh64"] = lev(k_pre)
        finally:
            torch.linalg.svd = real_svd
        real_mm = torch.matmul
        for tag, is_target in (("gram_f32", lambda a, b: a.shape[-2] == 48 and b.shape[-1] == 48 and a.shape[-1] > 48),
                               ("sketch_f32", lambda a, b: a.shape[-1] == D and b.shape[-1] == 48)):
            def mm(a, b, _t=is_target):
                if a.dtype == torch.bfloat16 and _t(a, b):
                    return real_mm(a.float(), b.float()).to(torch.bfloat16)
                return real_mm(a, b)
            torch.matmul = mm
            try:
                outs[f"{kind}/ref_{tag}"] = lev(k_pre)
            finally:
                torch.matmul = real_mm
        flag = torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction
        torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = not flag
        try:
            outs[f"{kind}/ref_splitk"] = lev(k_pre)
        finally:
            torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = flag
        # the reference's post-RoPE scorer, for the end-to-end blend of every variant (run by the parent's
        # request below): keep its raw per-sequence z-scored chunk scores WITHOUT a leverage term
        cu = torch.tensor([0, n_tok], dtype=torch.int32, device=dev)
        f = lambda: non_causal_attn_scores(q, k, v, cu, n_tok, chunk_size=128, sm_scale=1.0, normalize=True,
                                           context_lens=[n_tok], protected_first_tokens=[16],
                                           protected_last_tokens=[64])
        _first_call(f, ["compactor_vllm.compression.compactor"], notes)
        outs[f"{kind}/ref_chunk"] = f().float().cpu()
        pre = approximate_leverage_scores(k_pre, [n_tok], phi, normalize=True, chunk_size=512)
        outs[f"{kind}/ref_e2e"] = non_causal_attn_scores(
            q, k, v, cu, n_tok, chunk_size=128, sm_scale=1.0, normalize=True, accum_scores=pre,
            context_lens=[n_tok], protected_first_tokens=[16], protected_last_tokens=[64],
            accum_blending=0.5).float().cpu()
    outs["notes"] = notes
    torch.save(outs, f"/tmp/levself_{side}.pt")


def _stats(torch, x, y):
    fin = torch.isfinite(y) & torch.isfinite(x)
    d = (x[fin] - y[fin]).abs() / y[fin].abs().clamp_min(1.0)
    kk = x.numel() // 2
    tx = torch.topk(torch.nan_to_num(x.flatten(), nan=float("inf")), kk).indices
    ty = torch.topk(torch.nan_to_num(y.flatten(), nan=float("inf")), kk).indices
    m = torch.zeros(x.numel(), dtype=torch.bool)
    m[tx] = True
    return {"frac_within_2e-2": float((d <= 2e-2).float().mean()), "median_scaled_err": float(d.median()),
            "p99_scaled_err": float(d.kthvalue(max(1, int(0.99 * d.numel()))).values),
            "max_scaled_err": float(d.max()), "nonfinite_pattern_equal": bool(torch.equal(torch.isfinite(x), torch.isfinite(y))),
            "top50_overlap": float(m[ty].float().mean())}


def _e2e(torch, lev, chunk):
    """blend as compactor.py:586-598: chunk (z-scored, protected = +inf) + 0.5 * leverage"""
    return chunk + 0.5 * lev


def compare(n_tok):
    import torch
    a = torch.load("/tmp/levself_ref.pt")
    b = torch.load("/tmp/levself_ours.pt")
    rep = {"what": "leverage scores: distance of the unmodified reference to itself (unpinned choices varied) and of "
                   "this repo to the reference; B200, %d tokens, HKV=8, D=128, sketch 48, chunk 512, bf16" % n_tok,
           "notes": a.get("notes", {}), "inputs": {}}
    for kind in ("structured", "white"):
        ref = a[f"{kind}/ref"]
        chunk = a[f"{kind}/ref_chunk"]
        e2e_ref = _e2e(torch, ref, chunk)
        rows = {}
        for name, src in ([(k.split("/")[1], a[k]) for k in a if k.startswith(kind + "/ref_") and
                           k.split("/")[1] not in ("ref_chunk", "ref_e2e")] +
                          [(k.split("/")[1], b[k]) for k in b if k.startswith(kind + "/ours") and
                           not k.endswith("_e2e")]):
            st = _stats(torch, src, ref)
            st["e2e_kept_set_overlap"] = _stats(torch, _e2e(torch, src, chunk), e2e_ref)["top50_overlap"]
            rows[name + "_vs_ref"] = st
        # each side against the exact factorisation of the reference's own Gram
        ex = a[f"{kind}/ref_eigh64"]
        rows["ref_vs_eigh64"] = _stats(torch, ref, ex)
        rows["ours_vs_eigh64"] = _stats(torch, b[f"{kind}/ours"], ex)
        rows["ours_ulp_vs_ours"] = _stats(torch, b[f"{kind}/ours_ulp"], b[f"{kind}/ours"])
        # whole chains as shipped on both sides (blend computed by each side's own kernels)
        full = _stats(torch, b[f"{kind}/ours_e2e"], a[f"{kind}/ref_e2e"])
        rows["full_chain_ours_vs_ref"] = full
        rows["blend_formula_check_ref"] = _stats(torch, e2e_ref, a[f"{kind}/ref_e2e"])
        rep["inputs"][kind] = rows
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "r02_leverage_selfdistance.json").write_text(json.dumps(rep, indent=1))
    print(json.dumps(rep, indent=1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--side", choices=["ref", "ours"])
    ap.add_argument("--tokens", type=int, default=32768)
    a = ap.parse_args()
    if a.side:
        run_side(a.side, a.tokens)
        return
    for side in ("ref", "ours"):
        env = dict(os.environ)
        env.pop("PYTHONPATH", None)
        t0 = time.time()
        p = subprocess.run([sys.executable, __file__, "--side", side, "--tokens", str(a.tokens)], env=env,
                           capture_output=True, text=True, timeout=1200)
        log = ROOT / "gpurun_out" / f"levself_{side}.log"
        log.parent.mkdir(exist_ok=True)
        log.write_text(p.stdout[-20000:] + "\n--- stderr ---\n" + p.stderr[-20000:])
        print(f"[{side}] rc={p.returncode} {time.time() - t0:.0f}s", flush=True)
        if p.returncode != 0:
            print(p.stderr[-3000:])
            sys.exit(1)
    compare(a.tokens)


if __name__ == "__main__":
    main()
