"""Dump the device Gram / SV of one chunk and compare each stage with a CPU restatement."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "compactor-vllm_b200"), str(ROOT / "tests"), str(ROOT / "tools")]
import numpy as np
import torch
import oracle.scoring as osc
from compactor_vllm import _native as N
from compactor_vllm.compression.compactor import approximate_leverage_scores
from compactor_vllm.utils.arguments import make_phi
from test_gpu_scores import _structured_keys
from jacobi_emulation import jacobi

dtype, lens, H, chunk = torch.bfloat16, [1300, 700], 8, 512
Ntok = sum(lens)
keys = _structured_keys(Ntok, H, 128, torch.float32, seed=Ntok + H).to(dtype)
phi = make_phi(128, 48, dtype, "cpu")
raw = approximate_leverage_scores(keys.cuda(), lens, phi.cuda(), normalize=False, chunk_size=chunk).cpu().float()
torch.cuda.synchronize()
ws = [v for k, v in N._WORKSPACES.items() if k[2] == "leverage"][0].cpu()
_, cl = osc.split_into_chunks(lens, chunk)
n_chunks = len(cl)
al = lambda x: (x + 255) & ~255
o_gram = al(Ntok * H * 4)
mats = n_chunks * H
o_sv = o_gram + al(mats * 48 * 48 * 4)
o_mean = o_sv + al(mats * 48 * 48 * 4)
gram = ws[o_gram:o_gram + mats * 48 * 48 * 4].view(torch.float32).view(n_chunks, H, 48, 48)
sv = ws[o_sv:o_sv + mats * 48 * 48 * 4].view(torch.float32).view(n_chunks, H, 48, 48)
mean = ws[o_mean:o_mean + mats * 48 * 4].view(torch.float32).view(n_chunks, H, 48)
X = torch.matmul(keys.transpose(0, 1), phi)  # [H, N, 48] bf16
offs = np.cumsum([0] + cl)
for ci in (2, 0, 4):
    s, e = offs[ci], offs[ci + 1]
    c = X[:, s:e]
    mu = c.mean(dim=-2, keepdim=True)
    cc = c - mu
    g = torch.matmul(cc.transpose(-1, -2), cc).float() + 5e-3 * torch.eye(48)
    print(f"chunk {ci} len {cl[ci]}: mean diff {(mean[ci] - mu[:, 0].float()).abs().max():.3e}  gram diff "
          f"{(gram[ci] - g).abs().max():.3e} (gram max {g.abs().max():.3e})")
    for h in (0, 1):
        lam64, V64 = np.linalg.eigh(g[h].double().numpy())
        lam, V, _ = jacobi(gram[ci, h].numpy(), 8)
        order = np.argsort(lam)
        print(f"   h{h}: eig64 smallest |lam| {np.sort(np.abs(lam64))[:4]}  jacobi(cpu) {np.sort(np.abs(lam))[:4]}")
        W = torch.from_numpy(V * (1.0 / np.sqrt(np.abs(lam)))[None, :]).to(dtype).float()
        dsv = (sv[ci, h].abs().sort(dim=0).values - W.abs().sort(dim=0).values)
        # column norms identify columns irrespective of order / sign
        n_gpu = np.sort(sv[ci, h].norm(dim=0).numpy())[::-1][:4]
        n_cpu = np.sort(W.norm(dim=0).numpy())[::-1][:4]
        n_64 = np.sort(1.0 / np.sqrt(np.abs(lam64)))[::-1][:4]
        print(f"        largest SV column norms: gpu {n_gpu}  cpu-jacobi {n_cpu}  eigh64 {n_64}")
