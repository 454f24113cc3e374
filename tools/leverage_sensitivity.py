"""How far the reference's own leverage scores move when only its 48x48 factorisation changes
(fp32 svd -> fp64 svd / eigh) on the golden fixture: the noise floor the GPU parity test uses."""
import sys; sys.path[:0]=['/root/repo','/root/repo/tests']
import torch, numpy as np
from conftest import load_golden
import oracle.scoring as osc
g=load_golden('leverage_bf16')
lens=g['context_lens'].tolist()
keys,phi=g['keys'],g['phi']
dt=keys.dtype
X=torch.matmul(keys.transpose(0,1).to(dt),phi.to(dt))
_,cl=osc.split_into_chunks(lens,128)
offs=np.cumsum([0]+cl)
for i,L in enumerate(cl):
    s,e=offs[i],offs[i+1]
    ch=X[:,s:e]; ch=ch-ch.mean(dim=-2,keepdim=True)
    G=torch.matmul(ch.transpose(-1,-2),ch).float()+5e-3*torch.eye(48)
    outs=[]
    for mode in ('svd32','svd64','eigh32'):
        if mode=='svd32': V,S,_=torch.linalg.svd(G)
        elif mode=='svd64':
            V,S,_=torch.linalg.svd(G.double()); V,S=V.float(),S.float()
        else:
            S,V=torch.linalg.eigh(G); S=S.abs()
        SV=(V*S.rsqrt().unsqueeze(-2)).to(dt)
        U=torch.matmul(ch,SV)
        outs.append((U*U).sum(-1).clamp_min(0).float())
    ref=g['raw'][s:e].float().T
    print(L, 'smin',S.min().item(), 'svd32-golden',(outs[0]-ref).abs().max().item(),'svd64-svd32',(outs[1]-outs[0]).abs().max().item(),'eigh-svd32',(outs[2]-outs[0]).abs().max().item(), 'max',ref.max().item())
