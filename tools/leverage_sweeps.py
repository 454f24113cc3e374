import sys, numpy as np, torch
sys.path[:0]=['/root/repo','/root/repo/tests','/root/repo/compactor-vllm_b200','/root/repo/tools']
import oracle.scoring as osc
from compactor_vllm.utils.arguments import make_phi
from test_gpu_scores import _structured_keys, _overlap
from jacobi_emulation import jacobi
def mine(keys, lens, phi, chunk, sweeps=10):
    dt=keys.dtype
    _, cl = osc.split_into_chunks(lens, chunk)
    X=torch.matmul(keys.transpose(0,1).to(dt), phi.to(dt)); H,N,k=X.shape
    out=torch.empty((N,H),dtype=dt); off=0
    for L in cl:
        c=X[:,off:off+L]; c=c-c.mean(dim=-2,keepdim=True)
        g=torch.matmul(c.transpose(-1,-2),c).float()+5e-3*torch.eye(k)
        for h in range(H):
            lam,V,_=jacobi(g[h].numpy(),sweeps)
            SV=(torch.from_numpy(V)*torch.from_numpy(np.abs(lam)).rsqrt().unsqueeze(-2)).to(dt)
            U=torch.matmul(c[h],SV); out[off:off+L,h]=(U*U).sum(-1).clamp_min(0)
        off+=L
    return out
if __name__=="__main__":
  for dtype,scale in ((torch.bfloat16,1.0),(torch.float16,0.1)):
    for lens,H,chunk in [([512],8,512),([1300,700],8,512),([100,49,600],2,512),([2048],4,512),([300],8,128)]:
        N=sum(lens); keys=(_structured_keys(N,H,128,torch.float32,seed=N+H)*scale).to(dtype); phi=make_phi(128,48,dtype,'cpu')
        ex=osc.approximate_leverage_scores(keys,lens,phi,chunk_size=chunk,factorization='eigh64').float()
        rf=osc.approximate_leverage_scores(keys,lens,phi,chunk_size=chunk,factorization='svd32').float()
        for sw in (8,10):
            m=mine(keys,lens,phi,chunk,sw).float()
            em=((m-ex).abs()/ex.abs().clamp_min(1)).flatten(); er=((rf-ex).abs()/ex.abs().clamp_min(1)).flatten()
            qs=[0.5,0.99,0.999]
            print(dtype,lens,'sweeps',sw,'mine q',[round(em.kthvalue(max(1,int(q*em.numel()))).values.item(),4) for q in qs],'max',round(em.max().item(),3),'| ref q',[round(er.kthvalue(max(1,int(q*er.numel()))).values.item(),4) for q in qs],'max',round(er.max().item(),3))
