import sys, numpy as np, torch
sys.path[:0]=['/root/repo','/root/repo/tests','/root/repo/compactor-vllm_b200']
import oracle.scoring as osc
from compactor_vllm.utils.arguments import make_phi
sys.path.append('/root/repo')
import bench
def jacobi(G, sweeps):
    n=G.shape[0]; G=G.astype(np.float32).copy(); V=np.eye(n,dtype=np.float32); hist=[]
    for sw in range(sweeps):
        for rnd in range(n-1):
            pairs=[]
            for tid in range(n//2):
                a = 0 if tid==0 else 1+(tid-1+rnd)%(n-1)
                b = 1+(n-2-tid+rnd)%(n-1)
                if tid==0: b=1+(n-2+rnd)%(n-1)
                pairs.append((min(a,b),max(a,b)))
            J=np.eye(n,dtype=np.float32)
            for p,q in pairs:
                app,aqq,apq=G[p,p],G[q,q],G[p,q]
                c,s=np.float32(1),np.float32(0)
                if abs(apq)>1e-30:
                    tau=(aqq-app)/(2*apq); t=np.sign(tau) if tau!=0 else 1.0
                    t=np.float32((1.0 if tau>=0 else -1.0)/(abs(tau)+np.sqrt(1+tau*tau)))
                    c=np.float32(1/np.sqrt(1+t*t)); s=np.float32(t*c)
                J[p,p]=c;J[q,q]=c;J[p,q]=s;J[q,p]=-s
            G=(J.T@G@J).astype(np.float32); V=(V@J).astype(np.float32)
        off=np.sqrt((G**2).sum()-(np.diag(G)**2).sum()); hist.append(off/np.sqrt((np.diag(G)**2).sum()))
    return np.diag(G).copy(),V,hist
if __name__=="__main__":
  qkv,k_pre=bench.make_inputs(2048,'cpu',torch.bfloat16,7)
  phi=make_phi(128,48,torch.bfloat16,'cpu')
  X=torch.matmul(k_pre.transpose(0,1),phi)
  for h in (0,3):
    for s in (0,512):
      c=X[h,s:s+512]; c=c-c.mean(0,keepdim=True)
      g=(c.T@c).float().numpy()+5e-3*np.eye(48,dtype=np.float32)
      lam_ref=np.linalg.eigvalsh(g.astype(np.float64))
      lam,V,hist=jacobi(g,10)
      print('h',h,'chunk',s,'lam range',lam_ref.min(),lam_ref.max(),'hist',['%.1e'%x for x in hist])
      # score error vs sweeps
      xs=c.float().numpy()
      def score(lam,V): W=V/np.sqrt(np.abs(lam))[None,:]; return ((xs@W)**2).sum(1)
      w,Vr=np.linalg.eigh(g.astype(np.float64)); sref=score(w,Vr)
      for sw in (4,5,6,7,8):
          l2,V2,_=jacobi(g,sw); sc=score(l2,V2); print('   sweeps',sw,'max rel score err',np.abs(sc-sref).max()/np.abs(sref).max(), 'max abs', np.abs(sc-sref).max())
