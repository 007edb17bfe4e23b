#!/usr/bin/env python
"""How far the float32-distance linear-RBF path is from scipy's fp64 result over a WHOLE slab (16 x 1024 pixels, the
bench geometry), not a sample: node residuals of the pcg weights against the exact matrix, the |w|_2 estimator the
C entry uses to report pixels for the fp64 kernels, and the tables of the worst remaining pixels against the oracle.

    python tools/rbf_tail_study.py [noise|lambert] [n]
"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
from rti_b200 import _capi, synthetic
import rti_oracle as orc
lib=_capi.load()
rows, cols, n = 16, 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 500
rng = np.random.default_rng(8)
cams = synthetic.camera_positions(1024, 1024, n, rng)
kind = sys.argv[1] if len(sys.argv)>1 else "noise"
if kind=="noise":
    frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)
else:
    frames,_ = synthetic.lambertian_stack(rows, cols, n, seed=3, rows_total=1024, row0=512)
d_f, d_c = torch.from_numpy(frames).cuda(), torch.from_numpy(cams).cuda()
px = rows*cols
lx = torch.empty((px, n), dtype=torch.float32, device="cuda"); ly = torch.empty_like(lx)
inten = torch.empty((px, n), dtype=torch.int32, device="cuda")
_capi.call("rti_gather_soa", d_f.data_ptr(), px, d_c.data_ptr(), n, 512, rows, cols, lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), None)
tiles=(rows//16)*(cols//16); nbytes=lib.rti_rbf_pcg_workspace_bytes(n,tiles)
work=torch.empty(nbytes,dtype=torch.uint8,device="cuda"); w=torch.empty((px,n),dtype=torch.float64,device="cuda")
bad=torch.zeros(2,dtype=torch.int32,device="cuda")
_capi.call("rti_rbf_fit_pcg", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), rows, cols, n, w.data_ptr(), work.data_ptr(), nbytes, 48, 0.01, None, bad.data_ptr(), None)
torch.cuda.synchronize()
res=torch.empty(px,device="cuda",dtype=torch.float64); minsep=torch.empty(px,device="cuda",dtype=torch.float64)
for p0 in range(0,px,512):
    x=lx[p0:p0+512].double(); y=ly[p0:p0+512].double()
    a=torch.sqrt((x[:,:,None]-x[:,None,:])**2+(y[:,:,None]-y[:,None,:])**2)
    res[p0:p0+512]=(torch.bmm(a,w[p0:p0+512,:,None])[:,:,0]-inten[p0:p0+512].double()).abs().amax(dim=1)
    minsep[p0:p0+512]=(a+torch.eye(n,device="cuda")[None]*9).amin(dim=(1,2))
w2=w.norm(dim=1); w1=w.abs().sum(dim=1)
r=res.cpu().numpy(); 
print(kind,"node residual: median %.3f p90 %.3f p99 %.3f max %.3f"%(np.median(r),np.percentile(r,90),np.percentile(r,99),r.max()))
print("||w||2: median %.2e p99 %.2e max %.2e ; sum|w| median %.2e max %.2e ; min sep min %.1e"%(w2.median().item(), np.percentile(w2.cpu().numpy(),99), w2.max().item(), w1.median().item(), w1.max().item(), minsep.min().item()))
est=(1.2e-7*w2).cpu().numpy()
for thr in (0.03,0.05,0.1):
    print("thr",thr,"flagged by est",int((est>thr).sum()),"of",px,"; residual>thr",int((r>thr).sum()),"; missed (res>2thr & not flagged)",int(((r>2*thr)&(est<=thr)).sum()))
# grid error of the worst pixels through the fast table vs oracle
ax=orc.light_axis(); d_ax=torch.from_numpy(ax).cuda()
iters=torch.empty(px,dtype=torch.int32,device="cuda")
_capi.call("rti_rbf_fit_pcg", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), rows, cols, n, w.data_ptr(), work.data_ptr(), nbytes, 48, 0.01, iters.data_ptr(), bad.data_ptr(), None)
torch.cuda.synchronize()
fl=(iters<0).cpu().numpy()
print("reported by the C entry for the fp64 kernels:",int(fl.sum()),"of",px,"; worst residual among the rest %.3f"%r[~fl].max())
rr=r.copy(); rr[fl]=0
worst=np.argsort(-rr)[:6]
tab=torch.empty((100,100,px),dtype=torch.int32,device="cuda"); tb=torch.zeros(1,dtype=torch.int32,device="cuda")
_capi.call("rti_rbf_table_fast", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px, n, d_ax.data_ptr(), 100, tab.data_ptr(), px, tb.data_ptr(), None)
torch.cuda.synchronize()
for q in worst:
    q=int(q); l1,l2,dd=lx[q].cpu().numpy(),ly[q].cpu().numpy(),inten[q].cpu().numpy()
    ref=orc.rbf_grid(l1,l2,orc.rbf_fit(l1,l2,dd),ax)
    got=orc.rbf_grid(l1,l2,w[q].cpu().numpy(),ax)
    t=tab[:,:,q].cpu().numpy()
    dif=t!=np.trunc(ref)
    beyond=int(np.count_nonzero(dif&(np.abs(ref-np.rint(ref))>1e-3*np.abs(ref).max())))
    print("pixel",q,"node res %.3f"%r[q],"||w||2 %.2e"%w2[q].item(),"grid err fit only /max %.2e"%(np.abs(got-ref).max()/np.abs(ref).max()),"table vs trunc(ref): max|diff| %d, frac differing %.4f, entries off although the fp64 value is > 1e-3 max from an integer: %d"%(np.abs(t-np.trunc(ref)).max(), dif.mean(), beyond))
