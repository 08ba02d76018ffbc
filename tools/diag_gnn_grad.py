import sys, os
sys.path[:0]=['.','cloth-splatting_b200','tests']
import numpy as np, torch
import test_gnn_timed_shape_gpu as T
from oracle import gnn_oracle
from meshnet import EncodeProcessDecode
torch.manual_seed(0)
m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).cuda()
x, ei, ef = T._batch(40, 8, 4)
rng = np.random.default_rng(7)
w = torch.from_numpy(rng.standard_normal((x.shape[0], 3)).astype(np.float32)).cuda()
eit = torch.from_numpy(ei).cuda()
def kern():
    m.zero_grad()
    xt = torch.from_numpy(x).cuda().requires_grad_(True); eft = torch.from_numpy(ef).cuda().requires_grad_(True)
    y = m(xt, eit, eft); (y*w).sum().backward()
    g = {k: p.grad.detach().clone() for k, p in m.named_parameters()}; g['x']=xt.grad.clone(); g['edge_features']=eft.grad.clone()
    return g, y.detach()
def orc(emu, perm=None):
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xr = torch.from_numpy(x).cuda().requires_grad_(True); er = torch.from_numpy(ef).cuda().requires_grad_(True)
    e_i, e_f = eit, er
    if perm is not None:
        e_i = eit[:, perm]; e_f = er[perm]
    yr = gnn_oracle.forward(sd, xr, e_i, e_f, emulate_bf16=emu, device="cuda", differentiable=True)
    (yr*w).sum().backward()
    r = {k: v.grad for k, v in sd.items()}; r['x']=xr.grad; r['edge_features']=er.grad
    return r, yr.detach()
r32,y32 = orc(False); r16,y16 = orc(True)
perm = torch.randperm(ei.shape[1], generator=torch.Generator().manual_seed(1)).cuda()
r16p,_ = orc(True, perm)
rel = T._rel
for rep in range(3):
    g,y = kern()
    rows=[]
    for k,r in r32.items():
        if r is None or float(r.norm())==0: continue
        rows.append((rel(g[k],r), rel(r16[k],r), rel(r16p[k],r), rel(g[k], r16[k]), k))
    rows.sort(reverse=True)
    print('rep',rep,'fwd e32 %.4f e16 %.4f'%(float((y-y32).abs().max()/y32.abs().max()), float((y-y16).abs().max()/y32.abs().max())))
    for e,f,fp,e16,k in rows[:6]: print('   %.3f floor %.3f floorperm %.3f vs16 %.3f %s'%(e,f,fp,e16,k))
    worst=max(r[0]/(1.5*r[1]+3e-2) for r in rows); print('   worst ratio to tol', worst, 'median e', np.median([r[0] for r in rows]), 'median floor', np.median([r[1] for r in rows]))
    keys=[k for k in r32 if k not in('x','edge_features') and r32[k] is not None]
    a=torch.cat([g[k].flatten() for k in keys]).double(); b=torch.cat([r32[k].flatten() for k in keys]).double(); c=torch.cat([r16[k].flatten() for k in keys]).double()
    print('   cos(kernel,fp32) %.4f cos(emul,fp32) %.4f rel total %.3f / %.3f'%(float(a@b/(a.norm()*b.norm())), float(c@b/(c.norm()*b.norm())), float((a-b).norm()/b.norm()), float((c-b).norm()/b.norm())))
