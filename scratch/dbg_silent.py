import sys, torch, collections
sys.path.insert(0, '.'); sys.path.insert(0, 'scratch')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
import importlib
C = 42
POS = {}
def aligned(T, B, L, g, boost, il, tl, tg, noise=0.1):
    x = torch.randn(T,B,C,generator=g,device='cuda')
    tgc, ilc, tlc = tg.cpu(), il.cpu(), tl.cpu()
    gc = torch.Generator().manual_seed(7)
    lab = torch.zeros(T,B,dtype=torch.long)
    for b in range(B):
        Tb, Lb = int(ilc[b]), int(tlc[b])
        pos = torch.sort(torch.randperm(max(Tb//2,1), generator=gc)[:Lb])[0]*2
        POS[b] = pos.tolist()
        lab[pos, b] = tgc[b,:Lb]
    lab = lab.cuda()
    flip = torch.rand(T,B,generator=g,device='cuda') < noise
    lab = torch.where(flip, torch.randint(0,C,(T,B),generator=g,device='cuda'), lab)
    x.scatter_add_(2, lab[...,None], torch.full((T,B,1), float(boost), device='cuda'))
    return x
g = torch.Generator(device='cuda').manual_seed(3)
T,B,L = 1000, 512, 120
tg = torch.randint(1,C,(B,L),generator=g,device='cuda')
il = torch.full((B,),T,device='cuda'); tl = torch.full((B,),L,device='cuda')
lp = torch.log_softmax(aligned(T,B,L,g,25,il,tl,tg),-1)
n1,_,_ = ctc_loss_raw(lp,tg,il,tl,algo=1,validate=False); n1 = n1.clone()
n2,_,_ = ctc_loss_raw(lp,tg,il,tl,algo=2,validate=False)
fl = last_fallback_flags(B)
bad = ((fl == 0) & (((n1-n2)/n1).abs() > 1e-5)).nonzero().flatten().tolist()
print('flag counts', dict(collections.Counter((int(f)&255) for f in fl.cpu() if int(f))), 'silent-wrong', bad[:10])
torch.cuda.synchronize()
print('silent', len(bad))
for b in bad[:3]:
    pos = POS[b]
    ks = list(range(2, L+1))
    Bp = len(ks)
    lpp = lp[:, b:b+1].expand(T, Bp, C).contiguous()
    tgp = tg[b:b+1].expand(Bp, L).contiguous()
    ilp = torch.tensor([min(T, pos[k-1]+2) for k in ks], device='cuda'); tlp = torch.tensor(ks, device='cuda')
    m1,_,_ = ctc_loss_raw(lpp,tgp,ilp,tlp,algo=1,validate=False); m1 = m1.clone()
    m2,_,_ = ctc_loss_raw(lpp,tgp,ilp,tlp,algo=2,validate=False)
    m3,_,_ = ctc_loss_raw(lpp,tgp,ilp,tlp,algo=3,validate=False)
    f = last_fallback_flags(Bp).cpu()
    d = (m2 - m1).cpu(); d3 = (m3 - m1).cpu()
    print('utt', b, 'full nll log %.3f pair %.3f' % (float(n1[b]), float(n2[b])))
    first = None
    for i, k in enumerate(ks):
        if abs(float(d[i])) > 1e-2 and first is None: first = i
    for i in range(max(0,(first or 0)-3), min(Bp,(first or 0)+4)):
        print('   k=%d T=%d nll log %.3f  pair-log %.4f (flag %d at %d)  fast-log %.4f' % (ks[i], int(ilp[i]), float(m1[i]), float(d[i]), int(f[i])&255, int(f[i])>>8, float(d3[i])))
    if first is not None:
        import numpy as np
        np.savez_compressed('gpurun_out/silent_%d.npz' % b, lp=lp[:int(ilp[first]), b].cpu().numpy(), tg=tg[b,:ks[first]].cpu().numpy(),
                            nll_log=float(m1[first]), nll_pair=float(m2[first]))
        k = ks[first]
        t_hi = int(ilp[first]); t_lo = pos[k-3] if k >= 3 else 0
        print('   labels around k:', tg[b, max(0,k-4):k+1].tolist(), 'positions', pos[max(0,k-4):k+1])
        for t in range(max(0,t_lo-2), t_hi):
            row = lp[t, b]; top = int(row.argmax())
            pass; #print('     t=%d boosted class %d | blank %.1f | lp of labels k-3..k: %s' % (t, top, float(row[0]), [round(float(row[int(c)]),1) for c in tg[b, max(0,k-4):k+1]]))
