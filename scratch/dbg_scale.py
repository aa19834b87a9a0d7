import sys, torch, collections
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
C = 42
def run(name, x, tg, il, tl):
    T, B, _ = x.shape
    lp = torch.log_softmax(x,-1)
    n1,g1,_ = ctc_loss_raw(lp,tg,il,tl,algo=1,validate=False); n1, g1 = n1.clone(), g1.clone()
    n2,g2,_ = ctc_loss_raw(lp,tg,il,tl,algo=2,validate=False)
    torch.cuda.synchronize()
    fl = last_fallback_flags(B)
    cnt = collections.Counter((int(f)&255) for f in fl.cpu() if int(f))
    ok = (fl == 0) & torch.isfinite(n1)
    gd = (g1-g2).abs().amax(dim=(0,2))
    nrel = ((n1-n2)/n1).abs()
    if ok.any() and (float(gd[ok].max()) > 9e-5 or float(nrel[ok].max()) > 1e-5):
        score = torch.where(ok, gd + nrel, torch.zeros_like(gd))
        worst = torch.topk(score, 3)[1]
        for b in worst.tolist():
            xx = lp[:int(il[b]), b:b+1].double().cpu().requires_grad_(True)
            ref = torch.nn.functional.ctc_loss(xx, tg[b:b+1,:int(tl[b])].cpu(), il[b:b+1].cpu(), tl[b:b+1].cpu(), reduction='sum')
            ref.backward()
            Tb = int(il[b])
            print('   b=%d flag(b)=%d flag(partner)=%d nll fp64 %.6f log %.6f pair %.6f | grad err: log %.2e pair %.2e' % (
                b, int(fl[b]), int(fl[b^1]), float(ref), float(n1[b]), float(n2[b]),
                float((g1[:Tb,b].double().cpu()-xx.grad[:,0]).abs().max()), float((g2[:Tb,b].double().cpu()-xx.grad[:,0]).abs().max())), flush=True)
    e0,e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ctc_loss_raw(lp,tg,il,tl,algo=0,validate=False)
    e1.record(); torch.cuda.synchronize()
    n0,g0,_ = ctc_loss_raw(lp,tg,il,tl,algo=0,validate=False)
    torch.cuda.synchronize()
    fl0 = last_fallback_flags(B)
    fin = torch.isfinite(n1)
    print('      algo0: left for the log kernel %d | all utterances vs log kernel: grad %.2e nll rel %.2e' % (
        int((fl0 != 0).sum()), float((g0-g1).abs().amax(dim=(0,2))[fin].max()), float(((n0-n1)/n1).abs()[fin].max())))
    print('%-22s B=%d flagged %s | unflagged: max grad diff vs log kernel %.2e, nll rel %.2e | ms algo0 %.3f nll %.0f inf %d' % (
        name, B, dict(cnt), float(gd[ok].max()) if ok.any() else -1, float(nrel[ok].max()) if ok.any() else -1,
        e0.elapsed_time(e1)/5, float(n1[torch.isfinite(n1)].mean()), int((~torch.isfinite(n1)).sum())), flush=True)

def aligned(T, B, L, g, boost, il, tl, tg, noise=0.1):
    """posteriors of a 'trained model': a random valid alignment of each target gets +boost"""
    x = torch.randn(T,B,C,generator=g,device='cuda')
    fl = torch.zeros(T,B,dtype=torch.long,device='cuda')
    tgc, ilc, tlc = tg.cpu(), il.cpu(), tl.cpu()
    gc = torch.Generator().manual_seed(7)
    lab = torch.zeros(T,B,dtype=torch.long)
    for b in range(B):
        Tb, Lb = int(ilc[b]), int(tlc[b])
        if Lb == 0: continue
        # choose Lb distinct frames (sorted) for label onsets with gaps >= 2 where repeats need a blank
        pos = torch.sort(torch.randperm(max(Tb//2,1), generator=gc)[:Lb])[0]*2 if Tb >= 2*Lb else torch.arange(Lb)
        pos = pos.clamp(max=Tb-1)
        lab[pos, b] = tgc[b,:Lb]
    lab = lab.cuda()
    flip = torch.rand(T,B,generator=g,device='cuda') < noise
    lab = torch.where(flip, torch.randint(0,C,(T,B),generator=g,device='cuda'), lab)
    x.scatter_add_(2, lab[...,None], torch.full((T,B,1), float(boost), device='cuda'))
    return x

g = torch.Generator(device='cuda').manual_seed(0)
T,B,L = 1000, 1024, 120
tg = torch.randint(1,C,(B,L),generator=g,device='cuda')
il = torch.full((B,),T,device='cuda'); tl = torch.full((B,),L,device='cuda')
for k in (1.0, 2.0, 3.0, 4.0, 5.0):
    run('randn*%.1f' % k, torch.randn(T,B,C,generator=g,device='cuda')*k, tg, il, tl)
x = torch.randn(T,B,C,generator=g,device='cuda')
u = torch.rand(T,B,generator=g,device='cuda'); k = torch.randint(1,C,(T,B),generator=g,device='cuda')
x[:,:,0] += 8.0*(u<0.7); x.scatter_add_(2,k[...,None],(6.0*(u>=0.7))[...,None].float())
run('peaky-random', x, tg, il, tl)
for boost in (6, 10, 16, 20, 25):
    run('aligned+%d' % boost, aligned(T,B,L,g,boost,il,tl,tg), tg, il, tl)
run('aligned+10 clean', aligned(T,B,L,g,10,il,tl,tg,noise=0.0), tg, il, tl)
run('aligned+10 noise.3', aligned(T,B,L,g,10,il,tl,tg,noise=0.3), tg, il, tl)
run('aligned+16 noise.3', aligned(T,B,L,g,16,il,tl,tg,noise=0.3), tg, il, tl)
# ragged
il2 = torch.randint(300,T+1,(B,),generator=g,device='cuda'); tl2 = torch.randint(1,L+1,(B,),generator=g,device='cuda')
run('aligned+10 ragged', aligned(T,B,L,g,10,il2,tl2,tg), tg, il2, tl2)
# tight: T_b = 2*L_b + small slack
tl3 = torch.randint(20,L+1,(B,),generator=g,device='cuda'); il3 = (2*tl3 + torch.randint(0,6,(B,),generator=g,device='cuda')).clamp(max=T)
run('aligned+10 tight', aligned(T,B,L,g,10,il3,tl3,tg), tg, il3, tl3)
run('randn tight', torch.randn(T,B,C,generator=g,device='cuda'), tg, il3, tl3)
# short: config-2 like
T2,L2 = 200, 30
tg2 = torch.randint(1,C,(B,L2),generator=g,device='cuda'); il4 = torch.full((B,),T2,device='cuda'); tl4 = torch.full((B,),L2,device='cuda')
run('c2 aligned+10', aligned(T2,B,L2,g,10,il4,tl4,tg2), tg2, il4, tl4)
run('c2 aligned+16', aligned(T2,B,L2,g,16,il4,tl4,tg2), tg2, il4, tl4)
T3,L3 = 400, 50
tg3 = torch.randint(1,C,(B,L3),generator=g,device='cuda'); il5 = torch.full((B,),T3,device='cuda'); tl5 = torch.full((B,),L3,device='cuda')
run('L50 aligned+10', aligned(T3,B,L3,g,10,il5,tl5,tg3), tg3, il5, tl5)
run('L50 aligned+16', aligned(T3,B,L3,g,16,il5,tl5,tg3), tg3, il5, tl5)
run('L50 randn*2', torch.randn(T3,B,C,generator=g,device='cuda')*2, tg3, il5, tl5)
T4,L4,B4 = 1500, 200, 256
tg4 = torch.randint(1,C,(B4,L4),generator=g,device='cuda'); il6 = torch.full((B4,),T4,device='cuda'); tl6 = torch.full((B4,),L4,device='cuda')
run('L200 aligned+10', aligned(T4,B4,L4,g,10,il6,tl6,tg4), tg4, il6, tl6)
run('L200 aligned+16', aligned(T4,B4,L4,g,16,il6,tl6,tg4), tg4, il6, tl6)
run('L200 randn*2', torch.randn(T4,B4,C,generator=g,device='cuda')*2, tg4, il6, tl6)
run('c2 randn*3', torch.randn(T2,B,C,generator=g,device='cuda')*3, tg2, il4, tl4)
