import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
from mddctc import _rt
torch.manual_seed(0)
def case(T,B,C,L, il=None, tl=None, tg=None, tag=""):
    g = torch.Generator().manual_seed(0)
    lp = torch.log_softmax(torch.randn(T,B,C,generator=g),-1).cuda()
    tg = (torch.randint(1,C,(B,L),generator=g) if tg is None else tg).cuda()
    il = (torch.full((B,),T) if il is None else torch.tensor(il)).cuda(); tl = (torch.full((B,),min(L, T//2)) if tl is None else torch.tensor(tl)).cuda()
    n1,g1,_ = ctc_loss_raw(lp,tg,il,tl,algo=1); n1, g1 = n1.clone(), g1.clone()
    n2,g2,_ = ctc_loss_raw(lp,tg,il,tl,algo=2, validate=False)
    torch.cuda.synchronize()
    ws = _rt._ws_cache[(0,'ctc')]
    off = (B*8 + 255)//256*256
    flags = ws[off:off+B*4].view(torch.int32).cpu().tolist()
    d = (g1-g2).abs()
    per = d.amax(dim=(0,2)).cpu()
    print(tag, (T,B,C,L), 'NFLAGGED', sum(1 for f in flags if f), [(f&255, f>>8) for f in flags if f][:6], 'nll rel %.2e' % float(((n1-n2)/n1).abs().max()), 'grad diff %.3e' % float(d.max()), 'nan', int(torch.isnan(g2).sum()))
    for b in range(B):
        if per[b] > 1e-4 or per[b] != per[b]:
            idx = (d[:,b,:] > 1e-4).nonzero().cpu()
            ts = sorted(set(idx[:,0].tolist())); cs = sorted(set(idx[:,1].tolist()))
            print('   utt', b, 'T', int(il[b]), 'L', int(tl[b]), 'err %.3e' % float(per[b]), 'n_bad', len(idx), 't', ts[:12], 'c', cs[:12])
            t = ts[0]
            print('     row t=%d ref' % t, [round(float(x),4) for x in g1[t,b,:8]], 'got', [round(float(x),4) for x in g2[t,b,:8]])
case(37,5,43,9, tag="oddB")
case(37,4,42,9, tag="evenB")
case(37,4,42,9, il=[37,37,30,30], tag="raggedpairs-eqT")
case(37,4,42,9, il=[37,30,37,21], tag="ragged")
case(25,4,7,6, il=[25,17,9,12], tl=[6,6,0,4], tg=torch.tensor([[3, 3, 3, 2, 2, 1], [1, 2, 3, 4, 5, 6], [0, 0, 0, 0, 0, 0], [5, 5, 1, 1, 0, 0]]), tag="repeat_ragged")
case(25,4,7,6, tl=[6,6,0,4], tg=torch.tensor([[3, 3, 3, 2, 2, 1], [1, 2, 3, 4, 5, 6], [0, 0, 0, 0, 0, 0], [5, 5, 1, 1, 0, 0]]), tag="repeat")
case(1000,64,42,30, tag="")
case(1000,300,42,120, tag="")
case(1500,16,45,180, tag="")
case(200,32,42,30, tag="c2")
