import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
from mddctc import _rt
for (T,B,C,L) in ((1000,64,42,30),(400,4,42,30),(300,4,42,30),(1000,300,42,120),(1500,16,45,180),(37,5,43,9)):
    g = torch.Generator().manual_seed(0)
    lp = torch.log_softmax(torch.randn(T,B,C,generator=g),-1).cuda()
    tg = torch.randint(1,C,(B,L),generator=g).cuda()
    il = torch.full((B,),T).cuda(); tl = torch.full((B,),min(L, T//2)).cuda()
    n1,g1,_ = ctc_loss_raw(lp,tg,il,tl,algo=1); n1, g1 = n1.clone(), g1.clone()
    n2,g2,_ = ctc_loss_raw(lp,tg,il,tl,algo=2, validate=False)
    torch.cuda.synchronize()
    ws = _rt._ws_cache[(0,'ctc')]
    off = (B*8 + 255)//256*256
    flags = ws[off:off+B*4].view(torch.int32)
    bad = [(int(f)&255, int(f)>>8) for f in flags if int(f)]
    print(T,B,L,'flagged',len(bad),bad[:4],'nll rel', float(((n1-n2)/n1).abs().max()), 'grad diff', float((g1-g2).abs().max()), 'nan', int(torch.isnan(g2).sum()))
