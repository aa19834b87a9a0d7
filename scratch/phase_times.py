import sys, ctypes, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
L = mddctc._lib.lib()
import os
T,B,C,Lb = 1000,int(os.environ.get("PT_B","4096")),42,120
g = torch.Generator().manual_seed(1)
lp = torch.log_softmax(torch.randn(T,B,C,generator=g),-1).cuda()
tg = torch.randint(1,C,(B,Lb),generator=g).cuda()
il = torch.full((B,),T).cuda(); tl = torch.full((B,),Lb).cuda()
buf = (ctypes.c_ulonglong*16)()
for _ in range(2): ctc_loss_raw(lp,tg,il,tl,algo=2,validate=False)
L.mdd_debug_phase_cycles(buf, 1)
ctc_loss_raw(lp,tg,il,tl,algo=2,validate=False)
L.mdd_debug_phase_cycles(buf, 1)
names = ["setup","fwd loads+renorm","fwd ckpt+alpha","bwd bookkeeping","bwd alpha","bwd beta+scatter","bwd scan","bwd grad","fwd commit","bwd commit+loads","bwd wait ckpt","fwd wait rows"]
tot = sum(buf[i] for i in range(12))
npairs = B//2
for i,n in enumerate(names):
    print("%-20s %6.1f%%  %8.0f cycles/pair  %7.1f cycles/chunk" % (n, 100.0*buf[i]/tot, buf[i]/npairs, buf[i]/npairs/125))
print("total cycles/pair %.0f = %.3f ms at 1.965 GHz" % (tot/npairs, tot/npairs/1.965e6))
