"""One CTC shape, a few calls (for ncu launch lists): python scratch/one_shape.py T B C L [calls]"""
import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
T, B, C, L = (int(x) for x in sys.argv[1:5])
n = int(sys.argv[5]) if len(sys.argv) > 5 else 3
g = torch.Generator(device="cuda").manual_seed(1)
lp = torch.log_softmax(torch.randn(T, B, C, generator=g, device="cuda"), -1)
tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
il = torch.full((B,), T, device="cuda"); tl = torch.full((B,), L, device="cuda")
nll = torch.empty(B, device="cuda"); grad = torch.empty((T, B, C), device="cuda"); st = torch.empty(B, dtype=torch.int32, device="cuda")
ls = torch.empty(1, device="cuda")
for _ in range(n):
    ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st)
torch.cuda.synchronize()
print("ok", float(nll.sum()))
