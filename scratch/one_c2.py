import sys, torch
sys.path.insert(0, '.')
from mddctc.loss import ctc_loss_raw
T, B, C, L = (int(x) for x in sys.argv[1:5])
g = torch.Generator(device="cuda").manual_seed(1)
lp = torch.log_softmax(torch.randn(T, B, C, generator=g, device="cuda"), -1)
tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
il = torch.full((B,), T, device="cuda"); tl = torch.full((B,), L, device="cuda")
for _ in range(3):
    ctc_loss_raw(lp, tg, il, tl, validate=False)
torch.cuda.synchronize()
