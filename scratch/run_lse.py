import sys, torch
sys.path.insert(0, '.')
from mddctc.loss import ctc_loss_raw
T, B, C, L = 1000, 4096, 42, 120
g = torch.Generator(device='cuda').manual_seed(5)
x = torch.randn(T, B, C, device='cuda', generator=g)
tg = torch.randint(1, C, (B, L), device='cuda', generator=g)
il = torch.full((B,), T, device='cuda'); tl = torch.full((B,), L, device='cuda')
am = torch.empty((T, B), dtype=torch.uint8, device='cuda')
for _ in range(3):
    ctc_loss_raw(x, tg, il, tl, validate=False, logits=True, argmax=am)
torch.cuda.synchronize()
