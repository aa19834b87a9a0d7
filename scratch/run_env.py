"""aligned +16 posteriors with 10 % confidently wrong frames (B=1024, T=1000, L=120): a step in which all three tiers have work"""
import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from mddctc.loss import ctc_loss_raw, last_fallback_flags
from test_gpu_parity import _aligned_logits
g = torch.Generator(device='cuda').manual_seed(0)
T, B, C, L = 1000, 1024, 42, 120
tg = torch.randint(1, C, (B, L), generator=g, device='cuda')
il = torch.full((B,), T, device='cuda'); tl = torch.full((B,), L, device='cuda')
boost = float(sys.argv[1]) if len(sys.argv) > 1 else 16.0
lp = torch.log_softmax(_aligned_logits(T, B, C, L, il, tl, tg, boost, 0.1, g), -1)
for _ in range(3):
    ctc_loss_raw(lp, tg, il, tl, validate=False)
torch.cuda.synchronize()
print('left for the log-space kernel:', int((last_fallback_flags(B) != 0).sum()))
