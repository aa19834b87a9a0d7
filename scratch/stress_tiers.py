"""Peaky (trained-model-like) posteriors at batch sizes that send the flagged utterances to each second-tier kernel;
every utterance of the product path against the fp64 kernel."""
import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
from test_gpu_parity import _aligned_logits
for (T, B, C, L, boost, noise) in ((600, 4096, 42, 80, 16, 0.1), (600, 4096, 62, 80, 20, 0.1), (1000, 2048, 42, 120, 16, 0.1), (333, 900, 42, 60, 25, 0.1),
                                   (600, 64, 42, 80, 20, 0.1), (200, 700, 20, 30, 16, 0.3)):
    g = torch.Generator(device="cuda").manual_seed(T + B)
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.randint(T * 2 // 3, T + 1, (B,), generator=g, device="cuda")
    tl = torch.randint(L // 4, L + 1, (B,), generator=g, device="cuda")
    lp = torch.log_softmax(_aligned_logits(T, B, C, L, il, tl, tg, boost, noise, g), -1)
    n2, g2, _ = ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
    f1 = int((last_fallback_flags(B) != 0).sum())
    n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
    n0, g0 = n0.clone(), g0.clone()
    f3 = int((last_fallback_flags(B) != 0).sum())
    n0b, g0b, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
    assert torch.equal(n0, n0b) and torch.equal(g0, g0b)
    n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1, validate=False)
    ne = float(((n0 - n1).abs() / n1.abs()).max()); ge = float((g0 - g1).abs().max())
    print("T=%d B=%d C=%d L=%d +%d: first pass flagged %d, left for fp64 %d | loss rel %.2e grad abs %.2e" % (T, B, C, L, boost, f1, f3, ne, ge), flush=True)
    assert ne < 1e-5 and ge < 1e-4
print("tiers ok")
