import sys, torch
sys.path.insert(0, '.')
from mddctc.loss import ctc_loss_raw, last_fallback_flags
g = torch.Generator(device='cuda').manual_seed(2)
for (T, B, C, L, scale) in ((200, 63, 42, 30, 4.0), (300, 33, 42, 100, 3.0), (150, 17, 45, 50, 5.0), (400, 9, 43, 200, 3.0)):
    x = torch.randn(T, B, C, generator=g, device='cuda') * scale
    tg = torch.randint(1, C, (B, L), generator=g, device='cuda')
    il = torch.randint(T // 2, T + 1, (B,), generator=g, device='cuda'); tl = torch.randint(0, L + 1, (B,), generator=g, device='cuda')
    am = torch.empty((T, B), dtype=torch.uint8, device='cuda')
    for logits in (False, True):
        inp = x if logits else torch.log_softmax(x, -1)
        n0, g0, _ = ctc_loss_raw(inp, tg, il, tl, algo=0, validate=False, logits=logits, argmax=am if logits else None)
        n1, g1, _ = ctc_loss_raw(torch.log_softmax(x, -1), tg, il, tl, algo=1, validate=False)
        torch.cuda.synchronize()
        fin = torch.isfinite(n1)
        print(T, B, L, 'logits' if logits else 'lp', 'left', int((last_fallback_flags(B) != 0).sum()) if False else '-', 'nll rel %.2e grad %.2e' % (
            float(((n0 - n1) / n1).abs()[fin].max()), float((g0 - g1).abs().amax(dim=(0, 2))[fin].max())))
# 1-D concatenated targets through all tiers
T, B, C, L = 200, 21, 42, 30
x = torch.randn(T, B, C, generator=g, device='cuda') * 4
tl = torch.randint(1, L + 1, (B,), generator=g, device='cuda'); il = torch.full((B,), T, device='cuda')
tg2 = torch.randint(1, C, (B, L), generator=g, device='cuda')
tg1 = torch.cat([tg2[b, :int(tl[b])] for b in range(B)])
a = ctc_loss_raw(torch.log_softmax(x, -1), tg1, il, tl, algo=0); a = (a[0].clone(), a[1].clone())
b = ctc_loss_raw(torch.log_softmax(x, -1), tg2, il, tl, algo=0)
print('1-D vs 2-D targets equal:', bool(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])))
# loss only
n, gg, _ = ctc_loss_raw(torch.log_softmax(x, -1), tg2, il, tl, algo=0, want_grad=False)
print('loss-only equal:', bool(torch.allclose(n, b[0], rtol=1e-6)))
