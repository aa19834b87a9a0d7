import sys, torch
sys.path.insert(0, '.')
from mddctc.loss import ctc_loss_raw, ctc_loss_streamed, last_fallback_flags
g = torch.Generator().manual_seed(15)
T, B, C, L = 300, 37, 42, 40
lp_h = torch.log_softmax(torch.randn(T, B, C, generator=g), -1).pin_memory()
tg = torch.randint(1, C, (B, L), generator=g).cuda()
il = torch.randint(150, T + 1, (B,), generator=g).cuda()
tl = torch.randint(0, L + 1, (B,), generator=g).cuda()
for trial in range(3):
    nll, grad, _ = ctc_loss_raw(lp_h.cuda(), tg, il, tl, grad_scalar=0.5)
    for slices in (1, 4, 5):
        total, nll_s, grads = ctc_loss_streamed(lp_h, tg, il, tl, slices=slices, grad_scalar=0.5)
        gs = torch.cat(grads, dim=1)
        d = (gs - grad).abs()
        print(trial, slices, 'equal', bool(torch.equal(gs, grad)), 'max diff', float(d.max()), 'nan', int(torch.isnan(gs).sum()), int(torch.isnan(grad).sum()),
              'utts', d.amax(dim=(0, 2)).nonzero().flatten().tolist()[:6], 'nll eq', bool(torch.equal(nll_s, nll)))
