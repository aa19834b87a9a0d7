import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
g = torch.Generator().manual_seed(1)
T, B, C, L = 1000, 4096, 42, 120
lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
tg = torch.randint(1, C, (B, L), generator=g)
il = torch.full((B,), T)
il[::5] = torch.randint(500, T, (len(il[::5]),), generator=g)
tl = torch.full((B,), L)
nll, grad, st = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), algo=2, validate=False)
torch.cuda.synchronize()
f = last_fallback_flags(B).cpu()
idx = torch.nonzero(f).flatten()
print("flagged", len(idx), [(int(i), int(f[i]) & 255, int(f[i]) >> 8, int(il[i]), int(il[i ^ 1])) for i in idx[:12]])
for sl in (slice(3558, 3560), slice(3556, 3564)):
    lps = lp[:, sl].contiguous().cuda(); tgs = tg[sl].cuda(); ils = il[sl].cuda(); tls = tl[sl].cuda()
    n1, g1, _ = ctc_loss_raw(lps, tgs, ils, tls, algo=1); n1, g1 = n1.clone(), g1.clone()
    n2, g2, _ = ctc_loss_raw(lps, tgs, ils, tls, algo=2, validate=False)
    torch.cuda.synchronize()
    f = last_fallback_flags(lps.shape[1]).cpu()
    d = (g1 - g2).abs()
    print("slice", sl, "flags", f.tolist(), "nll", n1.tolist()[:2], n2.tolist()[:2], "grad diff", float(d.max()), "rowsum max", float(g2.sum(-1).abs().max()),
          "argmax t", int(d.amax(dim=(1, 2)).argmax()))
    rs = g2.sum(-1)  # (T, b)
    print("  worst rowsum per utt", rs.abs().amax(0).tolist())
sl = slice(3558, 3560)
lps = lp[:, sl].contiguous().cuda(); tgs = tg[sl].cuda(); ils = il[sl].cuda(); tls = tl[sl].cuda()
n1, g1, _ = ctc_loss_raw(lps, tgs, ils, tls, algo=1); n1, g1 = n1.clone(), g1.clone()
n2, g2, _ = ctc_loss_raw(lps, tgs, ils, tls, algo=2, validate=False)
d = (g1 - g2)[:, 0, :]
bad_t = torch.nonzero(d.abs().amax(1) > 1e-3).flatten().tolist()
print("bad frames", bad_t[:20], len(bad_t))
for t in bad_t[:3]:
    cs = torch.nonzero(d[t].abs() > 1e-3).flatten().tolist()
    print(t, [(c, round(float(g1[t, 0, c]), 5), round(float(g2[t, 0, c]), 3), int((tgs[0] == c).sum())) for c in cs])
