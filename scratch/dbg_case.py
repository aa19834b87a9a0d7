import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
T, B, C, L = (int(v) for v in sys.argv[1:5])
g = torch.Generator(device="cuda").manual_seed(5)
x = torch.randn(T, B, C, generator=g, device="cuda")
lp = torch.log_softmax(x, -1)
tg = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
if len(sys.argv) > 5 and sys.argv[5] == "full":
    il[:] = T; tl[:] = min(L, (T + 1) // 2)
n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
fl = last_fallback_flags(B)
n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1, validate=False)
fin = torch.isfinite(n1)
ge = (g0[:, fin] - g1[:, fin]).abs().amax(dim=(0, 2))
bad = (ge > 1e-4).nonzero().flatten()
idx = fin.nonzero().flatten()[bad]
print(sys.argv[1:], "max grad err %.3e, bad utterances %d of %d:" % (float(ge.max()), bad.numel(), int(fin.sum())), idx[:12].tolist(),
      "T_b", il[idx[:6]].tolist(), "L_b", tl[idx[:6]].tolist(), "flags", fl[idx[:6]].tolist())
if bad.numel():
    b = int(idx[0]); d = (g0[:, b] - g1[:, b]).abs()
    tt = d.amax(dim=1).nonzero().flatten()
    print("  utterance", b, "frames with error:", tt[:10].tolist(), "...", tt[-3:].tolist(), "classes:", d.amax(dim=0).nonzero().flatten()[:12].tolist())
