import sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
seed = int(sys.argv[1]); stop = int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(seed)
ri = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g, device="cuda"))
for it in range(stop + 1):
    C = [5, 20, 42, 43, 62, 63][ri(0, 6)]
    T = ri(1, 420)
    B = ri(1, 40) if it % 3 == 0 else (ri(1, 600) if it % 3 == 1 else ri(600, 1500))
    L = ri(0, min(127, max(1, T // 2)) + 1)
    x = torch.randn(T, B, C, generator=g, device="cuda") * [1.0, 1.0, 2.0, 0.3][it % 4]
    lp = torch.log_softmax(x, -1)
    tg = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
    il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
    tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
print("case", it, T, B, C, L, "scale", [1.0, 1.0, 2.0, 0.3][it % 4])
for rep in range(3):
    n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
    n0, g0 = n0.clone(), g0.clone()
    fl = last_fallback_flags(B)
    na, ga, _ = ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
    fa = last_fallback_flags(B)
    n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1, validate=False)
    fin = torch.isfinite(n1)
    ge = (g0[:, fin] - g1[:, fin]).abs().amax(dim=(0, 2))
    bad = (ge > 1e-4).nonzero().flatten()
    idx = fin.nonzero().flatten()[bad]
    print("rep", rep, "max grad err %.3e, bad %d of %d:" % (float(ge.max()), bad.numel(), int(fin.sum())), idx[:12].tolist(),
          "T_b", il[idx[:6]].tolist(), "L_b", tl[idx[:6]].tolist(), "flags after all tiers", fl[idx[:6]].tolist(), "first-pass flags", fa[idx[:6]].tolist(),
          "n flagged first pass", int((fa != 0).sum()))
    if bad.numel():
        b = int(idx[0]); d = (g0[:, b] - g1[:, b]).abs()
        tt = d.amax(dim=1).nonzero().flatten()
        print("  utterance", b, "partner T/L", int(il[b ^ 1]), int(tl[b ^ 1]), "frames with error:", tt[:6].tolist(), "...", tt[-3:].tolist(), "n", tt.numel())
        d2 = (ga[:, b] - g1[:, b]).abs()
        print("  first pass alone (algo=2) err for it: %.3e" % float(d2.max()))
b = 46
d = (g0[:, b] - g1[:, b])
Tb = int(il[b])
torch.set_printoptions(precision=4, linewidth=200, sci_mode=False)
print("labels", tg[b, :int(tl[b])].tolist())
for t in (0, 1, 2, 25, 47, 48, 49, 50, 51):
    row = d[t]
    nz = (row.abs() > 1e-5).nonzero().flatten().tolist()
    print("t", t, "rowsum ours %.2e ref %.2e" % (float(g0[t, b].sum()), float(g1[t, b].sum())), "classes off:", nz[:12], [round(float(row[c]), 4) for c in nz[:8]])
