"""Which flags the dual path raises on the envelope families; per-kernel times at C2 / B=512."""
import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
from test_gpu_parity import _aligned_logits, _torch_cpu_ctc
for name, boost, par in (("aligned+6", 6, 0.1), ("aligned+10", 10, 0.1), ("aligned+10 clean", 10, 0.0), ("randn*2", None, 2.0)):
    g = torch.Generator(device="cuda").manual_seed(21)
    T, B, C, L = 600, 256, 42, 80
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.randint(400, T + 1, (B,), generator=g, device="cuda")
    tl = torch.randint(20, L + 1, (B,), generator=g, device="cuda")
    x = torch.randn(T, B, C, generator=g, device="cuda") * par if boost is None else _aligned_logits(T, B, C, L, il, tl, tg, boost, par, g)
    lp = torch.log_softmax(x, -1)
    n2, g2, _ = ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
    fl = last_fallback_flags(B)
    bad = fl.nonzero().flatten().tolist()
    print(name, "flagged:", [(b, int(fl[b]) & 255, int(fl[b]) >> 8, int(il[b]), int(tl[b])) for b in bad][:10], flush=True)
    ref_nll, ref_grad = _torch_cpu_ctc(lp, tg, il, tl, dtype=torch.float64)
    ge = (g2.double().cpu() - ref_grad).abs().amax(dim=(0, 2))
    keep = (fl == 0).cpu()
    print("   grad err kept: %.3e  nll rel %.3e" % (float(ge[keep].max()), float(((n2.double().cpu() - ref_nll) / ref_nll).abs()[keep].max())))
