"""Where does the packed kernel's gradient error against fp64 sit?  (class blank / label / absent, frame position)"""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
def run(T, B, C, L, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g) * scale, -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.full((B,), T); tl = torch.full((B,), L)
    x = lp.double().clone().requires_grad_(True)
    nll = torch.nn.functional.ctc_loss(x, tg, il, tl, reduction="none"); nll.sum().backward()
    ref = x.grad
    for algo in (0, 1):
        n, gr, _ = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), algo=algo)
        torch.cuda.synchronize(); t0 = time.time()
        n, gr, _ = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), algo=algo)
        torch.cuda.synchronize(); dt = time.time() - t0
        e = (gr.cpu().double() - ref).abs()
        flat = int(e.argmax()); t, b, c = flat // (B * C), (flat // C) % B, flat % C
        present = torch.zeros(B, C, dtype=torch.bool); present.scatter_(1, tg, True)
        eb = float(e[:, :, 0].max()); 
        el = float(e[:, present.unsqueeze(0).expand(T, B, C)[0]].max()) if False else float((e * present[None].double()).max())
        absent = ~present; absent[:, 0] = False
        ea = float((e * absent[None].double()).max())
        print("T=%d L=%d scale %.1f algo %d: max %.2e at t=%d b=%d c=%d (ref %.3e) | blank %.2e label %.2e absent %.2e | mean %.2e | nll rel %.1e | %.2f ms"
              % (T, L, scale, algo, float(e.max()), t, b, c, float(ref[t, b, c]), eb, el, ea, float(e.mean()),
                 float(((n.cpu().double() - nll.detach()) / nll.detach()).abs().max()), dt * 1e3))
        if algo == 0:
            # per-frame row sum of the error (common-mode?) at the worst utterance
            rs = (gr.cpu().double() - ref)[:, b, :].sum(-1)
            print("   worst utterance: max |row-sum of error| %.2e ; err by frame quartile %s" % (
                float(rs.abs().max()), [float(e[q * T // 4:(q + 1) * T // 4, b].max()) for q in range(4)]))
run(1000, 32, 42, 120, 1)
run(1500, 16, 42, 180, 2)
run(1000, 32, 42, 120, 3, 2.0)
run(200, 32, 42, 30, 4)
