"""Per-phase cycles of the team kernel (debug build with -DMDD_TEAM_TIMING): CTA 0 only."""
import ctypes, sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
L_ = mddctc._lib.lib()
names = ["fwd other", "fwd renorm+ckpt", "fwd wait e_full", "fwd alpha", "bwd other", "bwd wait ckpt", "bwd take+bookkeeping", "bwd wait eb_full",
         "bwd alpha rec", "bwd wait occ_free", "bwd beta+atomics"] + ["-"] * 5 + ["h fwd wait rows", "h fwd wait e_free", "h fwd commit+issue", "h bwd commit+issue",
         "h bwd wait occ_full", "h bwd gradient"] + ["-"] * 10
def run(T, B, C, L):
    g = torch.Generator(device="cuda").manual_seed(1)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g, device="cuda"), -1)
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.full((B,), T, device="cuda"); tl = torch.full((B,), L, device="cuda")
    for _ in range(2): ctc_loss_raw(lp, tg, il, tl, validate=False)
    buf = (ctypes.c_ulonglong * 32)()
    L_.mdd_debug_team_cycles(buf, 1)
    ctc_loss_raw(lp, tg, il, tl, validate=False)
    L_.mdd_debug_team_cycles(buf, 1)
    nch = (T + 7) // 8
    print("T=%d B=%d L=%d: cycles per chunk (CTA 0)" % (T, B, L))
    for i, n in enumerate(names):
        if n != "-": print("   %-24s %8.0f" % (n, buf[i] / nch))
    print("   lattice total %8.0f   helper0 total %8.0f" % (sum(buf[:16]) / nch, sum(buf[16:]) / nch))
run(200, 32, 42, 30)
run(1000, 512, 42, 120)
