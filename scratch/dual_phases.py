"""Cycle stamps of the dual kernel's CTA 0 (debug build with -DMDD_DUAL_TIMING)."""
import ctypes, sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
L_ = mddctc._lib.lib()
names = ["setup done", "alpha loop start", "alpha chunk 0 done", "alpha chunk 1 done", "alpha loop done", "beta loop start", "beta chunk 0 done", "beta loop done", "alpha chunk 2 done"]
def run(T, B, C, L):
    g = torch.Generator(device="cuda").manual_seed(1)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g, device="cuda"), -1)
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.full((B,), T, device="cuda"); tl = torch.full((B,), L, device="cuda")
    for _ in range(3): ctc_loss_raw(lp, tg, il, tl, validate=False)
    buf = (ctypes.c_ulonglong * 16)()
    L_.mdd_debug_dual_cycles(buf)
    print("T=%d B=%d L=%d" % (T, B, L))
    for i, n in enumerate(names): print("   %-22s %8d" % (n, buf[i]))
run(200, 32, 42, 30)
run(1000, 512, 42, 120)
