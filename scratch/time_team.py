"""Timing of mdd_ctc_loss at small / sharded batch sizes (team kernel vs the one-warp-per-pair kernel)."""
import os, sys, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
def run(T, B, C, L, steps=20):
    g = torch.Generator(device="cuda").manual_seed(1)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g, device="cuda"), -1)
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.full((B,), T, device="cuda"); tl = torch.full((B,), L, device="cuda")
    nll = torch.empty(B, device="cuda"); grad = torch.empty((T, B, C), device="cuda"); st = torch.empty(B, dtype=torch.int32, device="cuda")
    f = lambda: ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): f()
    e1.record(); torch.cuda.synchronize()
    print("T=%d B=%d L=%d: %.4f ms per call, flagged %d, team_max=%s" % (T, B, L, e0.elapsed_time(e1) / steps,
          int((last_fallback_flags(B) != 0).sum()), os.environ.get("MDD_CTC_TEAM_MAX_PAIRS")), flush=True)
for shape in ((200, 32, 42, 30), (200, 8, 42, 30), (200, 64, 42, 30), (1000, 512, 42, 120), (1000, 1024, 42, 120), (1000, 2048, 42, 120), (1000, 4096, 42, 120)):
    run(*shape)
