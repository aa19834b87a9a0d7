import sys, os, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
T, C, L = 1000, 42, 120
for B in [int(x) for x in sys.argv[1:]]:
    g = torch.Generator().manual_seed(1)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1).cuda()
    tg = torch.randint(1, C, (B, L), generator=g).cuda()
    il = torch.full((B,), T).cuda(); tl = torch.full((B,), L).cuda()
    for _ in range(3): ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("duo=%s B=%d pairs/SM=%.2f ms=%.3f  pairs/ms/SM=%.2f" % (os.environ.get("MDD_CTC_DUO", "0"), B, B / 2 / 148, ms, B / 2 / 148 / ms))
