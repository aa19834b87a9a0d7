import sys, ctypes, os, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw
Lb = mddctc._lib.lib()
T, B, C, L = 1000, int(os.environ.get("PT_B", "4096")), 42, 120
g = torch.Generator().manual_seed(1)
lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1).cuda()
tg = torch.randint(1, C, (B, L), generator=g).cuda()
il = torch.full((B,), T).cuda(); tl = torch.full((B,), L).cuda()
buf = (ctypes.c_ulonglong * 16)()
for _ in range(2): ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
Lb.mdd_debug_duo_cycles(buf, 1)
ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)
Lb.mdd_debug_duo_cycles(buf, 1)
names = ["lattice wait fw_full", "lattice wait bw_full", "lattice wait occ_free", "-", "lattice fwd total", "lattice bwd total", "-", "-",
         "rows wait fw_free", "rows wait occ_full", "rows fwd total", "rows bwd total", "rows commit", "rows issue loads", "rows grad rows", "rows setup"]
n = B // 2
for i, nm in enumerate(names):
    if nm != "-": print("%-24s %9.0f cycles/pair %8.1f /chunk" % (nm, buf[i] / n, buf[i] / n / 125))
