"""Randomised stress of the small-batch CTC path against the fp64 kernel (algo=1): shapes, ragged lengths, class counts,
logits / log-prob input, repeated calls (determinism), back-to-back calls of different shapes on one stream."""
import sys, time, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
g = torch.Generator(device="cuda").manual_seed(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 150
worst = (0.0, 0.0)
t_start = time.time()
for it in range(n_cases):
    C = [5, 20, 42, 43, 63][int(torch.randint(0, 5, (1,), generator=g, device="cuda"))]
    T = int(torch.randint(1, 420, (1,), generator=g, device="cuda"))
    B = int(torch.randint(1, 600, (1,), generator=g, device="cuda")) if it % 3 else int(torch.randint(1, 40, (1,), generator=g, device="cuda"))
    L = int(torch.randint(0, min(127, max(1, T // 2)) + 1, (1,), generator=g, device="cuda"))
    scale = [1.0, 1.0, 2.0, 0.3][it % 4]
    x = torch.randn(T, B, C, generator=g, device="cuda") * scale
    lp = torch.log_softmax(x, -1)
    tg = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
    il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
    tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
    n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
    n0, g0 = n0.clone(), g0.clone()
    fl = last_fallback_flags(B)
    n0b, g0b, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
    assert torch.equal(n0, n0b) and torch.equal(g0.nan_to_num(7.0), g0b.nan_to_num(7.0)), ("not reproducible", it, T, B, C, L)
    n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1, validate=False)
    fin = torch.isfinite(n1)
    assert torch.equal(torch.isfinite(n0), fin), ("finiteness", it, T, B, C, L)
    if fin.any():
        ne = float(((n0[fin] - n1[fin]).abs() / n1[fin].abs().clamp(min=1e-3)).max())
        ge = float((g0[:, fin] - g1[:, fin]).abs().max())
        worst = (max(worst[0], ne), max(worst[1], ge))
        assert ne < 1e-5 and ge < 1e-4, ("mismatch", it, T, B, C, L, ne, ge, int((fl != 0).sum()))
torch.cuda.synchronize()
print("stress ok: %d cases, worst nll rel %.2e grad abs %.2e, %.1f s" % (n_cases, worst[0], worst[1], time.time() - t_start))
