"""How much does the rounding of the exp argument (x*log2e in fp32, ex2.approx 2 ulp) move the CTC gradient?"""
import sys, numpy as np
sys.path.insert(0, '.')
from oracle import api as O
rng = np.random.default_rng(0)
def run(T, L, C, B, fam):
    x = rng.standard_normal((T, B, C)).astype(np.float32)
    tg = rng.integers(1, C, (B, L))
    if fam == 'randn':
        pass
    else:
        k = float(fam)
        # aligned peaky: a valid alignment path gets +k
        for b in range(B):
            # random monotone alignment
            pos = np.sort(rng.choice(T, L, replace=False))
            for i, t in enumerate(pos):
                x[t, b, tg[b, i]] += k
            mask = np.ones(T, bool); mask[pos] = False
            x[mask, b, 0] += k
    x64 = x.astype(np.float64)
    lp = (x64 - np.log(np.exp(x64 - x64.max(-1, keepdims=True)).sum(-1, keepdims=True)) - x64.max(-1, keepdims=True)).astype(np.float32)
    il = np.full(B, T); tl = np.full(B, L)
    ref = O.ctc_loss(lp, tg, il, tl)
    # perturbed: y = fl32(lp * log2e); p = 2^y (exact)  => lp' = y / log2e
    LOG2E = np.float32(1.4426950408889634)
    y = (lp * LOG2E).astype(np.float32)
    lp1 = y.astype(np.float64) / 1.4426950408889634
    r1 = O.ctc_loss(lp1, tg, il, tl)
    # plus 2-ulp relative noise on p
    noise = rng.uniform(-2, 2, lp.shape) * 2.0**-23
    lp2 = lp1 + np.log1p(noise)
    r2 = O.ctc_loss(lp2, tg, il, tl)
    lp3 = lp.astype(np.float64) + np.log1p(noise)
    r3 = O.ctc_loss(lp3, tg, il, tl)
    print(fam, "T", T, "L", L, "arg-rounding only: %.2e  + ex2 noise: %.2e  ex2 noise only: %.2e  nll/frame %.2f" % (
        np.abs(r1['grad'] - ref['grad']).max(), np.abs(r2['grad'] - ref['grad']).max(), np.abs(r3['grad'] - ref['grad']).max(), ref['nll'].mean() / T))
run(1000, 120, 42, 8, 'randn')
run(1500, 180, 42, 8, 'randn')
run(1000, 120, 42, 8, '10')
run(1000, 120, 42, 8, '16')
run(1000, 120, 42, 8, '25')
