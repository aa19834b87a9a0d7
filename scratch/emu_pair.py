"""numpy emulation of the pair kernel's forward sweep (one utterance, NSL=8): where does the block-floating arithmetic lose mass?"""
import sys, numpy as np
f32 = np.float32
NSL, KF, GAP, SHIFT, E_EMPTY = 8, 8, int(sys.argv[2]) if len(sys.argv) > 2 else 36, int(sys.argv[3]) if len(sys.argv) > 3 else 0, -(1 << 20)
MINN = f32(2.0 ** -126)
def ftz(x):
    x = np.asarray(x, dtype=f32)
    return np.where(np.abs(x) < MINN, f32(0), x).astype(f32)
def pow2i(x):
    x = np.clip(x, -127, 127)
    return np.where(x <= -127, f32(0), np.exp2(np.maximum(x, -126).astype(np.float64))).astype(f32)
z = np.load(sys.argv[1]); lp = z['lp'].astype(np.float64); tg = z['tg']; T, C = lp.shape; L = len(tg); S = 2 * L + 1
ext = np.zeros(S, dtype=int); ext[1::2] = tg
# fp64 truth in log space (alpha, beta)
la = np.full((T, S), -np.inf); la[0, 0] = lp[0, 0]; la[0, 1] = lp[0, ext[1]]
skip = np.zeros(S, bool); skip[3::2] = ext[3::2] != ext[1:-2:2]
def lse(*a):
    m = np.maximum.reduce(a); m2 = np.where(np.isinf(m), 0, m)
    return m2 + np.log(sum(np.exp(x - m2) for x in a))
for t in range(1, T):
    p = la[t - 1]; p1 = np.concatenate([[-np.inf], p[:-1]]); p2 = np.concatenate([[-np.inf, -np.inf], p[:-2]]); p2 = np.where(skip, p2, -np.inf)
    la[t] = lse(p, p1, p2) + lp[t, ext]
lb = np.full((T, S), -np.inf); lb[T - 1, S - 1] = lp[T - 1, 0]; lb[T - 1, S - 2] = lp[T - 1, ext[S - 2]]
skipb = np.zeros(S, bool); skipb[:-2] = skip[2:]
for t in range(T - 2, -1, -1):
    p = lb[t + 1]; p1 = np.concatenate([p[1:], [-np.inf]]); p2 = np.concatenate([p[2:], [-np.inf, -np.inf]]); p2 = np.where(skipb, p2, -np.inf)
    lb[t] = lse(p, p1, p2) + lp[t, ext]
logP = float(lse(la[T - 1, S - 1], la[T - 1, S - 2]))
gam = la + lb - lp[:, ext] - logP           # log occupancy
print('truth nll %.4f (log kernel %.4f, pair kernel %.4f)' % (-logP, z['nll_log'], z['nll_pair']))
# emulation
e = ftz(np.exp2((lp * 1.4426950408889634).astype(f32).astype(np.float64)).astype(f32))       # (T,C)
a = np.zeros((32, NSL), f32); a[0, 0] = 1; ea = np.zeros(32, int)
ma = np.zeros((32, NSL), f32)
for s in range(S):
    if skip[s]: ma[s // NSL, s % NSL] = 1
em = np.zeros((T, 32 * NSL), f32)
for s in range(32 * NSL): pass
def emis(t):
    r = np.zeros(32 * NSL, f32); r[:S] = e[t, ext]; return r.reshape(32, NSL)
t = 0; lost_report = 0
chunks = [(i * KF, KF) for i in range(T // KF)] + [(T // KF * KF + i, 1) for i in range(T % KF)]
for (t0, nf) in chunks:
    m = a.max(axis=1); some = (m >= MINN) & np.isfinite(m)
    k = np.floor(np.log2(np.where(some, m, 1).astype(np.float64))).astype(int)
    e_own = np.where(some, ea + k - SHIFT, E_EMPTY)
    tt = e_own + GAP * np.arange(32); tt = np.maximum.accumulate(tt); e_new = tt - GAP * np.arange(32)
    f = pow2i(ea - e_new); squash = np.where(some, e_new - e_own, 0)
    ea = e_new.copy(); sc = np.zeros(32, f32); sc[1:] = pow2i(e_new[:-1] - e_new[1:])
    a = ftz(a * f[:, None])
    for r in range(nf):
        t = t0 + r
        p = np.zeros(32, f32); p[1:] = a[:-1, NSL - 1]; p = ftz(p * sc)
        E = emis(t); n = np.zeros_like(a)
        for j in range(NSL):
            m1 = a[:, j - 1] if j else p
            if j & 1:
                m2 = a[:, j - 2] if j >= 2 else p
                n[:, j] = ftz(ftz(ftz(ma[:, j] * m2) + ftz(a[:, j] + m1)) * E[:, j])
            else:
                n[:, j] = ftz(ftz(a[:, j] + m1) * E[:, j])
        a = n
        # compare with the truth where occupancy matters
        with np.errstate(divide='ignore'):
            emu = np.log(a.reshape(-1)[:S].astype(np.float64)) + np.repeat(ea, NSL)[:S] * np.log(2.0)
        imp = gam[t] > np.log(1e-3)
        bad = imp & ~(np.abs(emu - la[t]) < 1e-2)
        if bad.any() and lost_report < 6:
            lost_report += 1
            s = int(np.nonzero(bad)[0][0]); ln = s // NSL
            print('t=%d (chunk t0=%d) state %d (lane %d, j=%d) occupancy %.3g: true log alpha %.2f emulated %.2f | lane exp %d, lane mantissas log2 %s | left lane exp %d right %d' % (
                t, t0, s, ln, s % NSL, np.exp(gam[t, s]), la[t, s], emu[s], ea[ln], np.round(np.log2(np.maximum(a[ln].astype(np.float64), 1e-300))).astype(int).tolist(), ea[ln - 1] if ln else 0, ea[ln + 1] if ln < 31 else 0))
            print('      true log2 alpha of the lane rel. to its exponent: %s' % np.round(la[t, ln * NSL:(ln + 1) * NSL] / np.log(2) - ea[ln]).astype(int).tolist())
v = a.reshape(-1)[[S - 1, S - 2]].astype(np.float64) * np.exp2(np.repeat(ea, NSL)[[S - 1, S - 2]].astype(np.float64))
print('emulated nll %.4f' % -np.log(v.sum()))
