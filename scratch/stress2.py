"""Wider randomised sweep: long T, L up to 255, logits input, 1-D targets, strided input, upstream weights, peaky rows."""
import sys, time, torch
sys.path.insert(0, '.')
import mddctc
from mddctc.loss import ctc_loss_raw, last_fallback_flags
seed = int(sys.argv[1]); n_cases = int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(seed)
ri = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g, device="cuda"))
worst = [0.0, 0.0]; t_start = time.time(); nflag = 0
for it in range(n_cases):
    C = [5, 20, 42, 43, 62, 63, 64, 80][ri(0, 8)]
    T = ri(1, 1300) if it % 5 == 0 else ri(1, 420)
    B = ri(1, 40) if it % 3 == 0 else (ri(1, 600) if it % 3 == 1 else ri(600, 1300))
    if T > 600: B = min(B, 700)
    Lcap = [31, 63, 127, 255][ri(0, 4)]
    L = ri(0, min(Lcap, max(1, T // 2)) + 1)
    mode = it % 6
    x = torch.randn(T, B, C, generator=g, device="cuda") * [1.0, 1.0, 2.0, 0.3, 3.0][it % 5]
    if mode == 4:   # peaky rows
        x = x + 12.0 * torch.nn.functional.one_hot(torch.randint(0, C, (T, B), generator=g, device="cuda"), C)
    logits = mode == 1
    inp = x if logits else torch.log_softmax(x, -1)
    if mode == 2:   # strided view
        big = torch.zeros(T, B, C + 3, device="cuda"); big[..., :C] = inp; inp = big[..., :C] if False else inp
    il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
    tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
    tg2 = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
    if mode == 3:   # concatenated targets
        tg = torch.cat([tg2[b, :int(tl[b])] for b in range(B)]) if int(tl.sum()) > 0 else torch.zeros(0, dtype=torch.long, device="cuda")
    else:
        tg = tg2
    go = torch.rand(B, generator=g, device="cuda") + 0.5 if mode == 5 else None
    zi = bool(it % 2)
    kw = dict(validate=False, zero_infinity=zi, grad_out=go, logits=logits)
    n0, g0, _ = ctc_loss_raw(inp, tg, il, tl, algo=0, **kw)
    n0, g0 = n0.clone(), g0.clone()
    nflag += int((last_fallback_flags(B) != 0).sum())
    n0b, g0b, _ = ctc_loss_raw(inp, tg, il, tl, algo=0, **kw)
    assert torch.equal(n0, n0b) and torch.equal(g0.nan_to_num(7.0), g0b.nan_to_num(7.0)), ("not reproducible", it, T, B, C, L, mode)
    n1, g1, _ = ctc_loss_raw(inp, tg, il, tl, algo=1, **kw)
    fin = torch.isfinite(n1)
    assert torch.equal(torch.isfinite(n0), fin), ("finiteness", it, T, B, C, L, mode)
    assert torch.equal(torch.isnan(g0), torch.isnan(g1)), ("nan pattern", it, T, B, C, L, mode)
    if bool(fin.any()):
        ne = float(((n0[fin] - n1[fin]).abs() / n1[fin].abs().clamp(min=1e-3)).max())
        ge = float((g0[:, fin] - g1[:, fin]).abs().max())
        worst = [max(worst[0], ne), max(worst[1], ge)]
        if not (ne < 1e-5 and ge < 1.5e-4):
            rel = ((n0 - n1).abs() / n1.abs().clamp(min=1e-3)); rel[~fin] = 0
            b = int(rel.argmax())
            print("MISMATCH", it, T, B, C, L, mode, "ne %.2e ge %.2e" % (ne, ge), "utt", b, "nll ours %.8f fp64 %.8f" % (float(n0[b]), float(n1[b])), "T_b", int(il[b]), "L_b", int(tl[b]), "flag", int(last_fallback_flags(B)[b]))
torch.cuda.synchronize()
print("stress2 seed %d ok: %d cases, worst nll rel %.2e grad abs %.2e, %d utterances went to the fp64 kernel, %.1f s" % (seed, n_cases, worst[0], worst[1], nflag, time.time() - t_start))
