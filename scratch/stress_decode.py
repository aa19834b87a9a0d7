"""Wider random shapes for the integer kernels (Levenshtein + traceback, greedy collapse) against the C oracle."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import mddctc
from oracle import api as O
from mddctc.decoder import edit_align_tokens, GreedyDecoder
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n = 0
for it in range(120):
    B = int(rng.integers(1, 400)); Lh = int(rng.integers(1, [8, 40, 130, 300, 700][it % 5])); Lr = int(rng.integers(1, [8, 40, 130, 300, 700][(it // 5) % 5])); V = int(rng.integers(2, 45))
    if Lh * Lr > 100000: B = min(B, 24)
    hl = rng.integers(0, Lh + 1, size=B).astype(np.int32); rl = rng.integers(0, Lr + 1, size=B).astype(np.int32)
    H = rng.integers(0, V, size=(B, Lh)).astype(np.int32); R = rng.integers(0, V, size=(B, Lr)).astype(np.int32)
    for b in range(0, B, 2):
        k = min(hl[b], rl[b]); H[b, :k] = R[b, :k]
        flip = rng.random(k) < 0.15; H[b, :k][flip] = rng.integers(0, V, size=int(flip.sum()))
    dist, paths, st = O.edit_align([H[b, :hl[b]].tolist() for b in range(B)], [R[b, :rl[b]].tolist() for b in range(B)])
    r = edit_align_tokens(torch.from_numpy(H).cuda(), torch.from_numpy(hl).cuda(), torch.from_numpy(R).cuda(), torch.from_numpy(rl).cuda(), Lh, Lr, want_ops=True)
    assert r["dist"].cpu().numpy().tolist() == dist.tolist(), (it, B, Lh, Lr, V)
    assert r["status"].cpu().numpy().tolist() == st.tolist(), (it, B, Lh, Lr, V)
    ops, ol = r["ops"].cpu().numpy(), r["ops_len"].cpu().numpy()
    for b in range(B):
        assert "".join(chr(c) for c in ops[b, :ol[b]]) == "".join(paths[b]), (it, B, Lh, Lr, b)
    n += B
print("edit_align ok:", n, "pairs")
vocab = {i: ("blank" if i == 0 else "p%d" % i) for i in range(43)}
dec = GreedyDecoder(vocab, space_idx=-1, blank_index=0)
for it in range(40):
    T = int(rng.integers(1, 700)); B = int(rng.integers(1, 300))
    x = np.round(rng.standard_normal((T, B, 43)).astype(np.float32) * 2) / 2
    run = rng.integers(0, 43, size=(T, B))
    for t in range(1, T):
        keep = rng.random(B) < 0.7; run[t, keep] = run[t - 1, keep]
    x[np.arange(T)[:, None], np.arange(B)[None, :], run] += 2.5
    lens = rng.integers(0, T + 1, size=B)
    assert dec.decode(torch.from_numpy(x).cuda(), lens.tolist()) == O.greedy_decode(x, lens, vocab), (it, T, B)
print("greedy ok")
