"""Random configurations of the two beam decoders against the C oracle (shapes, beam widths, class counts, eos on/off)."""
import sys, os, numpy as np, torch
sys.path.insert(0, '.')
import mddctc
from oracle import api as O
from mddctc.decoder import BeamDecoder
from mddctc.prefix import joint_beam_search
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
rng = np.random.default_rng(seed)
g = torch.Generator().manual_seed(seed)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
assets = os.path.join(ROOT, "tests", "golden", "assets")
vocab43 = {0: "blank", 1: "UNK"}
for line in open(os.path.join(assets, "units")):
    if line.strip(): vocab43[len(vocab43)] = line.strip()
arpa = os.path.join(assets, "lm_phone_bg.arpa")
lm = O.lm_table(arpa, vocab43, 43)
for it in range(14):
    W = int(rng.choice([1, 2, 5, 10, 37])); alpha = float(rng.choice([0.0, 0.3, 1.0]))
    B = int(rng.integers(1, 40)); T = int(rng.integers(1, 260))
    x = rng.standard_normal((B, T, 43)).astype(np.float32) * float(rng.choice([0.5, 1.0, 3.0]))
    if it % 2: x[:, :, 0] += 5.0 * (rng.random((B, T)) < 0.6)
    if it % 5 == 0: x = np.round(x)
    P = torch.softmax(torch.from_numpy(x), -1).numpy()
    lens = rng.integers(0 if it % 3 == 0 else 1, T + 1, size=B)
    dec = BeamDecoder(vocab43, beam_width=W, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=alpha)
    toks, olen, st, score = O.beam_tokens(P, lens, W, 0, lm, alpha)
    t2, l2, s2, sc2 = dec.decode_probs_tokens(torch.from_numpy(P), lens.tolist())
    t2, l2, s2, sc2 = t2.cpu().numpy(), l2.cpu().numpy(), s2.cpu().numpy(), sc2.cpu().numpy()
    assert s2.tolist() == st.tolist(), ("status", it, W, B, T)
    assert l2.tolist() == olen.tolist(), ("len", it, W, B, T)
    for b in range(B):
        assert t2[b, :l2[b]].tolist() == toks[b, :olen[b]].tolist(), ("tokens", it, W, B, T, b)
print("beam ok")
for it in range(10):
    C = int(rng.choice([12, 43, 50])); W = int(rng.choice([1, 3, 10])); T = int(rng.integers(5, 220)); B = int(rng.integers(1, 20))
    steps = int(rng.integers(1, 14)); eos = None if it % 2 else C - 1; lam = float(rng.choice([0.0, 0.3, 1.0]))
    xl = torch.randn(T, B, C, generator=g); xl[:, :, 0] += 3.0
    x = torch.log_softmax(xl, -1)
    lens = torch.randint(1, T + 1, (B,), generator=g)
    table = torch.log_softmax(torch.randn(5, C + 1, C, generator=g) * 2.0, -1); td = table.cuda()
    tok, sc, gl, fl = joint_beam_search(x.cuda(), lens.cuda(), lambda s, stt: td[s % 5][(stt["last"] + 1).long()], beam=W, ctc_weight=lam, max_len=steps, eos=eos)
    wt, ws, wg, wf = O.joint_beam_search(x.numpy(), lens.numpy(), lambda s, stt: table[s % 5][torch.from_numpy(stt["last"] + 1).long()].numpy(), beam=W, ctc_weight=lam, max_len=steps, eos=eos)
    assert np.array_equal(fl.cpu().numpy(), wf) and np.array_equal(gl.cpu().numpy(), wg) and np.array_equal(tok.cpu().numpy(), wt), ("joint", it, C, W, T, B, steps, eos, lam)
    got = sc.cpu().numpy(); fin = np.isfinite(ws)
    assert np.array_equal(np.isfinite(got), fin)
    np.testing.assert_allclose(got[fin], ws[fin], rtol=1e-9, atol=1e-9)
print("joint ok")
