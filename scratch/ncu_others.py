"""Runs every product kernel that is not the CTC pair kernel once at its BASELINE shape (for ncu --set full)."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from mddctc.decoder import BeamDecoder, GreedyDecoder
from mddctc.mdd import MDDScorer
from mddctc.prefix import joint_beam_search
from mddctc.corpus import CorpusMDD
from mddctc.loss import ctc_loss_raw, row_lse
dev = torch.device("cuda", 0)
vocab = bench.phone_vocab()
reps = 2
# greedy (argmax_rows + collapse) at the config-3 shape, then config 5 alignment (L ~ 40 and L ~ 165)
g = torch.Generator(device=dev).manual_seed(3)
T, B, C = 1000, 4096, 42
lp = torch.log_softmax(torch.randn(T, B, C, device=dev, generator=g), -1)
gd42 = GreedyDecoder({i: "p%d" % i for i in range(C)}, space_idx=-1, blank_index=0)
lens = torch.full((B,), T, device=dev)
for _ in range(reps): gd42.decode_tokens(lp, lens)
x = torch.randn(T, B, C, device=dev, generator=g) * 2
for _ in range(reps): row_lse(x, want_argmax=True)
# the fp64 kernel on a 256-utterance slice, the team kernel at B = 512 (strong-scaling share of config 3)
tg = torch.randint(1, C, (B, 120), device=dev, generator=g)
tl = torch.full((B,), 120, device=dev)
for _ in range(reps): ctc_loss_raw(lp[:, :256], tg[:256], lens[:256], tl[:256], algo=1, validate=False)
for _ in range(reps): ctc_loss_raw(lp[:, :512], tg[:512], lens[:512], tl[:512], validate=False)
del lp, x
sc = MDDScorer(vocab)
for Lh in (40, 165):
    Bn = 2048
    dec = torch.randint(2, 43, (Bn, Lh + 10), device=dev, generator=g, dtype=torch.int32)
    can = torch.randint(2, 43, (Bn, Lh + 10), device=dev, generator=g, dtype=torch.int32)
    dl = torch.randint(Lh - 5, Lh + 6, (Bn,), device=dev, generator=g, dtype=torch.int32)
    cl = torch.randint(Lh - 5, Lh + 6, (Bn,), device=dev, generator=g, dtype=torch.int32)
    for _ in range(reps): sc.align_tokens(dec, dl, can, cl)
# config 4: both joint decoders
lp4, lens4 = bench.c4_inputs(1024, 500, 43, 4, dev)
arpa = os.path.join(ROOT, "tests", "golden", "assets", "lm_phone_bg.arpa")
bd = BeamDecoder(vocab, beam_width=10, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=0.3)
probs = torch.exp(lp4.transpose(0, 1)).contiguous()
for _ in range(reps): bd.decode_probs_tokens(probs, lens4)
table = bench.synthetic_attention(43).to(dev)
for _ in range(reps): joint_beam_search(lp4, lens4, lambda s, st: table[s % 7][(st["last"] + 1).long()], beam=10, ctc_weight=0.3, max_len=3, eos=42)
# corpus counts
ev = CorpusMDD(vocab)
Bn, Lc = 2048, 60
rows = lambda: (torch.randint(3, 42, (Bn, Lc), device=dev, generator=g, dtype=torch.int32), torch.randint(20, Lc + 1, (Bn,), device=dev, generator=g, dtype=torch.int32))
(D, dl), (Lb, ll), (Cn, cl) = rows(), rows(), rows()
for _ in range(reps): ev.evaluate_tokens(D, dl, Lb, ll, Cn, cl)
torch.cuda.synchronize()
print("done")
