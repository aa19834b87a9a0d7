"""Writes tests/golden/* by EXECUTING the real reference (and torch CPU ctc_loss) in the build container.

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so the outputs are committed as small
fixtures next to this script.  What each file pins (reference file:line in parentheses):

  assets/units, assets/lm_phone_bg.arpa   data copies of egs/attention_aug/data/{units,lm_phone_bg.arpa} (model assets)
  ctc_small.npz        torch 2.11 CPU F.ctc_loss fp32/fp64 nll+grad+alpha (call site steps/train_ctc.py:186,72,96)
  ctc_logits_small.npz torch 2.11 CPU log_softmax -> ctc_loss -> autograd to the logits (models/model_ctc.py:214-218 +
                       steps/train_ctc.py:72,96), fp32/fp64 nll + logit gradient + logsumexp + argmax
  textio.json.gz       create_input lengths + train_ctc.py:68 truncation, nosil strings / log lines of test_ctc_nosil.py,
                       CMU -> IPA rendering of infer.py:331-333
  greedy.npz           GreedyDecoder.decode strings (utils/ctcDecoder.py:186-200)
  beam_small.npz       BeamDecoder.decode strings / exceptions, 160 random small cases (utils/BeamSearch.py:73-153)
  beam_c4.npz          same at the C4 shape (T=500, C=43, W=10, lm_alpha 0 and 0.3), 6 utterances
  lm_table.npz         LanguageModel.get_bi_prob over the 43-class vocab (utils/NgramLM.py:65-78)
  wer_corpus.json.gz   900-utterance corpora from result/ logs + totals + per-pair wer() outputs
  mdd_post.json.gz     wer -> align_canonical_decoded -> stastics -> score (infer.py:324-342) on 600 noisy pairs + KATs
  corpus_mdd.json.gz   test_ctc_nosil.py:218-297 (TA/FR/FA/TR counters, PER) per utterance and in total on 500 synthetic
                       (canonical, human label, recognised) triples; the d maps of print_align_space_canonical_origin
  c1_words.json.gz     BASELINE config 1: the 20 words of egs/vocabulary/single, canonical phones by the rule at
                       infer.py:549-551 from dict/cmudict.dict, synthetic posteriors, greedy + MDD outputs
"""
from __future__ import annotations

import gzip
import json
import math
import os
import shutil
import string
import sys

import numpy as np
import torch

from . import ref_loader as RL

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def jdump(name, obj):
    with gzip.open(os.path.join(GOLD, name), "wt") as f:
        json.dump(obj, f)


# ----------------------------------------------------------------------------------------------- CTC
def gen_ctc():
    g = torch.Generator().manual_seed(1234)
    cases = {}

    def run(name, T, B, C, Lmax, in_len, tgt_len, targets=None, one_d=False, zero_inf=False):
        lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
        if targets is None:
            targets = torch.randint(1, C, (B, Lmax), generator=g)
        in_len_t = torch.tensor(in_len, dtype=torch.long)
        tgt_len_t = torch.tensor(tgt_len, dtype=torch.long)
        tg = targets
        if one_d:
            tg = torch.cat([targets[b, :tgt_len[b]] for b in range(B)])
        out = {"lp": lp.numpy(), "targets": tg.numpy(), "in_len": in_len_t.numpy(), "tgt_len": tgt_len_t.numpy(),
               "zero_inf": np.array(int(zero_inf))}
        for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            x = lp.to(dt).clone().requires_grad_(True)
            nll = torch.nn.functional.ctc_loss(x, tg, in_len_t, tgt_len_t, blank=0, reduction="none",
                                               zero_infinity=zero_inf)
            nll.sum().backward()
            out["nll_" + tag] = nll.detach().numpy()
            out["grad_" + tag] = x.grad.numpy()
        # lattice (alpha beyond input_length is uninitialised in aten: masked to -inf here)
        nll2, la = torch._ctc_loss(lp.double(), tg, in_len, tgt_len, 0, zero_inf)
        la = la.clone()
        for b in range(B):
            la[b, in_len[b]:] = -math.inf
            la[b, :, 2 * tgt_len[b] + 1:] = -math.inf
        out["alpha_f64"] = la.numpy()
        for k, v in out.items():
            cases[name + "/" + k] = v

    run("basic", 25, 4, 7, 6, [25, 25, 25, 25], [6, 6, 6, 6])
    # forced repeats, L=0, ragged T_b
    tg = torch.tensor([[3, 3, 3, 2, 2, 1], [1, 2, 3, 4, 5, 6], [0, 0, 0, 0, 0, 0], [5, 5, 1, 1, 0, 0]])
    run("repeat_ragged", 25, 4, 7, 6, [25, 17, 9, 12], [6, 6, 0, 4], targets=tg)
    run("one_d", 20, 3, 9, 5, [20, 11, 15], [5, 2, 3], one_d=True)
    # infeasible row (T_b < L_b + repeats) -> inf loss, NaN grad (zero_infinity False) / zeros (True)
    tg2 = torch.tensor([[2, 2, 2, 2], [1, 2, 3, 4], [4, 3, 2, 1]])
    run("infeasible", 12, 3, 6, 4, [5, 12, 3], [4, 4, 4], targets=tg2)
    run("infeasible_zi", 12, 3, 6, 4, [5, 12, 3], [4, 4, 4], targets=tg2, zero_inf=True)
    run("c42", 60, 5, 42, 30, [60, 47, 33, 60, 31], [30, 12, 9, 1, 15])
    run("tiny", 1, 2, 5, 1, [1, 1], [1, 0])
    np.savez_compressed(os.path.join(GOLD, "ctc_small.npz"), **cases)


def gen_ctc_logits():
    """The reference's own tail: out = F.log_softmax(fc(x), -1) (model_ctc.py:214-218), loss = CTCLoss('sum')(out, ...),
    loss.backward() -- gradient taken at the LOGITS."""
    g = torch.Generator().manual_seed(4321)
    cases = {}

    def run(name, T, B, C, Lmax, in_len, tgt_len, targets=None, zero_inf=False, scale=3.0, weights=False):
        x0 = torch.randn(T, B, C, generator=g) * scale
        if targets is None:
            targets = torch.randint(1, C, (B, Lmax), generator=g)
        il, tl = torch.tensor(in_len, dtype=torch.long), torch.tensor(tgt_len, dtype=torch.long)
        go = torch.rand(B, generator=g) + 0.5 if weights else torch.ones(B)
        out = {"logits": x0.numpy(), "targets": targets.numpy(), "in_len": il.numpy(), "tgt_len": tl.numpy(),
               "zero_inf": np.array(int(zero_inf)), "grad_out": go.numpy()}
        for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            x = x0.to(dt).clone().requires_grad_(True)
            lp = torch.log_softmax(x, -1)
            nll = torch.nn.functional.ctc_loss(lp, targets, il, tl, blank=0, reduction="none", zero_infinity=zero_inf)
            (nll * go.to(dt)).sum().backward()
            out["nll_" + tag] = nll.detach().numpy()
            out["grad_" + tag] = x.grad.numpy()
            out["lse_" + tag] = torch.logsumexp(x.detach(), -1).numpy()
        out["argmax"] = torch.max(x0, dim=-1)[1].numpy()
        for k, v in out.items():
            cases[name + "/" + k] = v

    run("basic", 25, 4, 7, 6, [25, 25, 25, 25], [6, 6, 6, 6])
    tg = torch.tensor([[3, 3, 3, 2, 2, 1], [1, 2, 3, 4, 5, 6], [0, 0, 0, 0, 0, 0], [5, 5, 1, 1, 0, 0]])
    run("repeat_ragged", 25, 4, 7, 6, [25, 17, 9, 12], [6, 6, 0, 4], targets=tg, weights=True)
    tg2 = torch.tensor([[2, 2, 2, 2], [1, 2, 3, 4], [4, 3, 2, 1]])
    run("infeasible", 12, 3, 6, 4, [5, 12, 3], [4, 4, 4], targets=tg2)
    run("infeasible_zi", 12, 3, 6, 4, [5, 12, 3], [4, 4, 4], targets=tg2, zero_inf=True)
    run("c42", 60, 5, 42, 30, [60, 47, 33, 60, 31], [30, 12, 9, 1, 15], weights=True)
    run("c45_wide", 40, 3, 45, 10, [40, 40, 25], [10, 4, 7], scale=12.0)
    np.savez_compressed(os.path.join(GOLD, "ctc_logits_small.npz"), **cases)


# ----------------------------------------------------------------------------------------------- greedy
def gen_greedy(vocab43):
    D = RL.decoder_module()
    rng = np.random.default_rng(7)
    out, strings = {}, {}
    vocab45 = {0: "blank", 1: "UNK", 2: "sil"}
    for w in ["sh", "iy", "hh", "ae", "d", "y", "er0", "aa", "r", "k", "s", "uw", "t", "ih", "n", "g", "w", "ao", "dh",
              "l", "ow", "m", "eh", "oy", "ay", "b", "er", "v", "f", "z", "th", "ah", "ah0", "p", "ey", "ng", "ch", "uh",
              "zh", "jh", "aw", "err"]:
        vocab45[len(vocab45)] = w           # utils/tools.py:58-104
    dup_vocab = {0: "blank", 1: "a", 2: "b", 3: "a", 4: "c"}     # two indices sharing one word string
    for i, (voc, T, B) in enumerate([(vocab43, 50, 6), (vocab45, 80, 5), (vocab43, 33, 7), (dup_vocab, 40, 8),
                                      (vocab43, 1, 3), (vocab45, 64, 4)]):
        C = len(voc)
        x = rng.standard_normal((T, B, C)).astype(np.float32)
        if i % 2 == 0:          # heavy exact ties + long runs
            x = np.round(x * 1.5).astype(np.float32)
        if i == 3:
            x[:, :, 0] -= 1.0
        run = rng.integers(0, C, size=(T, B))
        for t in range(1, T):
            keep = rng.random(B) < 0.6
            run[t, keep] = run[t - 1, keep]
        x[np.arange(T)[:, None], np.arange(B)[None, :], run] += 3.0
        lp = torch.log_softmax(torch.from_numpy(x), -1)
        lens = rng.integers(0, T + 1, size=B).tolist()
        lens[0] = T
        dec = D.GreedyDecoder(voc, space_idx=-1, blank_index=0)
        res = dec.decode(lp, lens)
        out["%d/lp" % i] = lp.numpy()
        out["%d/lens" % i] = np.array(lens, dtype=np.int64)
        strings[str(i)] = {"vocab": {str(k): v for k, v in voc.items()}, "out": res}
    np.savez_compressed(os.path.join(GOLD, "greedy.npz"), **out)
    jdump("greedy.json.gz", strings)


# ----------------------------------------------------------------------------------------------- beam
def _ref_beam(D, voc, probs_btc, lens, W, alpha):
    dec = D.BeamDecoder(voc, beam_width=W, blank_index=0, space_idx=-1, lm_path=RL.arpa_path(), lm_alpha=alpha)
    res = []
    for b in range(probs_btc.shape[0]):             # utterance by utterance so each exception is recorded
        p = torch.from_numpy(probs_btc[b:b + 1])
        try:
            res.append(dec._decoder.decode(p, [lens[b]])[0])
        except (IndexError, ValueError, KeyError) as e:
            res.append({"exc": type(e).__name__})
    return res


def gen_beam_small(vocab43):
    D = RL.decoder_module()
    rng = np.random.default_rng(11)
    names = [vocab43[i] for i in range(2, 43)]
    arrs, meta = {}, []
    for i in range(160):
        C = int(rng.integers(4, 20)); T = int(rng.integers(3, 40)); W = int(rng.integers(1, 11))
        alpha = [0.0, 0.01, 0.3][i % 3]
        pick = rng.choice(len(names), size=C - 1, replace=False)
        voc = {0: "blank"}
        for j, k in enumerate(pick):
            voc[j + 1] = names[int(k)]
        if i % 16 == 5:
            voc[1] = "UNK"
        kind = i % 5
        x = rng.standard_normal((1, T, C))
        if kind == 1:       # peaky
            for t in range(T):
                if rng.random() < 0.7:
                    x[0, t, 0] += 8
                else:
                    x[0, t, int(rng.integers(1, C))] += 6
        elif kind == 2:     # quantised ties
            x = np.round(x)
        elif kind == 3:     # uniform
            x[:] = 0
        elif kind == 4:     # mid-blank frames around the 0.9 / 0.1 thresholds
            for t in range(T):
                x[0, t, 0] += rng.choice([0.0, 2.0, 4.0, 6.0])
        lp = torch.log_softmax(torch.from_numpy(x).float(), -1)
        probs = torch.exp(lp).numpy()           # ctcDecoder.py:223-225 (transpose is a no-op for B=1 here)
        if i % 40 == 17:
            probs[0, T // 2, int(rng.integers(0, C))] = 0.0        # math.log(0) -> ValueError (if that frame is processed)
        L = T if i % 4 else int(rng.integers(1, T + 1))
        if i == 150:
            voc[C - 1] = "ah0"     # not in the shipped ARPA -> KeyError
        res = _ref_beam(D, voc, probs, [L], W, alpha)[0]
        arrs["%d" % i] = probs[0]
        meta.append({"C": C, "T": T, "W": W, "alpha": alpha, "len": L, "vocab": [voc[k] for k in range(C)], "out": res})
    np.savez_compressed(os.path.join(GOLD, "beam_small.npz"), **arrs)
    jdump("beam_small.json.gz", meta)


def peaky_probs(rng, B, T, C):
    """SURVEY.md 8(d) C4 generator: w.p. 0.7 add +8 to blank, else +6 to a random non-blank; exp(log_softmax) float32."""
    x = rng.standard_normal((B, T, C)).astype(np.float32)
    u = rng.random((B, T))
    k = rng.integers(1, C, size=(B, T))
    bi, ti = np.nonzero(u < 0.7)
    x[bi, ti, 0] += 8
    bi, ti = np.nonzero(u >= 0.7)
    x[bi, ti, k[bi, ti]] += 6
    lp = torch.log_softmax(torch.from_numpy(x), -1)
    return torch.exp(lp).numpy()


def gen_beam_c4(vocab43):
    D = RL.decoder_module()
    rng = np.random.default_rng(23)
    probs = peaky_probs(rng, 6, 500, 43)
    lens = [500, 500, 431, 500, 77, 500]
    meta = {"lens": lens, "vocab": [vocab43[k] for k in range(43)]}
    for alpha in (0.0, 0.3):
        meta["out_alpha_%g" % alpha] = _ref_beam(D, vocab43, probs, lens, 10, alpha)
    np.savez_compressed(os.path.join(GOLD, "beam_c4.npz"), probs=probs)
    jdump("beam_c4.json.gz", meta)


def gen_lm_table(vocab43):
    D = RL.decoder_module()
    import utils.NgramLM as NG  # noqa: imported by decoder_module
    lm = NG.LanguageModel(arpa_file=RL.arpa_path())
    C = 43
    tab = np.full((C + 1, C + 1), np.nan)
    names = [vocab43[i] for i in range(C)] + [""]
    for i, a in enumerate(names):
        for j, b in enumerate(names):
            try:
                tab[i, j] = lm.get_bi_prob(a, b)
            except KeyError:
                pass
    np.savez_compressed(os.path.join(GOLD, "lm_table.npz"), table=tab)


# ----------------------------------------------------------------------------------------------- Levenshtein corpora
def _pairs_from_log(path, a="decoded:", b="origin :"):
    dec, org = [], []
    for line in open(path, errors="replace"):
        if line.startswith(a):
            dec.append(line[len(a):].strip())
        elif line.startswith(b):
            org.append(line[len(b):].strip())
    return dec, org


def _pairs_from_files(hyp, ref):
    def rd(p):
        d = {}
        for line in open(p):
            parts = line.strip().split(" ", 1)
            d[parts[0]] = parts[1] if len(parts) > 1 else ""
        return d
    h, r = rd(hyp), rd(ref)
    keys = [k for k in h if k in r]
    return [h[k] for k in keys], [r[k] for k in keys]


def gen_wer_corpus():
    D = RL.decoder_module()
    dec = D.Decoder({0: "blank"}, space_idx=-1, blank_index=0)
    A, Cn = RL.eg_dir("attention_aug"), RL.eg_dir("cnn-rnn-ctc")
    out = {}

    def last_block(d, o, n=900):
        return d[-n:], o[-n:]

    d, o = _pairs_from_log(os.path.join(A, "result", "train.log"))
    corpora = {"attention_aug_trainlog": last_block(d, o) + (4976,)}
    d, o = _pairs_from_log(os.path.join(Cn, "nohup.out"))
    corpora["cnn_rnn_nohup"] = last_block(d, o) + (7923,)
    corpora["attention_aug_hyp_human"] = _pairs_from_files(os.path.join(A, "result", "hyp"),
                                                           os.path.join(A, "result", "human_seq")) + (4423,)
    corpora["cnn_rnn_hyp_human"] = _pairs_from_files(os.path.join(Cn, "result", "hyp"),
                                                     os.path.join(Cn, "result", "human_seq")) + (7934,)
    for name, (hyp, ref, total) in corpora.items():
        dists, paths = [], []
        for h, r in zip(hyp, ref):
            dd, pp = dec.wer(h, r)
            dists.append(dd)
            paths.append("".join(pp))
        assert sum(dists) == total, (name, sum(dists), total)
        ntok = sum(len(r.split()) for r in ref)
        out[name] = {"hyp": hyp, "ref": ref, "total": total, "ref_tokens": ntok, "dist": dists,
                     "paths": paths if name == "attention_aug_hyp_human" else None}
        print(name, sum(dists), ntok)
    jdump("wer_corpus.json.gz", out)


# ----------------------------------------------------------------------------------------------- MDD post-processing
def ref_mdd(D, F, decoded, canonical):
    """The body of the per-utterance loop of infer.py:310-342 on one (decoded, canonical) string pair."""
    dec = D.Decoder({0: "blank"}, space_idx=-1, blank_index=0)
    hyp = [i for i in decoded.split(" ") if i != "sil"]
    ref = [i for i in canonical.split(" ") if i != "sil"]
    dn, cn = " ".join(hyp), " ".join(ref)
    dn = dn.replace("err", "").replace("  ", " ")
    try:
        _, path = dec.wer(dn, cn)
    except TypeError:
        return {"exc": "TypeError"}
    pd = [c for c in dn.split(" ") if c]
    pc = [c for c in cn.split(" ") if c]
    path0 = list(path)
    pd, pc, path = F["align_canonical_decoded"](pd, pc, path)
    s1, s2, s3 = F["print_aligned_string"](pd, pc, path)
    ds = sum(1 if c in "DS" else 0 for c in path)
    cc = sum(1 if c == "-" else 0 for c in path)
    ins, sub, dele = F["stastics"](path, pc, pd)
    tmp = min(len(ins) / 4, 0.1 * (cc + ds))
    score = math.ceil((1 - (ds + tmp) / (ds + cc)) * 100)
    return {"path0": path0, "dec": pd, "can": pc, "path": path, "ins": ins, "sub": sub, "del": dele,
            "correct": cc, "del_sub": ds, "score": score, "printed": [s1, s2, s3]}


def noisy_pair(rng, phones, L):
    can = [phones[int(k)] for k in rng.integers(0, len(phones), size=L)]
    dec = []
    for p in can:
        u = rng.random()
        if u < 0.08:
            continue
        if u < 0.18:
            dec.append(phones[int(rng.integers(0, len(phones)))])
        else:
            dec.append(p)
        if rng.random() < 0.07:
            dec.append(phones[int(rng.integers(0, len(phones)))])
    lead = int(rng.integers(0, 4)) if rng.random() < 0.4 else 0
    head = [phones[int(k)] for k in rng.integers(0, len(phones), size=lead)]
    if lead and rng.random() < 0.5 and dec:
        head[-1] = dec[0]                       # duplicated leading phone (infer.py:203-207)
    if rng.random() < 0.15:
        dec.insert(int(rng.integers(0, len(dec) + 1)), "sil")
    if rng.random() < 0.1:
        dec.insert(int(rng.integers(0, len(dec) + 1)), "err")
    return " " + " ".join(head + dec) if (head or dec) else "", " ".join(can)


def gen_mdd_post(vocab43):
    D, F = RL.decoder_module(), RL.infer_functions()
    rng = np.random.default_rng(5)
    phones = [vocab43[i] for i in range(3, 42)]
    cases = []
    kats = [(" t t v ow k b y uw l eh r iy iy", "v ow k ae b y ah l eh r iy"), (" s s g ae dh er", "g ae dh er"),
            (" g g ae dh er", "g ae dh er"), (" ah b aw", "ah b aw t"), (" k ah ah b aw t s", "ah b aw t"),
            ("", "ah b"), (" sil", "sil ah")]
    for d, c in kats:
        cases.append({"decoded": d, "canonical": c, "out": ref_mdd(D, F, d, c)})
    for i in range(600):
        L = int(rng.integers(1, 14)) if i % 3 else int(rng.integers(14, 70))
        d, c = noisy_pair(rng, phones, L)
        cases.append({"decoded": d, "canonical": c, "out": ref_mdd(D, F, d, c)})
    jdump("mdd_post.json.gz", cases)


# ----------------------------------------------------------------------------------------------- config 1
def gen_c1(vocab43):
    D, F = RL.decoder_module(), RL.infer_functions()
    root = os.path.join(RL.REF_ROOT, "egs", "vocabulary", "single")
    cmu = {}
    for line in open(os.path.join(RL.eg_dir(), "dict", "cmudict.dict")):
        parts = line.split("#")[0].split()
        if parts and "(" not in parts[0]:
            cmu.setdefault(parts[0], parts[1:])
    vocab45 = {0: "blank", 1: "UNK", 2: "sil"}
    for w in ["sh", "iy", "hh", "ae", "d", "y", "er0", "aa", "r", "k", "s", "uw", "t", "ih", "n", "g", "w", "ao", "dh",
              "l", "ow", "m", "eh", "oy", "ay", "b", "er", "v", "f", "z", "th", "ah", "ah0", "p", "ey", "ng", "ch", "uh",
              "zh", "jh", "aw", "err"]:
        vocab45[len(vocab45)] = w
    w2i = {v: k for k, v in vocab45.items()}
    rng = np.random.default_rng(101)
    words, canon = [], []
    for i in range(1, 21):
        word = open(os.path.join(root, "%d.txt" % i)).read().strip()
        ph = cmu.get(word.lower())
        if ph is None:
            continue                                # OPPO, longtimenosee are OOV in the shipped dictionary
        parts_ = [p.rstrip(string.digits) if p not in ["ER0", "AH0"] else p for p in ph]     # infer.py:549
        words.append(word)
        canon.append([p.lower() for p in parts_])
    B = len(words)
    T = 112
    lens = rng.integers(60, 111, size=B)
    lp = np.zeros((T, B, 45), dtype=np.float32)
    for b in range(B):
        # recognised = canonical with some edits, rendered as peaky posteriors with 'sil' at both ends
        seq = ["sil"]
        for p in canon[b]:
            u = rng.random()
            if u < 0.1:
                continue
            seq.append(p if u > 0.25 else vocab45[int(rng.integers(3, 44))])
            if rng.random() < 0.08:
                seq.append(vocab45[int(rng.integers(3, 44))])
        seq.append("sil")
        frames = []
        for p in seq:
            frames += [w2i[p]] * int(rng.integers(1, 4)) + [0] * int(rng.integers(0, 3))
        frames = (frames + [0] * T)[:T]
        x = rng.standard_normal((T, 45)).astype(np.float32)
        x[np.arange(T), frames] += 7
        lp[:, b] = torch.log_softmax(torch.from_numpy(x), -1).numpy()
    dec = D.GreedyDecoder(vocab45, space_idx=-1, blank_index=0)
    decoded = dec.decode(torch.from_numpy(lp), lens.tolist())
    outs = [ref_mdd(D, F, decoded[b], "sil " + " ".join(canon[b]) + " sil") for b in range(B)]
    np.savez_compressed(os.path.join(GOLD, "c1_words.npz"), lp=lp, lens=lens.astype(np.int64))
    jdump("c1_words.json.gz", {"words": words, "canonical": [["sil"] + c + ["sil"] for c in canon],
                               "vocab": [vocab45[k] for k in range(45)], "decoded": decoded, "mdd": outs})


# ----------------------------------------------------------------------------------------------- corpus MDD evaluation
def gen_corpus(vocab43):
    """test_ctc_nosil.py:218-297 executed per utterance (and once over the whole list) on synthetic
    (canonical, human label, recognised) triples: canonical -> label by L2-style edits, label -> recognised by
    recogniser noise, with runs of insertions at one position (the remove-while-iterating branch, :276-287)."""
    D = RL.decoder_module()
    pr, run = RL.corpus_eval_functions()
    rng = np.random.default_rng(11)
    phones = [vocab43[i] for i in range(3, 42)]

    def edit(seq, p_del, p_sub, p_ins, burst):
        out = []
        for w in seq:
            u = rng.random()
            if u < p_del:
                pass
            elif u < p_del + p_sub:
                out.append(phones[int(rng.integers(0, len(phones)))])
            else:
                out.append(w)
            if rng.random() < p_ins:
                for _ in range(1 + (int(rng.integers(0, burst)) if burst else 0)):
                    out.append(phones[int(rng.integers(0, len(phones)))])
        return out

    utts, labs, cans, decs = [], [], [], []
    for n in range(500):
        L = int(rng.integers(1, 14)) if n % 5 else int(rng.integers(14, 60))
        can = [phones[int(k)] for k in rng.integers(0, len(phones), size=L)]
        lab = edit(can, 0.06, 0.10, 0.06, 3) or [can[0]]
        if n % 7 == 0:
            lab = [phones[int(k)] for k in rng.integers(0, len(phones), size=int(rng.integers(0, 3)))] + lab
        dec = edit(lab, 0.07, 0.12, 0.08, 3) or [lab[0]]
        if n % 11 == 0:
            dec = [phones[int(k)] for k in rng.integers(0, len(phones), size=int(rng.integers(1, 4)))] + dec
        utts.append(("TXHC_%04d" if n % 3 == 0 else "NJS_%04d") % n)
        labs.append(" ".join(lab)); cans.append(" ".join(can)); decs.append(" ".join(dec))
    dec0 = D.Decoder({0: "blank"}, space_idx=-1, blank_index=0)
    per = []
    for u, l, c, d in zip(utts, labs, cans, decs):
        r = run(dec0, [u], [l], [c], [d])
        _, lc = dec0.wer(l, c)
        _, dc = dec0.wer(d, c)
        d1 = pr(l, c, lc)[3]
        d2 = pr(d, c, dc)[3]
        per.append({"counts": r, "lc": "".join(lc), "dc": "".join(dc),
                    "d1": [[str(k), v] for k, v in d1.items()], "d2": [[str(k), v] for k, v in d2.items()]})
    total = run(dec0, utts, labs, cans, decs)
    jdump("corpus_mdd.json.gz", {"utt": utts, "labels": labs, "canonicals": cans, "decoded": decs, "per_utt": per,
                                 "total": total, "m_speaker": ["TXHC"]})
    print("corpus_mdd", total)


def gen_textio(vocab):
    """collate lengths (data_loader.py:151-181 + train_ctc.py:68), nosil strings and log lines
    (test_ctc_nosil.py:196-214), CMU -> IPA table (dict/phonetic_dict.py:21-64), all from the reference's own code."""
    create_input = RL.collate_function()
    rng = np.random.default_rng(5)
    batches = []
    for _ in range(60):
        B = int(rng.integers(1, 17))
        flen = rng.integers(8, 1600, size=B).tolist()
        llen = rng.integers(1, 40, size=B).tolist()
        batch = [(torch.zeros(f, 3), torch.ones(l, dtype=torch.long), torch.ones(max(1, l // 2), dtype=torch.long), "u%d" % i)
                 for i, (f, l) in enumerate(zip(flen, llen))]
        _, input_sizes, _, target_sizes, _, _, _ = create_input(batch)
        mx = max(flen)
        outs = {}
        for out_len in sorted({mx, mx // 4, (mx + 3) // 4, mx // 2, 1 + mx // 8}):
            if out_len > 0:
                outs[str(out_len)] = (input_sizes * out_len).long().tolist()          # train_ctc.py:68
        batches.append({"feature_lengths": flen, "label_lengths": llen, "input_sizes_f32": [float(v) for v in input_sizes],
                        "target_sizes": target_sizes.tolist(), "frame_lengths": outs})
    run = RL.nosil_log_loops()
    units = [vocab[i] for i in sorted(vocab) if i >= 2]
    logs = []
    for k in range(12):
        n = int(rng.integers(1, 6))
        def seq():
            toks = [("sil" if rng.random() < 0.25 else units[int(rng.integers(len(units)))]) for _ in range(int(rng.integers(0, 12)))]
            return " ".join(toks)
        decoded = [(" " + seq()) if rng.random() < 0.5 else seq() for _ in range(n)]     # greedy strings lead with a blank
        labels, canon = [seq() for _ in range(n)], [seq() for _ in range(n)]
        utts = ["spk%d_utt%03d" % (k, i) for i in range(n)]
        d, l, c, t1, t2, t3 = run(decoded, labels, canon, utts)
        logs.append({"decoded": decoded, "labels": labels, "canonicals": canon, "utt_list": utts, "decoded_nosil": d,
                     "labels_nosil": l, "canonicals_nosil": c, "decode_seq": t1, "human_seq": t2, "transcribe": t3})
    table = RL.cmu_ipa_table()
    probe = sorted(set(list(table) + [k.lower() for k in table] + units + ["sil", "err", "UNK", "xx"]))
    ipa = {c: table.get(c.upper(), c) for c in probe}                                     # infer.py:331-333
    jdump("textio.json.gz", {"batches": batches, "logs": logs, "ipa": ipa})


def gen_scoring():
    """The offline half of the corpus evaluation: the rates (recall / precision / f1, TA / FR / FA, Correct / Error
    Diag, FAR / FRR / DER) that result/utils/scoring/ins_del_sub_cor_analysis.py:193-219 prints from its eleven
    counters.  The counting loop above those lines reads Kaldi align-text output, which this image cannot produce;
    lines 193-219 are executed verbatim on injected counters and their printed text is recorded."""
    import contextlib
    import io
    path = os.path.join(RL.eg_dir(), "result", "utils", "scoring", "ins_del_sub_cor_analysis.py")
    src = "".join(open(path).readlines()[192:219])
    assert src.startswith("sum1 = cor_cor"), src[:40]
    names = ["cor_cor", "cor_nocor", "sub_sub", "sub_sub1", "sub_nosub", "ins_ins", "ins_ins1", "ins_noins", "del_del",
             "del_del1", "del_nodel"]
    rng = np.random.default_rng(17)
    cases = []
    for k in range(12):
        c = {n: int(rng.integers(1, 4000 if k % 3 else 40)) for n in names}
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            exec(compile(src, path, "exec"), dict(c))
        cases.append({"counters": c, "text": buf.getvalue()})
    jdump("scoring_rates.json.gz", {"names": names, "cases": cases})


def main():
    assert RL.available(), "reference tree not mounted"
    os.makedirs(os.path.join(GOLD, "assets"), exist_ok=True)
    shutil.copy(os.path.join(RL.eg_dir(), "data", "units"), os.path.join(GOLD, "assets", "units"))
    shutil.copy(RL.arpa_path(), os.path.join(GOLD, "assets", "lm_phone_bg.arpa"))
    vocab43 = RL.units_vocab()
    assert len(vocab43) == 43, len(vocab43)
    steps = sys.argv[1:] or ["ctc", "ctc_logits", "greedy", "beam_small", "beam_c4", "lm", "wer", "mdd", "c1", "corpus", "textio", "scoring"]
    for s in steps:
        print("==", s, flush=True)
        {"ctc": gen_ctc, "ctc_logits": gen_ctc_logits, "greedy": lambda: gen_greedy(vocab43), "beam_small": lambda: gen_beam_small(vocab43),
         "beam_c4": lambda: gen_beam_c4(vocab43), "lm": lambda: gen_lm_table(vocab43), "wer": gen_wer_corpus,
         "mdd": lambda: gen_mdd_post(vocab43), "c1": lambda: gen_c1(vocab43),
         "corpus": lambda: gen_corpus(vocab43), "textio": lambda: gen_textio(vocab43), "scoring": gen_scoring}[s]()


if __name__ == "__main__":
    main()
