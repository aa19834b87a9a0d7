"""Pins the CPU oracle (oracle/mdd_oracle.c) against outputs of the REAL reference recorded in tests/golden/
(written by oracle/gen_golden.py in the build container) and against torch CPU ctc_loss.  CPU-only."""
import numpy as np
import pytest
import torch

from conftest import jload, npz
from oracle import api as O


CTC_CASES = ["basic", "repeat_ragged", "one_d", "infeasible", "infeasible_zi", "c42", "tiny"]


@pytest.mark.parametrize("name", CTC_CASES)
def test_ctc_oracle_vs_torch_golden(name):
    z = npz("ctc_small.npz")
    g = lambda k: z[name + "/" + k]
    zi = bool(g("zero_inf"))
    r = O.ctc_loss(g("lp"), g("targets"), g("in_len"), g("tgt_len"), zero_infinity=zi, want_alpha=True)
    np.testing.assert_allclose(r["nll"], g("nll_f64"), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(r["grad"], g("grad_f64"), rtol=0, atol=1e-12, equal_nan=True)
    a, ga = r["alpha"], g("alpha_f64")
    S = a.shape[2]
    assert np.array_equal(np.isneginf(a), np.isneginf(ga[:, :, :S]))
    fin = np.isfinite(a)
    np.testing.assert_allclose(a[fin], ga[:, :, :S][fin], rtol=0, atol=1e-11)
    r32 = O.ctc_loss(g("lp"), g("targets"), g("in_len"), g("tgt_len"), zero_infinity=zi, dtype=np.float32)
    np.testing.assert_allclose(r32["nll"], g("nll_f32"), rtol=1e-5)
    np.testing.assert_allclose(r32["grad"], g("grad_f32"), rtol=0, atol=2e-5, equal_nan=True)   # fp32 vs fp32 rounding


LOGIT_CASES = ["basic", "repeat_ragged", "infeasible", "infeasible_zi", "c42", "c45_wide"]


@pytest.mark.parametrize("name", LOGIT_CASES)
def test_ctc_logits_oracle_vs_torch_golden(name):
    """log_softmax -> ctc_loss -> autograd to the logits, as torch CPU computed it (gen_ctc_logits)."""
    z = npz("ctc_logits_small.npz")
    g = lambda k: z[name + "/" + k]
    r = O.ctc_loss_logits(g("logits"), g("targets"), g("in_len"), g("tgt_len"), zero_infinity=bool(g("zero_inf")),
                          grad_out=g("grad_out"))
    np.testing.assert_allclose(r["nll"], g("nll_f64"), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(r["grad"], g("grad_f64"), rtol=0, atol=1e-12, equal_nan=True)
    np.testing.assert_allclose(r["row_lse"], g("lse_f64"), rtol=1e-13)
    assert np.array_equal(r["argmax"], g("argmax"))
    # the logit gradient of a finite utterance is (softmax - occupancy) * g: rows sum to zero, rows t >= T_b are zero
    gr, il = r["grad"], g("in_len")
    for b in range(gr.shape[1]):
        if np.isfinite(gr[:, b]).all():
            assert np.abs(gr[:, b].sum(-1)).max() < 1e-12
            assert not gr[il[b]:, b].any()


def test_ctc_oracle_vs_torch_live():
    g = torch.Generator().manual_seed(3)
    T, B, C, L = 40, 6, 11, 8
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1).double().requires_grad_(True)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.tensor([40, 40, 31, 22, 40, 17])
    tl = torch.tensor([8, 3, 8, 5, 0, 8])
    go = torch.rand(B, generator=g).double()
    nll = torch.nn.functional.ctc_loss(lp, tg, il, tl, reduction="none")
    (nll * go).sum().backward()
    r = O.ctc_loss(lp.detach().numpy(), tg.numpy(), il.numpy(), tl.numpy(), grad_out=go.numpy())
    np.testing.assert_allclose(r["nll"], nll.detach().numpy(), rtol=1e-12)
    np.testing.assert_allclose(r["grad"], lp.grad.numpy(), atol=1e-12)


def test_greedy_oracle_vs_reference():
    z, js = npz("greedy.npz"), jload("greedy.json.gz")
    for i, case in js.items():
        voc = {int(k): v for k, v in case["vocab"].items()}
        got = O.greedy_decode(z[i + "/lp"], z[i + "/lens"], voc)
        assert got == case["out"], i


def test_beam_oracle_vs_reference_small(arpa):
    z, meta = npz("beam_small.npz"), jload("beam_small.json.gz")
    n_exc = 0
    for i, m in enumerate(meta):
        voc = {k: w for k, w in enumerate(m["vocab"])}
        lm = O.lm_table(arpa, voc, m["C"])
        probs = z[str(i)][None]
        toks, olen, st, _ = O.beam_tokens(probs, [m["len"]], m["W"], 0, lm, m["alpha"])
        if isinstance(m["out"], dict):
            n_exc += 1
            assert O.BEAM_EXC[int(st[0])].__name__ == m["out"]["exc"], (i, st, m["out"])
        else:
            assert st[0] == 0, (i, st, m["out"])
            assert " ".join(voc[int(k)] for k in toks[0, :olen[0]]) == m["out"], i
    assert n_exc >= 5      # the fixture exercises the reference's crash modes


def test_beam_oracle_vs_reference_c4(arpa, vocab43):
    z, meta = npz("beam_c4.npz"), jload("beam_c4.json.gz")
    lm = O.lm_table(arpa, vocab43, 43)
    for alpha in (0.0, 0.3):
        want = meta["out_alpha_%g" % alpha]
        toks, olen, st, _ = O.beam_tokens(z["probs"], meta["lens"], 10, 0, lm, alpha)
        for b, w in enumerate(want):
            if isinstance(w, dict):
                assert O.BEAM_EXC[int(st[b])].__name__ == w["exc"]
            else:
                assert st[b] == 0 and " ".join(vocab43[int(k)] for k in toks[b, :olen[b]]) == w, (alpha, b)


def test_lm_table_vs_reference(arpa, vocab43):
    want = npz("lm_table.npz")["table"]
    got = O.lm_table(arpa, vocab43, 43)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])       # bit-exact fp64


def test_wer_corpus_totals():
    """4976 / 7923 / 4423 / 7934 over 29087 reference phones; per-pair paths for one corpus."""
    data = jload("wer_corpus.json.gz")
    for name, c in data.items():
        words = {}
        hyp = [[words.setdefault(w, len(words)) for w in h.split()] for h in c["hyp"]]
        ref = [[words.setdefault(w, len(words)) for w in r.split()] for r in c["ref"]]
        dist, paths, st = O.edit_align(hyp, ref)
        assert not st.any()
        assert int(dist.sum()) == c["total"], name
        assert c["ref_tokens"] == 29087
        assert dist.tolist() == c["dist"]
        if c["paths"]:
            assert ["".join(p) for p in paths] == c["paths"]


def _mdd_from_strings(decoded, canonical):
    hyp = [w for w in decoded.split(" ") if w != "sil"]
    ref = [w for w in canonical.split(" ") if w != "sil"]
    dn = " ".join(hyp).replace("err", "").replace("  ", " ")
    cn = " ".join(ref)
    words = {}
    d = [words.setdefault(w, len(words)) for w in dn.split()]
    c = [words.setdefault(w, len(words)) for w in cn.split()]
    inv = {v: k for k, v in words.items()}
    inv[O.PAD_D], inv[O.PAD_I] = "D", "I"
    dist, paths, st = O.edit_align([d], [c])
    if st[0]:
        return {"exc": "TypeError"}
    r = O.mdd_post(d, c, paths[0])
    return {"path0": paths[0], "dec": [inv[k] for k in r["dec"]], "can": [inv[k] for k in r["can"]], "path": r["ops"],
            "ins": [inv[k] for k in r["ins"]], "sub": [inv[k] for k in r["sub"]], "del": [inv[k] for k in r["del"]],
            "correct": r["correct"], "del_sub": r["del_sub"],
            "score": O.mdd_score(len(r["ins"]), r["correct"], r["del_sub"])}


def test_mdd_post_oracle_vs_reference():
    cases = jload("mdd_post.json.gz")
    n_exc = 0
    for i, c in enumerate(cases):
        got = _mdd_from_strings(c["decoded"], c["canonical"])
        want = dict(c["out"])
        want.pop("printed", None)
        n_exc += "exc" in want
        assert got == want, (i, c["decoded"], c["canonical"])
    assert n_exc >= 1


def test_prefix_score_invariants():
    """psi(l.eos) == -nll (torch fp64) and P(g...) = P(g ends) + sum_c P(g.c ...)."""
    g = torch.Generator().manual_seed(9)
    T, C = 30, 6
    x = torch.log_softmax(torch.randn(T, C, generator=g).double(), -1)
    for labels in ([1, 2, 3], [2, 2, 4], [5], [1, 1, 1, 2, 2]):
        r = O.prefix_init(x.numpy())
        last, psi_g = -1, 0.0
        for n, c in enumerate(labels):
            psi, rr, eos = O.prefix_score(x.numpy(), r, last, n)
            if n > 0:
                tot = np.logaddexp.reduce(np.append(psi[1:], eos))
                assert abs(tot - psi_g) < 1e-10
            psi_g, r, last = psi[c], rr[c], c
        _, _, eos = O.prefix_score(x.numpy(), r, last, len(labels))
        nll = torch.nn.functional.ctc_loss(x[:, None, :], torch.tensor([labels]), torch.tensor([T]),
                                           torch.tensor([len(labels)]), reduction="none")
        assert abs(eos + float(nll)) < 1e-10


def _corpus_rows(data):
    words = {}
    enc = lambda s: [words.setdefault(w, len(words)) for w in s.split()]
    return [enc(s) for s in data["labels"]], [enc(s) for s in data["canonicals"]], [enc(s) for s in data["decoded"]]


def test_corpus_counts_oracle_vs_reference():
    """test_ctc_nosil.py:218-297 executed by oracle/gen_golden.py on 500 triples: per-utterance counters (both the
    general and the m_speaker rule), paths, PER numerators, and the corpus totals."""
    data = jload("corpus_mdd.json.gz")
    lab, can, dec = _corpus_rows(data)
    _, lc, st1 = O.edit_align(lab, can)
    dl, _, st2 = O.edit_align(dec, lab)
    _, dc, st3 = O.edit_align(dec, can)
    assert not (st1.any() or st2.any() or st3.any())
    keys = ["true_accept", "false_rejection", "false_accept", "true_rejection_correct_diagnose",
            "true_rejection_wrong_diagnose", "total_phonemes_in_canonical"]
    tot = {k: 0 for k in keys + ["total_wer", "num_word"]}
    quirk = 0
    for b, ref in enumerate(data["per_utt"]):
        assert "".join(lc[b]) == ref["lc"] and "".join(dc[b]) == ref["dc"]
        c = O.corpus_counts(lab[b], lc[b], dec[b], dc[b])
        want = ref["counts"]
        assert c[:6] == [want[k] for k in keys], (b, c, want)
        assert int(dl[b]) == want["total_wer"] and len(lab[b]) == want["num_word"]
        if data["utt"][b][:4] in data["m_speaker"]:
            assert [c[0], c[6], c[7], c[3], c[4], c[5]] == [want["m_" + k] for k in keys], (b, c, want)
            quirk += c[6] != c[1] or c[7] != c[2]
        for k, v in zip(keys, c[:6]):
            tot[k] += v
        tot["total_wer"] += int(dl[b])
        tot["num_word"] += len(lab[b])
    assert quirk > 0                                                    # the m_speaker '+= 1' branch is exercised
    for k, v in tot.items():
        assert v == data["total"][k], k


def test_corpus_insertion_keys_unique():
    """print_align_space_canonical_origin keys an insertion by str(j-1)+str(j) (:61); the oracle stores j."""
    keys = [str(j - 1) + str(j) for j in range(0, 5000)]
    assert len(set(keys)) == len(keys)


def test_joint_beam_step_c_vs_python_restatement():
    """The C restatement of the joint CTC/attention beam step (orc_joint_step; SURVEY Appendix A.6, reference-unpinned)
    against an independent pure-Python loop over O.prefix_score: same tokens, scores, flags, with and without eos,
    including hypotheses that die because the utterance is shorter than the search."""
    rng = np.random.default_rng(0)
    T, B, C, W, steps = 14, 3, 6, 3, 5
    x = np.log(rng.dirichlet(np.ones(C), (T, B)))
    lens = np.array([14, 9, 4])
    tab = np.log(rng.dirichlet(np.ones(C), (4, C + 1))).astype(np.float32)

    def att(step, st):
        return tab[step % 4][st["last"] + 1]

    for eos in (None, 5):
        tok, sc, gl, fl = O.joint_beam_search(x, lens, att, beam=W, ctc_weight=0.3, max_len=steps, eos=eos)
        for b in range(B):
            xb = x[:lens[b], b]
            hyps = [dict(r=O.prefix_init(xb), last=-1, glen=0, psi=0.0, score=0.0, flag=0, tok=[])]
            for step in range(steps):
                cands = []
                for w, h in enumerate(hyps):
                    if h["flag"] == 1:
                        cands.append((h["score"], w, eos, h, None))
                        continue
                    psi, r, pe = O.prefix_score(xb, h["r"], h["last"], h["glen"])
                    a = tab[step % 4][h["last"] + 1]
                    for c in range(1, C):
                        pc = pe if c == eos else psi[c]
                        if np.isfinite(pc):
                            cands.append((h["score"] + 0.3 * (pc - h["psi"]) + 0.7 * float(a[c]), w, c, h, (r[c], pc)))
                cands.sort(key=lambda z: (-z[0], z[1] * C + z[2]))
                new = []
                for v, w, c, h, ex in cands[:W]:
                    if h["flag"] == 1:
                        new.append(h)
                    else:
                        new.append(dict(r=ex[0], last=c, glen=h["glen"] + 1, psi=ex[1], score=v,
                                        flag=1 if c == eos else 0, tok=h["tok"] + [c]))
                hyps = new
            for k, h in enumerate(hyps):
                assert list(tok[b, k, :len(h["tok"])]) == h["tok"]
                assert abs(sc[b, k] - h["score"]) < 1e-9 and fl[b, k] == h["flag"] and gl[b, k] == len(h["tok"])
            assert all(fl[b, k] == 2 for k in range(len(hyps), W))
