"""ctypes front-end of the C oracle (oracle/mdd_oracle.c).  TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.

Function names mirror the reference symbols they restate:

* ``ctc_loss``            nn.CTCLoss call site, egs/attention_aug/steps/train_ctc.py:186,72,96
* ``greedy_decode``       GreedyDecoder.decode, egs/attention_aug/utils/ctcDecoder.py:186-200
* ``edit_align``/``wer``  Decoder._edit_distance/printChanges/wer, utils/ctcDecoder.py:118-184
* ``compute_wer``         CTC_Model.compute_wer, egs/attention_aug/models/model_ctc.py:227-244
* ``beam_search``         ctcBeamSearch.decode, egs/attention_aug/utils/BeamSearch.py:73-153
* ``lm_table``            LanguageModel, egs/attention_aug/utils/NgramLM.py:16-78
* ``mdd_post``            align_canonical_decoded/stastics/score, egs/attention_aug/infer.py:162-209,405-433,338-342
* ``prefix_score``        SURVEY.md Appendix A.6 (not in the reference)
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmdd_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile oracle/mdd_oracle.c with gcc (seconds).  Building the checker is not using it."""
    src = os.path.join(_HERE, "mdd_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


# ------------------------------------------------------------------------------------------------- CTC loss
def ctc_loss(lp, targets, in_len, tgt_len, blank=0, zero_infinity=False, grad_out=None, want_grad=True,
             want_alpha=False, dtype=np.float64):
    """Returns dict(nll[B], grad[T,B,C] or None, alpha[B,T,2Lmax+1] or None).  targets: (B,Lmax) or 1-D."""
    lp = np.ascontiguousarray(lp, dtype=dtype)
    T, B, Cc = lp.shape
    targets = np.ascontiguousarray(targets, dtype=np.int64)
    in_len = np.ascontiguousarray(in_len, dtype=np.int64)
    tgt_len = np.ascontiguousarray(tgt_len, dtype=np.int64)
    stride = targets.shape[1] if targets.ndim == 2 else 0
    Lmax = int(tgt_len.max()) if B else 0
    if targets.ndim == 2:
        Lmax = max(Lmax, 0)
    nll = np.zeros(B, dtype=dtype)
    grad = np.zeros_like(lp) if want_grad else None
    alpha = np.full((B, T, 2 * Lmax + 1), -np.inf, dtype=dtype) if want_alpha else None
    go = None if grad_out is None else np.ascontiguousarray(grad_out, dtype=dtype)
    fn = lib().orc_ctc_loss_f64 if dtype == np.float64 else lib().orc_ctc_loss_f32
    ty = C.c_double if dtype == np.float64 else C.c_float
    rc = fn(_p(lp, ty), T, B, Cc, _p(targets, C.c_int64), C.c_int64(stride), _p(in_len, C.c_int64),
            _p(tgt_len, C.c_int64), blank, int(zero_infinity), _p(go, ty), _p(nll, ty), _p(grad, ty),
            _p(alpha, ty), Lmax)
    if rc:
        raise RuntimeError("oracle ctc_loss: bad lengths")
    return {"nll": nll, "grad": grad, "alpha": alpha}


def ctc_loss_logits(logits, targets, in_len, tgt_len, blank=0, zero_infinity=False, grad_out=None, dtype=np.float64):
    """CTC loss on the output layer's logits: log_softmax (models/model_ctc.py:214-218) -> ctc_loss -> the chain rule
    through log_softmax, g_x = g_lp - softmax * sum_c g_lp (what torch autograd computes at steps/train_ctc.py:96).
    Returns dict(nll[B], grad[T,B,C] w.r.t. the logits, row_lse[T,B], argmax[T,B])."""
    x = np.asarray(logits, dtype=dtype)
    m = x.max(axis=-1, keepdims=True)
    lse = (m + np.log(np.exp(x - m).sum(axis=-1, keepdims=True)))
    lp = x - lse
    r = ctc_loss(lp, targets, in_len, tgt_len, blank, zero_infinity, grad_out, dtype=dtype)
    g = r["grad"]
    gx = g - np.exp(lp) * g.sum(axis=-1, keepdims=True)
    return {"nll": r["nll"], "grad": gx, "row_lse": lse[..., 0], "argmax": np.asarray(logits).argmax(axis=-1)}


# ------------------------------------------------------------------------------------------------- greedy
def greedy_tokens(lp, in_len, blank=0, word_id=None):
    lp = np.ascontiguousarray(lp, dtype=np.float32)
    T, B, Cc = lp.shape
    in_len = np.ascontiguousarray(in_len, dtype=np.int64)
    tokens = np.zeros((B, max(T, 1)), dtype=np.int32)
    out_len = np.zeros(B, dtype=np.int32)
    amax = np.zeros((B, max(T, 1)), dtype=np.int32)
    wid = None if word_id is None else np.ascontiguousarray(word_id, dtype=np.int32)
    rc = lib().orc_greedy(_p(lp, C.c_float), T, B, Cc, _p(in_len, C.c_int64), blank, _p(wid, C.c_int32),
                          _p(tokens, C.c_int32), _p(out_len, C.c_int32), _p(amax, C.c_int32))
    if rc:
        raise RuntimeError("oracle greedy: bad lengths")
    return tokens, out_len, amax


def word_ids(int2char):
    """ids such that two class indices compare equal iff their word strings are equal (ctcDecoder.py:83-84)."""
    n = max(int2char.keys()) + 1 if isinstance(int2char, dict) else len(int2char)
    seen, out = {}, np.zeros(n, dtype=np.int32)
    for i in range(n):
        w = int2char[i] if (not isinstance(int2char, dict) or i in int2char) else None
        out[i] = seen.setdefault(w, len(seen))
    return out


def greedy_decode(lp, in_len, int2char, blank=0):
    tokens, out_len, _ = greedy_tokens(lp, in_len, blank, word_ids(int2char))
    return ["".join(" " + int2char[int(k)] for k in tokens[b, :out_len[b]]) for b in range(len(out_len))]


# ------------------------------------------------------------------------------------------------- edit distance
def edit_align(hyp, ref):
    """hyp, ref: lists of int lists.  Returns (dist[B], ops list[str-list], status[B])."""
    B = len(hyp)
    hl = np.array([len(h) for h in hyp], dtype=np.int32)
    rl = np.array([len(r) for r in ref], dtype=np.int32)
    hs, rs = max(int(hl.max()) if B else 0, 1), max(int(rl.max()) if B else 0, 1)
    H = np.zeros((B, hs), dtype=np.int32)
    R = np.zeros((B, rs), dtype=np.int32)
    for b in range(B):
        H[b, :hl[b]] = hyp[b]
        R[b, :rl[b]] = ref[b]
    dist = np.zeros(B, dtype=np.int32)
    ops = np.zeros((B, hs + rs), dtype=np.uint8)
    ol = np.zeros(B, dtype=np.int32)
    st = np.zeros(B, dtype=np.int32)
    lib().orc_edit_align(_p(H, C.c_int32), hs, _p(hl, C.c_int32), _p(R, C.c_int32), rs, _p(rl, C.c_int32), B,
                         _p(dist, C.c_int32), _p(ops, C.c_uint8), hs + rs, _p(ol, C.c_int32), _p(st, C.c_int32))
    paths = [[chr(x) for x in ops[b, :ol[b]]] for b in range(B)]
    return dist, paths, st


def wer(s1: str, s2: str):
    """Decoder.wer (attention_aug variant): (distance, op path).  Empty side -> TypeError like the reference."""
    words = {}
    w1 = [words.setdefault(w, len(words)) for w in s1.split()]
    w2 = [words.setdefault(w, len(words)) for w in s2.split()]
    dist, paths, st = edit_align([w1], [w2])
    if st[0]:
        raise TypeError("cannot unpack non-iterable int object")
    return int(dist[0]), paths[0]


def compute_wer(index, in_len, targets, tgt_len):
    index = np.ascontiguousarray(index, dtype=np.int64)
    B, T = index.shape
    targets = np.ascontiguousarray(targets, dtype=np.int64)
    in_len = np.ascontiguousarray(in_len, dtype=np.int64)
    tgt_len = np.ascontiguousarray(tgt_len, dtype=np.int64)
    e, n = C.c_int64(0), C.c_int64(0)
    lib().orc_compute_wer(_p(index, C.c_int64), B, T, _p(in_len, C.c_int64), _p(targets, C.c_int64),
                          targets.shape[1], _p(tgt_len, C.c_int64), C.byref(e), C.byref(n))
    return int(e.value), int(n.value)


# ------------------------------------------------------------------------------------------------- LM table
def parse_arpa(path):
    """NgramLM.py:25-58: tab-split ARPA, log10 -> ln; unigram['UNK'] = unigram['<unk>']."""
    scale = math.log(10)
    uni, bi = {}, {}
    rec = 0
    with open(path, "r") as f:
        for lines in f.readlines():
            line = lines.strip("\n")
            if line == "\\1-grams:":
                rec = 1
                continue
            if line == "\\2-grams:":
                rec = 2
                continue
            parts = line.split("\t")
            tgt = uni if rec == 1 else bi if rec == 2 else None
            if tgt is None:
                continue
            if len(parts) == 3:
                tgt[parts[1]] = [scale * float(parts[0]), scale * float(parts[2])]
            elif len(parts) == 2:
                tgt[parts[1]] = [scale * float(parts[0]), 0.0]
    uni["UNK"] = uni["<unk>"]
    return uni, bi


def lm_table(path, int2char, n_class):
    """(C+1)x(C+1) fp64 table of get_bi_prob (NgramLM.py:65-78); index C = ''; NaN where the lookup raises KeyError."""
    uni, bi = parse_arpa(path)

    def get(w1, w2):
        if w1 == "":
            w1 = "<s>"
        if w2 == "":
            w2 = "</s>"
        key = w1 + " " + w2
        if key not in bi:
            return uni[w1][1] + uni[w2][0]
        return bi[key][0]

    tab = np.full((n_class + 1, n_class + 1), np.nan, dtype=np.float64)
    names = [int2char[i] for i in range(n_class)] + [""]
    for i, a in enumerate(names):
        for j, b in enumerate(names):
            try:
                tab[i, j] = get(a, b)
            except KeyError:
                pass
    return tab


# ------------------------------------------------------------------------------------------------- beam search
BEAM_EXC = {1: IndexError, 2: ValueError, 3: KeyError}


def beam_tokens(probs, in_len, beam_width, blank=0, lm=None, lm_alpha=0.0):
    probs = np.ascontiguousarray(probs, dtype=np.float32)
    B, T, Cc = probs.shape
    il = None if in_len is None else np.ascontiguousarray(in_len, dtype=np.int64)
    tokens = np.zeros((B, max(T, 1)), dtype=np.int32)
    out_len = np.zeros(B, dtype=np.int32)
    st = np.zeros(B, dtype=np.int32)
    score = np.zeros(B, dtype=np.float64)
    lmv = None if lm is None else np.ascontiguousarray(lm, dtype=np.float64)
    lib().orc_beam_search(_p(probs, C.c_float), B, T, Cc, _p(il, C.c_int64), int(beam_width), blank,
                          _p(lmv, C.c_double), C.c_double(lm_alpha), _p(tokens, C.c_int32), max(T, 1),
                          _p(out_len, C.c_int32), _p(st, C.c_int32), _p(score, C.c_double))
    return tokens, out_len, st, score


def beam_decode(probs, in_len, int2char, beam_width, lm, lm_alpha, blank=0):
    """ctcBeamSearch.decode on PROBABILITIES (B,T,C): list[str]; raises like the reference on the first bad utterance."""
    tokens, out_len, st, _ = beam_tokens(probs, in_len, beam_width, blank, lm, lm_alpha)
    res = []
    for b in range(len(out_len)):
        if st[b]:
            raise BEAM_EXC[int(st[b])]("oracle beam search status %d at utterance %d" % (st[b], b))
        res.append(" ".join(int2char[int(k)] for k in tokens[b, :out_len[b]]))
    return res


# ------------------------------------------------------------------------------------------------- MDD post
PAD_D, PAD_I = -2, -3


def mdd_post(dec, can, ops):
    """Integer form of infer.py:335-342.  Returns aligned rows, fault lists and counts."""
    dec = np.asarray(dec, dtype=np.int32)
    can = np.asarray(can, dtype=np.int32)
    o = np.frombuffer("".join(ops).encode(), dtype=np.uint8).copy()
    n = len(o)
    a_dec = np.zeros(max(n, 1), np.int32)
    a_can = np.zeros(max(n, 1), np.int32)
    a_ops = np.zeros(max(n, 1), np.uint8)
    ins = np.zeros(max(n, 1), np.int32)
    sub = np.zeros(max(n, 1), np.int32)
    dele = np.zeros(max(n, 1), np.int32)
    cnt = np.zeros(6, np.int32)
    lib().orc_mdd_post(_p(dec, C.c_int32), len(dec), _p(can, C.c_int32), len(can), _p(o, C.c_uint8), n,
                       _p(a_dec, C.c_int32), _p(a_can, C.c_int32), _p(a_ops, C.c_uint8), _p(ins, C.c_int32),
                       _p(sub, C.c_int32), _p(dele, C.c_int32), _p(cnt, C.c_int32))
    m, ni, ns, nd, cc, ds = (int(x) for x in cnt)
    return {"dec": a_dec[:m].tolist(), "can": a_can[:m].tolist(), "ops": [chr(x) for x in a_ops[:m]],
            "ins": ins[:ni].tolist(), "sub": sub[:ns].tolist(), "del": dele[:nd].tolist(),
            "correct": cc, "del_sub": ds}


def corpus_counts(lab, lc_ops, dec, dc_ops):
    """Per-utterance counters of test_ctc_nosil.py:218-293 on integer tokens: lab/dec = human / recognised token
    lists, lc_ops/dc_ops = op paths of wer(labels, canonical) / wer(decoded, canonical).
    Returns [TA, FR, FA, TR_correct, TR_wrong, n_canonical, FR_m, FA_m] (the _m pair under the m_speaker rule)."""
    lab = np.asarray(lab, dtype=np.int32)
    dec = np.asarray(dec, dtype=np.int32)
    o1 = np.frombuffer("".join(lc_ops).encode(), dtype=np.uint8).copy()
    o2 = np.frombuffer("".join(dc_ops).encode(), dtype=np.uint8).copy()
    cnt = np.zeros(12, np.int32)
    lib().orc_corpus_counts(_p(lab, C.c_int32), len(lab), _p(o1, C.c_uint8), len(o1), _p(dec, C.c_int32), len(dec),
                            _p(o2, C.c_uint8), len(o2), _p(cnt, C.c_int32))
    return [int(x) for x in cnt[:8]]


def mdd_score(n_ins, correct, del_sub):
    """infer.py:341-342 (ZeroDivisionError on an empty canonical side, like the reference)."""
    tmp = min(n_ins / 4, 0.1 * (correct + del_sub))
    return math.ceil((1 - (del_sub + tmp) / (del_sub + correct)) * 100)


# ------------------------------------------------------------------------------------------------- prefix scorer
def prefix_init(x, blank=0):
    x = np.ascontiguousarray(x, dtype=np.float64)
    T, Cc = x.shape
    r0 = np.zeros((T, 2), dtype=np.float64)
    lib().orc_prefix_init(_p(x, C.c_double), T, Cc, blank, _p(r0, C.c_double))
    return r0


def prefix_score(x, r_prev, last, glen, blank=0):
    """One scoring step for all C candidates: (psi[C], r[C,T,2], psi_eos)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    T, Cc = x.shape
    r_prev = np.ascontiguousarray(r_prev, dtype=np.float64)
    psi = np.zeros(Cc, dtype=np.float64)
    r = np.zeros((Cc, T, 2), dtype=np.float64)
    eos = C.c_double(0)
    lib().orc_prefix_score(_p(x, C.c_double), T, Cc, blank, _p(r_prev, C.c_double), int(last), int(glen),
                           _p(psi, C.c_double), _p(r, C.c_double), C.byref(eos))
    return psi, r, float(eos.value)


def joint_beam_search(x, lens, att_fn, beam=10, ctc_weight=0.3, max_len=None, blank=0, eos=None):
    """fp64 restatement of mddctc.prefix.joint_beam_search: x (T,B,C) log-probs, lens[B],
    att_fn(step, state) -> (B,W,C) float32 attention log-probs; state = dict of numpy arrays (tokens, last, glen, flag).
    Returns (tokens[B,W,max_len] int32, scores[B,W] fp64, glen[B,W], flag[B,W])."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    T, B, Cc = x.shape
    W = beam
    max_len = max_len or T
    lens = np.ascontiguousarray(lens, dtype=np.int64)
    r = np.full((B, W, T, 2), -np.inf, dtype=np.float64)
    for b in range(B):
        r0 = prefix_init(x[:lens[b], b], blank)
        r[b, :, :lens[b]] = r0[None]
    last = np.full((B, W), -1, dtype=np.int32)
    glen = np.zeros((B, W), dtype=np.int32)
    psi = np.zeros((B, W), dtype=np.float64)
    score = np.full((B, W), -np.inf, dtype=np.float64)
    flag = np.full((B, W), 2, dtype=np.int32)
    score[:, 0] = 0.0
    flag[:, 0] = 0
    tokens = np.full((B, W, max_len), -1, dtype=np.int32)
    for step in range(max_len):
        att = np.ascontiguousarray(att_fn(step, {"tokens": tokens, "last": last, "glen": glen, "flag": flag}),
                                   dtype=np.float32)
        assert att.shape == (B, W, Cc)
        rc = lib().orc_joint_step(_p(x, C.c_double), T, B, Cc, _p(lens, C.c_int64), blank, -1 if eos is None else eos,
                                  W, C.c_double(ctc_weight), max_len, _p(att, C.c_float), _p(r, C.c_double),
                                  _p(last, C.c_int32), _p(glen, C.c_int32), _p(psi, C.c_double), _p(score, C.c_double),
                                  _p(flag, C.c_int32), _p(tokens, C.c_int32))
        if rc:
            raise RuntimeError("orc_joint_step failed")
    return tokens, score, glen, flag
