"""MDD alignment / scoring drop-ins (egs/attention_aug/infer.py) and train-time PER (models/model_ctc.py).

* ``align_canonical_decoded(s1, s2, l)``   infer.py:162-209   (single-utterance call surface)
* ``stastics(dc_path, canonicals, decoded)`` infer.py:405-433
* ``print_aligned_string(s1, s2, l)``      infer.py:155-160   (pure string formatting)
* ``compute_wer(index, input_sizes, targets, target_sizes)``  models/model_ctc.py:227-244
* ``MDDScorer``                            the batched per-utterance loop of infer.py:304-342: sil/err filtering,
                                           wer path, aligned triples, fault lists, counts and the 0-100 score for a
                                           whole batch in one kernel launch.

All alignment arithmetic runs in the CUDA library; the host only maps words <-> ints and formats strings.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib, _rt
from .decoder import _words_to_ids, edit_align_tokens

PAD_D, PAD_I = -2, -3


def print_aligned_string(s1, s2, l):
    s1 = [p + " " if len(p) == 1 else p for p in s1]
    s2 = [p + " " if len(p) == 1 else p for p in s2]
    l = [s + " " for s in l]
    return " ".join(s1), " ".join(s2), " ".join(l)


def _post_align(dec_lists, can_lists, paths):
    """mdd_post_align on a batch of (decoded tokens, canonical tokens, op path)."""
    L = _lib.lib()
    dev = _rt.device()
    A, la, Bm, lb, inv = _words_to_ids(dec_lists, can_lists)
    B = len(paths)
    pl = np.array([len(p) for p in paths], dtype=np.int32)
    stride = max(A.shape[1] + Bm.shape[1], int(pl.max()) if B else 1, 1)
    P = np.zeros((B, stride), dtype=np.uint8)
    for i, p in enumerate(paths):
        P[i, :pl[i]] = [ord(c) for c in p]
    t = lambda a: torch.from_numpy(a).to(dev)
    dA, dla, dB, dlb, dP, dpl = t(A), t(la), t(Bm), t(lb), t(P), t(pl)
    out = {k: torch.empty((B, stride), dtype=torch.int32, device=dev) for k in ("a_dec", "a_can", "ins", "sub", "del")}
    out["a_ops"] = torch.empty((B, stride), dtype=torch.uint8, device=dev)
    out["counts"] = torch.zeros((B, 8), dtype=torch.int32, device=dev)
    post = _lib.PostOut(out["a_dec"].data_ptr(), out["a_can"].data_ptr(), out["a_ops"].data_ptr(),
                        out["ins"].data_ptr(), out["sub"].data_ptr(), out["del"].data_ptr(), out["counts"].data_ptr())
    rc = L.mdd_post_align(dA.data_ptr(), dA.stride(0), dla.data_ptr(), dB.data_ptr(), dB.stride(0), dlb.data_ptr(), B,
                          A.shape[1], stride - A.shape[1], dP.data_ptr(), stride, dpl.data_ptr(), _lib.C.byref(post),
                          stride, _rt.stream_ptr(dev))
    _lib.check(rc, "mdd_post_align")
    return {k: v.cpu().numpy() for k, v in out.items()}, inv


def _name(inv, k):
    return "D" if k == PAD_D else "I" if k == PAD_I else inv[k]


def align_canonical_decoded(s1, s2, l):
    """Pads the decoded list with 'D' and the canonical list with 'I' along the op path, keeps at most one leading
    insertion and drops it if it duplicates the next decoded phone.  Returns (s1, s2, l) like infer.py:162-209."""
    if len(l) == 0:
        raise IndexError("list index out of range")         # l[0] at infer.py:204
    r, inv = _post_align([list(s1)], [list(s2)], [list(l)])
    n = int(r["counts"][0, 0])
    return ([_name(inv, int(k)) for k in r["a_dec"][0, :n]], [_name(inv, int(k)) for k in r["a_can"][0, :n]],
            [chr(c) for c in r["a_ops"][0, :n]])


def stastics(dc_path, phones_canonicals, phones_decoded):
    """(insertion_fault, substution_fault, deletion_fault) lists from three equal-length aligned lists
    (infer.py:405-433): a pure gather by op code, done on the host from the caller's own lists."""
    ins, sub, dele = [], [], []
    for op, can, dec in zip(dc_path, phones_canonicals, phones_decoded):
        if op == "S":
            sub.append(can)
        elif op == "I":
            ins.append(dec)
        elif op != "-":
            dele.append(can)
    return ins, sub, dele


def mdd_score(n_ins, correct_cnt, del_sub_cnt):
    """infer.py:341-342, evaluated in Python floats exactly like the reference (ZeroDivisionError included)."""
    tmp_score = min(n_ins / 4, 0.1 * (correct_cnt + del_sub_cnt))
    return math.ceil((1 - (del_sub_cnt + tmp_score) / (del_sub_cnt + correct_cnt)) * 100)


def compute_wer(index, input_sizes, targets, target_sizes):
    """CTC_Model.compute_wer: collapse the frame argmax ids, sum edit distances against the labels.
    index (B,T) ints, input_sizes (B), targets (B,Lmax), target_sizes (B) -- numpy arrays or tensors.
    Returns (batch_errs, batch_tokens) as Python ints."""
    L = _lib.lib()
    dev = _rt.device(index if isinstance(index, torch.Tensor) else None)
    idx = torch.as_tensor(np.asarray(index) if not isinstance(index, torch.Tensor) else index).to(dev)
    B, T = idx.shape
    amax = idx.t().contiguous().to(torch.uint8)                       # (T,B) layout of the collapse kernel
    il = torch.as_tensor(np.asarray(input_sizes) if not isinstance(input_sizes, torch.Tensor) else input_sizes)
    il = il.to(device=dev, dtype=torch.int64).contiguous()
    tg = torch.as_tensor(np.asarray(targets) if not isinstance(targets, torch.Tensor) else targets)
    tg = tg.to(device=dev, dtype=torch.int32).contiguous()
    tl = torch.as_tensor(np.asarray(target_sizes) if not isinstance(target_sizes, torch.Tensor) else target_sizes)
    tl = tl.to(device=dev, dtype=torch.int32).contiguous()
    tokens = torch.empty((B, max(T, 1)), dtype=torch.int32, device=dev)
    out_len = torch.empty(B, dtype=torch.int32, device=dev)
    rc = L.mdd_collapse(amax.data_ptr(), T, B, il.data_ptr(), 0, None, tokens.data_ptr(), tokens.stride(0),
                        out_len.data_ptr(), _rt.stream_ptr(dev))
    _lib.check(rc, "mdd_collapse")
    r = edit_align_tokens(tg, tl, tokens, out_len, tg.size(1), max(T, 1), want_ops=False)
    return int(r["dist"].sum().item()), int(tl.sum().item())


class MDDScorer:
    """Batched form of the per-utterance loop of infer.py:304-342.

    ``vocab`` maps class index -> phone string.  Decoded tokens (class ids from the greedy / beam kernels) and
    canonical tokens (class ids of the canonical transcript) stay on the device; 'sil' is dropped from both sides
    (:310-320), 'err' is removed from the decoded side with the reference's *substring* rule
    (``str.replace('err','')``, :324) so a word merely containing 'err' maps to its remainder.
    """

    def __init__(self, vocab, sil="sil", err="err"):
        n = max(vocab.keys()) + 1 if isinstance(vocab, dict) else len(vocab)
        self.names = [vocab[i] for i in range(n)]
        table = {}
        self.can_map = np.full(n, -1, dtype=np.int32)
        self.dec_map = np.full(n, -1, dtype=np.int32)
        for i, w in enumerate(self.names):
            if w == sil:
                continue
            self.can_map[i] = table.setdefault(w, len(table))
            w2 = w.replace(err, "")
            if w2.split():
                self.dec_map[i] = table.setdefault(" ".join(w2.split()), len(table))
        self.words = [None] * len(table)
        for w, k in table.items():
            self.words[k] = w
        self._dev_maps = {}

    def _maps(self, dev):
        if dev.index not in self._dev_maps:
            self._dev_maps[dev.index] = (torch.from_numpy(self.dec_map).to(dev), torch.from_numpy(self.can_map).to(dev))
        return self._dev_maps[dev.index]

    def align_tokens(self, dec, dec_len, can, can_len, max_dec=None, max_can=None):
        """dec (B,*) / can (B,*) int32 CUDA class ids.  Returns the dict of device tensors of edit_align_tokens."""
        dm, cm = self._maps(dec.device)
        return edit_align_tokens(dec, dec_len, can, can_len, max_dec or dec.size(1), max_can or can.size(1), hyp_map=dm,
                                 ref_map=cm, want_ops=True, want_post=True)

    def name(self, k):
        return "D" if k == PAD_D else "I" if k == PAD_I else self.words[k]

    def to_python(self, r):
        """Host-side rendering of align_tokens output: one dict per utterance, or a TypeError instance where the
        reference's decoder.wer would have raised it (an empty side after filtering)."""
        h = {k: v.cpu().numpy() for k, v in r.items()}
        out = []
        for b in range(h["dist"].shape[0]):
            if h["status"][b]:
                out.append(TypeError("cannot unpack non-iterable int object"))
                continue
            n, ni, ns, nd, cc, ds = (int(x) for x in h["counts"][b, :6])
            try:
                score = mdd_score(ni, cc, ds)
            except ZeroDivisionError as e:
                score = e
            out.append({"path0": [chr(c) for c in h["ops"][b, :h["ops_len"][b]]],
                        "dec": [self.name(int(k)) for k in h["a_dec"][b, :n]],
                        "can": [self.name(int(k)) for k in h["a_can"][b, :n]],
                        "path": [chr(c) for c in h["a_ops"][b, :n]],
                        "ins": [self.name(int(k)) for k in h["ins"][b, :ni]],
                        "sub": [self.name(int(k)) for k in h["sub"][b, :ns]],
                        "del": [self.name(int(k)) for k in h["del"][b, :nd]],
                        "correct": cc, "del_sub": ds, "score": score, "dist": int(h["dist"][b])})
        return out
