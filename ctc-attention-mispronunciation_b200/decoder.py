"""Decoder / GreedyDecoder / BeamDecoder drop-ins on the sm_100a kernels.

Mirrors egs/attention_aug/utils/ctcDecoder.py (same class names, constructor arguments, return types and exception
behaviour); the cnn-rnn-ctc recipe's variant (egs/cnn-rnn-ctc/utils/ctcDecoder.py, wer() returns the int only) is
selected with ``variant='cnn-rnn-ctc'``.  Strings are assembled on the host from integer outputs of the kernels;
every number comes from the CUDA library (no CPU fallback).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, _rt
from .lm import LanguageModel

_EXC = {1: IndexError, 2: ValueError, 3: KeyError}


def _words_to_ids(seqs_a, seqs_b):
    """Per-pair word -> int maps (only equality matters, ctcDecoder.py:120-124); returns padded int32 arrays."""
    B = len(seqs_a)
    la = np.array([len(s) for s in seqs_a], dtype=np.int32)
    lb = np.array([len(s) for s in seqs_b], dtype=np.int32)
    A = np.zeros((B, max(int(la.max()) if B else 0, 1)), dtype=np.int32)
    Bm = np.zeros((B, max(int(lb.max()) if B else 0, 1)), dtype=np.int32)
    table = {}
    for i in range(B):
        A[i, :la[i]] = [table.setdefault(w, len(table)) for w in seqs_a[i]]
        Bm[i, :lb[i]] = [table.setdefault(w, len(table)) for w in seqs_b[i]]
    inv = [None] * len(table)
    for w, k in table.items():
        inv[k] = w
    return A, la, Bm, lb, inv


def edit_align_tokens(hyp, hyp_len, ref, ref_len, max_hyp, max_ref, hyp_map=None, ref_map=None, want_ops=True,
                      want_post=False):
    """Device-side batched Levenshtein (+ traceback, + MDD post-processing).  All arguments are CUDA int32 tensors.
    Returns a dict of device tensors."""
    L = _lib.lib()
    dev = hyp.device
    B = hyp.size(0)
    stride = max(max_hyp + max_ref, 1)
    out = {"dist": torch.empty(B, dtype=torch.int32, device=dev),
           "status": torch.empty(B, dtype=torch.int32, device=dev)}
    ops = ops_len = None
    if want_ops:
        ops = torch.empty((B, stride), dtype=torch.uint8, device=dev)
        ops_len = torch.empty(B, dtype=torch.int32, device=dev)
        out["ops"], out["ops_len"] = ops, ops_len
    post = None
    if want_post:
        for k in ("a_dec", "a_can", "ins", "sub", "del"):
            out[k] = torch.empty((B, stride), dtype=torch.int32, device=dev)
        out["a_ops"] = torch.empty((B, stride), dtype=torch.uint8, device=dev)
        out["counts"] = torch.zeros((B, 8), dtype=torch.int32, device=dev)
        post = _lib.PostOut(out["a_dec"].data_ptr(), out["a_can"].data_ptr(), out["a_ops"].data_ptr(),
                            out["ins"].data_ptr(), out["sub"].data_ptr(), out["del"].data_ptr(),
                            out["counts"].data_ptr())
    need = _lib.C.c_size_t(0)
    _lib.check(L.mdd_edit_align_workspace_bytes(B, max_hyp, max_ref, _lib.C.byref(need)), "mdd_edit_align_workspace_bytes")
    ws = _rt.workspace(need.value, dev, "edit")
    n_map = 0
    if hyp_map is not None:
        n_map = hyp_map.numel()
    if ref_map is not None:
        n_map = ref_map.numel() if n_map == 0 else min(n_map, ref_map.numel())
    rc = L.mdd_edit_align(hyp.data_ptr(), hyp.stride(0), hyp_len.data_ptr(), ref.data_ptr(), ref.stride(0),
                          ref_len.data_ptr(), B, max_hyp, max_ref, _rt.ptr(hyp_map), _rt.ptr(ref_map), n_map,
                          out["dist"].data_ptr(), _rt.ptr(ops), stride, _rt.ptr(ops_len), out["status"].data_ptr(),
                          _lib.C.byref(post) if post is not None else None, ws.data_ptr(), ws.numel(),
                          _rt.stream_ptr(dev))
    _lib.check(rc, "mdd_edit_align")
    return out


class Decoder(object):
    """Base class: converts model output into text and scores it (ctcDecoder.py:9-184)."""

    def __init__(self, int2char, space_idx=1, blank_index=0, variant="attention_aug"):
        self.int_to_char = int2char
        self.space_idx = space_idx
        self.blank_index = blank_index
        self.num_word = 0
        self.num_char = 0
        self.variant = variant

    def decode(self):
        raise NotImplementedError

    # ---------------------------------------------------------------------------------------------- scoring
    def wer_batch(self, s1_list, s2_list):
        """Batched wer(): list of (distance, op-path) -- one kernel launch for the whole list.
        Entries whose hypothesis or reference is empty are returned as TypeError instances (attention_aug variant)."""
        a = [s.split() for s in s1_list]
        b = [s.split() for s in s2_list]
        return self._edit_batch(a, b)

    def _edit_batch(self, a, b):
        dev = _rt.device()
        A, la, Bm, lb, _ = _words_to_ids(a, b)
        r = edit_align_tokens(torch.from_numpy(A).to(dev), torch.from_numpy(la).to(dev), torch.from_numpy(Bm).to(dev),
                              torch.from_numpy(lb).to(dev), A.shape[1], Bm.shape[1], want_ops=True)
        dist = r["dist"].cpu().numpy()
        st = r["status"].cpu().numpy()
        ops = r["ops"].cpu().numpy()
        ol = r["ops_len"].cpu().numpy()
        res = []
        for i in range(len(a)):
            if st[i]:
                res.append(TypeError("cannot unpack non-iterable int object") if self.variant == "attention_aug"
                           else int(dist[i]))
            elif self.variant == "attention_aug":
                res.append((int(dist[i]), [chr(c) for c in ops[i, :ol[i]]]))
            else:
                res.append(int(dist[i]))
        return res

    def wer(self, s1, s2):
        """Word-level edit distance of two space-separated strings.  attention_aug: (distance, op path)
        (ctcDecoder.py:118-127); cnn-rnn-ctc: the distance only.  An empty side raises TypeError in the
        attention_aug variant, as the reference does."""
        r = self.wer_batch([s1], [s2])[0]
        if isinstance(r, Exception):
            raise r
        return r

    def cer(self, s1, s2):
        """Character-level edit distance (ctcDecoder.py:129-132)."""
        r = self._edit_batch([list(s1)], [list(s2)])[0]
        if isinstance(r, Exception):
            raise r
        return r[0] if isinstance(r, tuple) else r

    # ---------------------------------------------------------------------------------------------- string helpers
    def _word_ids(self, n_class):
        """int32[C]: equal ids <=> equal word strings (the repeat rule compares strings, ctcDecoder.py:83-84)."""
        seen, out = {}, np.zeros(n_class, dtype=np.int32)
        for i in range(n_class):
            out[i] = seen.setdefault(self.int_to_char[i], len(seen))
        return out


class GreedyDecoder(Decoder):
    """Frame-wise argmax + collapse (ctcDecoder.py:186-200).  Returns strings with a LEADING space per token."""

    def decode_tokens(self, prob_tensor, frame_seq_len):
        """Device part: returns (tokens[B,T] int32, out_len[B] int32, argmax[T,B] uint8) CUDA tensors."""
        L = _lib.lib()
        dev = _rt.device(prob_tensor)
        lp = prob_tensor.to(device=dev, dtype=torch.float32).contiguous()
        T, B, C = lp.shape
        if isinstance(frame_seq_len, torch.Tensor):
            lens = frame_seq_len.to(device=dev, dtype=torch.int64).contiguous()
        else:
            host = [int(x) for x in frame_seq_len]
            if any(x > T for x in host):
                raise IndexError("index %d is out of bounds for axis 0 with size %d" % (T, T))
            lens = torch.tensor(host, dtype=torch.int64, device=dev)
        wid = torch.from_numpy(self._word_ids(C)).to(dev)
        amax = torch.empty((max(T, 1), B), dtype=torch.uint8, device=dev)
        tokens = torch.empty((B, max(T, 1)), dtype=torch.int32, device=dev)
        out_len = torch.empty(B, dtype=torch.int32, device=dev)
        rc = L.mdd_greedy_decode(lp.data_ptr(), T, B, C, lens.data_ptr(), self.blank_index, wid.data_ptr(),
                                 amax.data_ptr(), tokens.data_ptr(), tokens.stride(0), out_len.data_ptr(),
                                 _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_greedy_decode")
        return tokens, out_len, amax

    def decode(self, prob_tensor, frame_seq_len):
        tokens, out_len, _ = self.decode_tokens(prob_tensor, frame_seq_len)
        tk, ol = tokens.cpu().numpy(), out_len.cpu().numpy()
        i2c = self.int_to_char
        if self.space_idx == -1:
            return ["".join(" " + i2c[int(k)] for k in tk[b, :ol[b]]) for b in range(len(ol))]
        sp = i2c[self.space_idx]
        return ["".join(" " if i2c[int(k)] == sp else i2c[int(k)] for k in tk[b, :ol[b]]) for b in range(len(ol))]


class BeamDecoder(Decoder):
    """CTC prefix beam search with bigram-LM fusion (ctcDecoder.py:202-226 -> BeamSearch.py:73-153)."""

    def __init__(self, int2char, beam_width=200, blank_index=0, space_idx=-1, lm_path=None, lm_alpha=0.01):
        self.beam_width = beam_width
        super(BeamDecoder, self).__init__(int2char, space_idx=space_idx, blank_index=blank_index)
        self.lm = LanguageModel(arpa_file=lm_path)      # mandatory, as in the reference (open(None) -> TypeError)
        self.lm_alpha = lm_alpha
        self._tables = {}

    def _lm_table(self, n_class, dev):
        key = (n_class, dev.index)
        if key not in self._tables:
            self._tables[key] = torch.from_numpy(self.lm.dense_table(self.int_to_char, n_class)).to(dev)
        return self._tables[key]

    def decode_probs_tokens(self, probs, frame_seq_len=None):
        """ctcBeamSearch.decode boundary: probs (B,T,C) float32 PROBABILITIES.  Returns device tensors
        (tokens[B,T], out_len[B], status[B], score[B])."""
        L = _lib.lib()
        dev = _rt.device(probs)
        pr = probs.to(device=dev, dtype=torch.float32).contiguous()
        B, T, C = pr.shape
        lens = None
        if frame_seq_len is not None:
            lens = (frame_seq_len.to(device=dev, dtype=torch.int64).contiguous() if isinstance(frame_seq_len, torch.Tensor)
                    else torch.tensor([int(x) for x in frame_seq_len], dtype=torch.int64, device=dev))
        tab = self._lm_table(C, dev)
        tokens = torch.empty((B, max(T, 1)), dtype=torch.int32, device=dev)
        out_len = torch.empty(B, dtype=torch.int32, device=dev)
        status = torch.empty(B, dtype=torch.int32, device=dev)
        score = torch.empty(B, dtype=torch.float64, device=dev)
        need = _lib.C.c_size_t(0)
        _lib.check(L.mdd_beam_search_workspace_bytes(B, T, C, self.beam_width, _lib.C.byref(need)),
                   "mdd_beam_search_workspace_bytes")
        ws = _rt.workspace(need.value, dev, "beam")
        rc = L.mdd_beam_search(pr.data_ptr(), B, T, C, _rt.ptr(lens), self.beam_width, self.blank_index,
                               tab.data_ptr(), float(self.lm_alpha), tokens.data_ptr(), tokens.stride(0),
                               out_len.data_ptr(), status.data_ptr(), score.data_ptr(), ws.data_ptr(), ws.numel(),
                               _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_beam_search")
        return tokens, out_len, status, score

    def decode_probs(self, probs, frame_seq_len=None):
        tokens, out_len, status, _ = self.decode_probs_tokens(probs, frame_seq_len)
        tk, ol, st = tokens.cpu().numpy(), out_len.cpu().numpy(), status.cpu().numpy()
        res = []
        for b in range(len(ol)):
            if st[b]:       # the reference raises at the first failing utterance of the batch
                raise _EXC.get(int(st[b]), RuntimeError)("beam search: utterance %d (status %d)" % (b, st[b]))
            res.append(" ".join(self.int_to_char[int(k)] for k in tk[b, :ol[b]]))
        return res

    def decode(self, prob_tensor, frame_seq_len=None):
        """prob_tensor: (T,B,C) log-probs.  CPU tensors take torch's CPU exp exactly like the reference
        (ctcDecoder.py:223-225), CUDA tensors are exponentiated on the device."""
        if prob_tensor.is_cuda:
            L = _lib.lib()
            lp = prob_tensor.to(torch.float32).contiguous()
            T, B, C = lp.shape
            probs = torch.empty((B, T, C), dtype=torch.float32, device=lp.device)
            _lib.check(L.mdd_exp_transpose(lp.data_ptr(), T, B, C, probs.data_ptr(), _rt.stream_ptr(lp.device)),
                       "mdd_exp_transpose")
        else:
            probs = torch.exp(prob_tensor.transpose(0, 1))
        return self.decode_probs(probs, frame_seq_len)
