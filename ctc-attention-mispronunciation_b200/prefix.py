"""CTC prefix scorer for CTC/attention joint beam decoding (BASELINE.json north_star).

The reference has no attention decoder and no such function; its nearest code is the (prBlank, prNonBlank, prTotal)
recursion of egs/attention_aug/utils/BeamSearch.py:52-66,99-125.  This module provides the standard
Watanabe/ESPnet-style scorer on the CUDA kernels (mdd_prefix_init / mdd_prefix_score / mdd_prefix_advance) and a small
joint beam-search driver that fuses it with externally supplied attention log-probabilities:

    score(g.c) = ctc_weight * (psi(g.c) - psi(g)) + (1 - ctc_weight) * att_logp[c]
"""
from __future__ import annotations

import torch

from . import _lib, _rt


class CTCPrefixScorer:
    def __init__(self, log_probs, input_lengths=None, blank=0, beam=10):
        """log_probs: (T,B,C) log-softmax outputs; input_lengths: (B) or None."""
        self.dev = _rt.device(log_probs)
        self.x = log_probs.to(device=self.dev, dtype=torch.float32).contiguous()
        self.T, self.B, self.C = self.x.shape
        self.W = beam
        self.blank = blank
        self.lens = None if input_lengths is None else input_lengths.to(device=self.dev, dtype=torch.int64).contiguous()
        self._L = _lib.lib()

    def initial_state(self):
        r = torch.empty((self.B, self.W, self.T, 2), dtype=torch.float32, device=self.dev)
        _lib.check(self._L.mdd_prefix_init(self.x.data_ptr(), self.T, self.B, self.C, _rt.ptr(self.lens), self.blank,
                                           self.W, r.data_ptr(), _rt.stream_ptr(self.dev)), "mdd_prefix_init")
        return {"r": r, "last": torch.full((self.B, self.W), -1, dtype=torch.int32, device=self.dev),
                "glen": torch.zeros((self.B, self.W), dtype=torch.int32, device=self.dev),
                "psi": torch.zeros((self.B, self.W), dtype=torch.float32, device=self.dev)}

    def score(self, state):
        """psi[B,W,C] = log p_ctc(g.c, ...) for every candidate label, psi_eos[B,W] = log p_ctc(g)."""
        psi = torch.empty((self.B, self.W, self.C), dtype=torch.float32, device=self.dev)
        eos = torch.empty((self.B, self.W), dtype=torch.float32, device=self.dev)
        _lib.check(self._L.mdd_prefix_score(self.x.data_ptr(), self.T, self.B, self.C, _rt.ptr(self.lens), self.blank,
                                            self.W, state["r"].data_ptr(), state["last"].data_ptr(),
                                            state["glen"].data_ptr(), psi.data_ptr(), eos.data_ptr(),
                                            _rt.stream_ptr(self.dev)), "mdd_prefix_score")
        return psi, eos

    def advance(self, state, sel_hyp, sel_label, keep=None, psi=None):
        """New state for the chosen (parent hypothesis, label) pairs.  keep[b,w] False leaves slot w untouched
        (finished hypotheses)."""
        sel_hyp = sel_hyp.to(device=self.dev, dtype=torch.int32).contiguous()
        sel_label = sel_label.to(device=self.dev, dtype=torch.int32).contiguous()
        r_new = torch.empty_like(state["r"])
        _lib.check(self._L.mdd_prefix_advance(self.x.data_ptr(), self.T, self.B, self.C, _rt.ptr(self.lens), self.blank,
                                              self.W, state["r"].data_ptr(), state["last"].data_ptr(),
                                              state["glen"].data_ptr(), sel_hyp.data_ptr(), sel_label.data_ptr(),
                                              r_new.data_ptr(), _rt.stream_ptr(self.dev)), "mdd_prefix_advance")
        idx = sel_hyp.long().clamp_min(0)
        last = sel_label.clone()
        glen = torch.gather(state["glen"], 1, idx) + 1
        new_psi = state["psi"] if psi is None else torch.gather(psi.reshape(self.B, -1), 1,
                                                               idx * self.C + sel_label.long().clamp_min(0))
        if keep is not None:
            keep = keep.to(self.dev)
            r_new = torch.where(keep[:, :, None, None], r_new, state["r"])
            last = torch.where(keep, last, state["last"])
            glen = torch.where(keep, glen, state["glen"])
            new_psi = torch.where(keep, new_psi, state["psi"])
        return {"r": r_new, "last": last, "glen": glen.to(torch.int32), "psi": new_psi}


def joint_beam_search(log_probs, input_lengths, att_logp_fn, beam=10, ctc_weight=0.3, max_len=None, blank=0, eos=None):
    """Label-synchronous joint CTC/attention beam search (BASELINE config 4 shape: beam 10, CTC weight 0.3).

    att_logp_fn(step, prefixes[B,W,step]) -> (B,W,C) attention-decoder log-probs (synthetic in the benchmarks, since
    the reference has no attention decoder).  Ends after max_len steps; returns (tokens[B,W,max_len], scores[B,W]).
    """
    scorer = CTCPrefixScorer(log_probs, input_lengths, blank=blank, beam=beam)
    B, W, C, dev = scorer.B, beam, scorer.C, scorer.dev
    max_len = max_len or scorer.T
    state = scorer.initial_state()
    tokens = torch.zeros((B, W, max_len), dtype=torch.int64, device=dev)
    score = torch.full((B, W), -float("inf"), device=dev)
    score[:, 0] = 0.0                                    # a single live (empty) hypothesis per utterance
    for step in range(max_len):
        psi, _ = scorer.score(state)
        att = att_logp_fn(step, tokens[:, :, :step])
        delta = psi - state["psi"][:, :, None]
        cand = score[:, :, None] + ctc_weight * delta + (1.0 - ctc_weight) * att
        cand[:, :, blank] = -float("inf")
        if eos is not None:
            cand[:, :, eos] = -float("inf")
        top, flat = cand.reshape(B, W * C).topk(W, dim=1)
        sel_h, sel_l = (flat // C).to(torch.int32), (flat % C).to(torch.int32)
        state = scorer.advance(state, sel_h, sel_l, psi=psi)
        tokens = torch.gather(tokens, 1, sel_h.long()[:, :, None].expand(B, W, max_len)).clone()
        tokens[:, :, step] = sel_l.long()
        score = top
    return tokens, score
