"""CTC prefix scorer for CTC/attention joint beam decoding (BASELINE.json north_star).

The reference has no attention decoder and no such function; its nearest code is the (prBlank, prNonBlank, prTotal)
recursion and the sort + [:beamWidth] of egs/attention_aug/utils/BeamSearch.py:52-66,99-125.  This module provides the
standard Watanabe/ESPnet-style scorer on the CUDA kernels (mdd_prefix_probs / mdd_prefix_init / mdd_prefix_score /
mdd_prefix_select / mdd_prefix_advance) and the label-synchronous joint beam search that fuses it with externally supplied attention
log-probabilities:

    score(g.c) = score(g) + ctc_weight * (psi(g.c) - psi(g)) + (1 - ctc_weight) * att_logp[c]

Scores and CTC state are fp64 on the device (see csrc/prefix_score.cu); a decoding step is three kernel launches
(score, select, advance) and whatever the caller's attention decoder does -- no torch.topk / gather in between, and
joint_beam_search allocates its beam buffers once (two state sets used in turn), not per step.
"""
from __future__ import annotations

import torch

from . import _lib, _rt

LIVE, FINISHED, DEAD = 0, 1, 2


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
        # emission probabilities in fp64, utterance-major, exponentiated once for all the decoding steps
        self.p_lin = torch.empty((self.B, self.T, self.C), dtype=torch.float64, device=self.dev)
        _lib.check(self._L.mdd_prefix_probs(self.x.data_ptr(), self.T, self.B, self.C, self.p_lin.data_ptr(),
                                            _rt.stream_ptr(self.dev)), "mdd_prefix_probs")

    def initial_state(self, max_len=None, all_slots_live=False):
        """The empty prefix.  By default only slot 0 of every utterance is live (a beam search starts from ONE empty
        hypothesis); ``all_slots_live`` makes every slot an independent empty hypothesis (scoring given prefixes)."""
        B, W, T, dev = self.B, self.W, self.T, self.dev
        r = torch.empty((B, W, T, 2), dtype=torch.float64, device=dev)
        _lib.check(self._L.mdd_prefix_init(self.x.data_ptr(), T, B, self.C, _rt.ptr(self.lens), self.blank, W,
                                           r.data_ptr(), _rt.stream_ptr(dev)), "mdd_prefix_init")
        flag = torch.full((B, W), LIVE if all_slots_live else DEAD, dtype=torch.int32, device=dev)
        score = torch.full((B, W), 0.0 if all_slots_live else -float("inf"), dtype=torch.float64, device=dev)
        flag[:, 0] = LIVE
        score[:, 0] = 0.0
        return {"r": r, "last": torch.full((B, W), -1, dtype=torch.int32, device=dev),
                "glen": torch.zeros((B, W), dtype=torch.int32, device=dev),
                "psi": torch.zeros((B, W), dtype=torch.float64, device=dev),          # psi(empty prefix) = log 1
                "score": score, "flag": flag,
                "tokens": torch.full((B, W, max_len or T), -1, dtype=torch.int32, device=dev)}

    def empty_state(self, like):
        """Uninitialised buffers shaped like ``like`` (the second state set of a beam search)."""
        return {k: torch.empty_like(v) for k, v in like.items()}

    def score(self, state, out=None):
        """psi[B,W,C] = log p_ctc(g.c, ...) for every candidate label, psi_eos[B,W] = log p_ctc(g) (fp64).
        Finished / dead hypotheses are not extended (all -inf).  ``out`` = (psi, psi_eos) buffers to reuse."""
        if out is None:
            out = (torch.empty((self.B, self.W, self.C), dtype=torch.float64, device=self.dev),
                   torch.empty((self.B, self.W), dtype=torch.float64, device=self.dev))
        psi, eos = out
        _lib.check(self._L.mdd_prefix_score(self.p_lin.data_ptr(), self.T, self.B, self.C, _rt.ptr(self.lens),
                                            self.blank, self.W, state["r"].data_ptr(), state["last"].data_ptr(),
                                            state["glen"].data_ptr(), state["flag"].data_ptr(), psi.data_ptr(),
                                            eos.data_ptr(), _rt.stream_ptr(self.dev)), "mdd_prefix_score")
        return psi, eos

    def select(self, state, psi, psi_eos, att_logp, ctc_weight=0.3, eos=None, out=None):
        """One beam step: the W best of the W*C candidates of every utterance.  Returns the new state WITHOUT its CTC
        recursion state (call advance and store the result under "r") plus (sel_hyp, sel_label).  ``out`` = a state
        dict whose buffers are overwritten (its "r" is left alone)."""
        B, W, C, dev = self.B, self.W, self.C, self.dev
        att = att_logp
        if att.dtype != torch.float32 or att.device != dev or not att.is_contiguous():
            att = att.to(device=dev, dtype=torch.float32).contiguous()
        if tuple(att.shape) != (B, W, C):
            raise RuntimeError("att_logp must be (B, W, C)")
        max_len = state["tokens"].shape[2]
        new = out if out is not None else self.empty_state({k: v for k, v in state.items() if k not in ("r", "sel_hyp")})
        if "sel_hyp" not in new:
            new["sel_hyp"] = torch.empty((B, W), dtype=torch.int32, device=dev)
        sel_hyp = new["sel_hyp"]
        rc = self._L.mdd_prefix_select(B, W, C, self.blank, -1 if eos is None else int(eos), float(ctc_weight), max_len,
                                       psi.data_ptr(), psi_eos.data_ptr(), att.data_ptr(), state["score"].data_ptr(),
                                       state["psi"].data_ptr(), state["glen"].data_ptr(), state["flag"].data_ptr(),
                                       state["tokens"].data_ptr(), sel_hyp.data_ptr(), new["last"].data_ptr(),
                                       new["score"].data_ptr(), new["psi"].data_ptr(), new["glen"].data_ptr(),
                                       new["flag"].data_ptr(), new["tokens"].data_ptr(), _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_prefix_select")
        return new, sel_hyp, new["last"]

    def advance(self, state, sel_hyp, sel_label, new_flag=None, out=None):
        """CTC recursion state r (B,W,T,2) of the chosen (parent hypothesis, label) pairs; slots whose new flag is
        non-zero (finished / dead) or whose parent is -1 are left untouched.  ``out`` = the buffer to write."""
        sel_hyp = sel_hyp.to(device=self.dev, dtype=torch.int32).contiguous()
        sel_label = sel_label.to(device=self.dev, dtype=torch.int32).contiguous()
        r_new = out if out is not None else torch.empty_like(state["r"])
        _lib.check(self._L.mdd_prefix_advance(self.p_lin.data_ptr(), self.T, self.B, self.C, _rt.ptr(self.lens),
                                              self.blank, self.W, state["r"].data_ptr(), state["last"].data_ptr(),
                                              state["glen"].data_ptr(), sel_hyp.data_ptr(), sel_label.data_ptr(),
                                              _rt.ptr(new_flag), r_new.data_ptr(), _rt.stream_ptr(self.dev)),
                   "mdd_prefix_advance")
        return r_new

    def extend(self, state, labels):
        """Append labels[B,W] (int, -1 = leave the slot) to the hypotheses in place of a beam step: the state of
        g.c for given prefixes (used to score known label sequences)."""
        labels = labels.to(device=self.dev, dtype=torch.int32).contiguous()
        psi, _ = self.score(state)
        hyp = torch.arange(self.W, dtype=torch.int32, device=self.dev)[None, :].expand(self.B, self.W).contiguous()
        keep = labels >= 0
        hyp = torch.where(keep, hyp, torch.full_like(hyp, -1))
        r_new = self.advance(state, hyp, labels)
        new = dict(state)
        new["r"] = torch.where(keep[:, :, None, None], r_new, state["r"])
        new["last"] = torch.where(keep, labels, state["last"])
        new["glen"] = state["glen"] + keep.to(torch.int32)
        new["psi"] = torch.where(keep, torch.gather(psi, 2, labels.clamp_min(0).long()[..., None])[..., 0], state["psi"])
        return new


def joint_beam_search(log_probs, input_lengths, att_logp_fn, beam=10, ctc_weight=0.3, max_len=None, blank=0, eos=None):
    """Label-synchronous joint CTC/attention beam search (BASELINE config 4 shape: beam 10, CTC weight 0.3).

    att_logp_fn(step, state) -> (B,W,C) attention-decoder log-probs for the hypotheses of ``state`` (its "tokens"
    (B,W,max_len) int32, "last" and "glen" describe them; synthetic in the benchmarks, since the reference has no
    attention decoder).  With ``eos`` a hypothesis ends when eos is chosen (scored with psi_eos = log p_ctc(g)) and is
    carried unchanged from then on; without it every hypothesis runs to max_len labels.
    Returns (tokens[B,W,max_len] int32 (-1 padded), scores[B,W] fp64, lengths[B,W] int32, flags[B,W] int32), best first.
    """
    scorer = CTCPrefixScorer(log_probs, input_lengths, blank=blank, beam=beam)
    max_len = max_len or scorer.T
    state = scorer.initial_state(max_len=max_len)
    state["sel_hyp"] = torch.empty((scorer.B, beam), dtype=torch.int32, device=scorer.dev)
    other = scorer.empty_state(state)
    out = (torch.empty((scorer.B, beam, scorer.C), dtype=torch.float64, device=scorer.dev),
           torch.empty((scorer.B, beam), dtype=torch.float64, device=scorer.dev))
    for step in range(max_len):
        psi, psi_eos = scorer.score(state, out=out)
        att = att_logp_fn(step, state)
        new, sel_h, sel_l = scorer.select(state, psi, psi_eos, att, ctc_weight=ctc_weight, eos=eos, out=other)
        scorer.advance(state, sel_h, sel_l, new_flag=new["flag"], out=new["r"])
        state, other = new, state
    return state["tokens"], state["score"], state["glen"], state["flag"]
