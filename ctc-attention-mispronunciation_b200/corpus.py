"""Corpus-level MDD evaluation: the loop of egs/attention_aug/steps/test_ctc_nosil.py:188-343 on the CUDA library.

* ``print_align_space_canonical_origin(s1, s2, l)``   test_ctc_nosil.py:33-82 (single-pair call surface; string
                                                      formatting and the dict ``d`` on the host from the caller's lists)
* ``CorpusMDD``                                       the batched evaluation: 'sil' filtering (:196-209), the three
                                                      alignments per utterance (:219-221), the TA / FR / FA / TR
                                                      counters incl. the m_speaker block (:236-297), PER, and the
                                                      Precision / Recall / F1 the reference prints (:299-343).

Alignment and counting run in libmddctc.so (``mdd_edit_align`` x3 + ``mdd_corpus_counts``); only 24 integers come
back to the host per corpus batch.  No CPU fallback.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, _rt
from .decoder import edit_align_tokens

COUNT_NAMES = ["true_accept", "false_rejection", "false_accept", "true_rejection_correct_diagnose",
               "true_rejection_wrong_diagnose", "total_phonemes_in_canonical"]


def _pad3(w):
    return w + " " * (3 - len(w)) if len(w) in (1, 2) else w


def print_align_space_canonical_origin(s1, s2, l):
    """(canonical line, hypothesis line, op line, d): d[j] = '-' | 'S'+hypothesis phone | 'D' for canonical position j
    and d['I'] = [str(j-1)+str(j)] per insertion (test_ctc_nosil.py:33-82).  s1 = hypothesis, s2 = canonical."""
    hyp, can = s1.split(" "), s2.split(" ")
    d = {j: "" for j in range(len(can))}
    d["I"] = []
    row_h, row_c = [], []
    ih = ic = 0
    for op in l:
        if op == "I":
            d["I"].append(str(ic - 1) + str(ic))
            row_h.append(hyp[ih])
            row_c.append("I")
            ih += 1
        elif op == "D":
            d[ic] = "D"
            row_h.append("D")
            row_c.append(can[ic])
            ic += 1
        else:
            d[ic] = op + (hyp[ih] if op == "S" else "")
            row_h.append(hyp[ih])
            row_c.append(can[ic])
            ih += 1
            ic += 1
    row_h += hyp[ih:]                                     # a path shorter than its rows leaves the tails in place
    row_c += can[ic:]
    return (" ".join(_pad3(w) for w in row_c), " ".join(_pad3(w) for w in row_h), " ".join(s + "  " for s in l), d)


class CorpusMDD:
    """Batched corpus evaluation over class-id rows on the device.

    ``vocab`` maps class index -> phone string (``Vocab.index2word``); ``sil`` is dropped from all three sides."""

    def __init__(self, vocab, sil="sil", m_speaker=("TXHC",)):
        n = max(vocab.keys()) + 1 if isinstance(vocab, dict) else len(vocab)
        self.names = [vocab[i] for i in range(n)]
        table = {}
        self.map = np.full(n, -1, dtype=np.int32)
        for i, w in enumerate(self.names):
            if w != sil:
                self.map[i] = table.setdefault(w, len(table))
        self.m_speaker = list(m_speaker)
        self._dev = {}
        self.reset()

    def reset(self):
        self.totals = np.zeros((2, 12), dtype=np.int64)

    def _map(self, dev):
        if dev.index not in self._dev:
            self._dev[dev.index] = torch.from_numpy(self.map).to(dev)
        return self._dev[dev.index]

    def evaluate_tokens(self, dec, dec_len, lab, lab_len, can, can_len, group=None, accumulate=True, token_map=None):
        """dec / lab / can: (B,*) int32 CUDA class ids with int32 lengths; ``group`` optional uint8[B] (m_speaker
        membership); ``token_map`` replaces the vocabulary's id map (id < 0 drops the token).  Returns device tensors
        {"counts": (B,12) int32, "status": (B,) int32, "totals": (2,12) int64, ...} and, when ``accumulate``, adds the
        totals to ``self.totals`` (one 192-byte D2H copy)."""
        L = _lib.lib()
        dev = dec.device
        m = self._map(dev) if token_map is None else token_map
        B = dec.size(0)
        lc = edit_align_tokens(lab, lab_len, can, can_len, lab.size(1), can.size(1), hyp_map=m, ref_map=m, want_ops=True)
        dl = edit_align_tokens(dec, dec_len, lab, lab_len, dec.size(1), lab.size(1), hyp_map=m, ref_map=m, want_ops=False)
        dc = edit_align_tokens(dec, dec_len, can, can_len, dec.size(1), can.size(1), hyp_map=m, ref_map=m, want_ops=True)
        counts = torch.empty((B, 12), dtype=torch.int32, device=dev)
        status = torch.empty(B, dtype=torch.int32, device=dev)
        totals = torch.zeros((2, 12), dtype=torch.int64, device=dev)
        rc = L.mdd_corpus_counts(lab.data_ptr(), lab.stride(0), lab_len.data_ptr(), dec.data_ptr(), dec.stride(0),
                                 dec_len.data_ptr(), B, lab.size(1), dec.size(1), can.size(1), m.data_ptr(), m.data_ptr(),
                                 m.numel(), lc["ops"].data_ptr(), lc["ops"].stride(0), lc["ops_len"].data_ptr(),
                                 dc["ops"].data_ptr(), dc["ops"].stride(0), dc["ops_len"].data_ptr(),
                                 dl["dist"].data_ptr(), _rt.ptr(group), counts.data_ptr(), totals.data_ptr(),
                                 status.data_ptr(), _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_corpus_counts")
        if accumulate:
            self.totals += totals.cpu().numpy()
        return {"counts": counts, "status": status, "totals": totals, "lc": lc, "dc": dc, "dl": dl}

    def evaluate_strings(self, utt_list, labels, canonicals, decoded, sil="sil"):
        """The reference's call surface: lists of space-joined phone strings as they are before the 'sil' filter
        (:188-209).  Words are compared as strings, so they get per-call ids.  Raises TypeError where
        ``decoder.wer`` would (a side that is empty after filtering)."""
        dev = _rt.device()
        table = {}

        def enc(strings):
            rows = [[table.setdefault(w, len(table)) for w in s.split(" ")] for s in strings]
            A = np.zeros((len(rows), max(max((len(r) for r in rows), default=1), 1)), dtype=np.int32)
            for i, r in enumerate(rows):
                A[i, :len(r)] = r
            return A, np.array([len(r) for r in rows], dtype=np.int32)

        (D, dl), (Lb, ll), (Cn, cl) = enc(decoded), enc(labels), enc(canonicals)
        tmap = np.array([-1 if w in (sil, "") else k for w, k in table.items()], dtype=np.int32)
        t = lambda a: torch.from_numpy(a).to(dev)
        grp = t(np.array([1 if u[:4] in self.m_speaker else 0 for u in utt_list], dtype=np.uint8))
        r = self.evaluate_tokens(t(D), t(dl), t(Lb), t(ll), t(Cn), t(cl), group=grp, accumulate=False, token_map=t(tmap))
        if bool((r["status"] != 0).any().item()):
            raise TypeError("cannot unpack non-iterable int object")
        self.totals += r["totals"].cpu().numpy()
        return r

    def scoring(self, m_speaker=False):
        """FAR / FRR / DER and the other rates of the offline scorer from this evaluator's totals (see scoring_rates):
        its per-error-type counters collapse onto TA / FR / FA / correct / wrong diagnosis."""
        ta, fr, fa, trc, trw = (int(x) for x in self.totals[1 if m_speaker else 0][:5])
        return scoring_rates({"cor_cor": ta, "cor_nocor": fr, "sub_nosub": fa, "sub_sub": trc, "sub_sub1": trw})

    def summary(self):
        """Reference-named counters and the rates printed at test_ctc_nosil.py:299-343 (ZeroDivisionError included)."""
        out = {}
        for tag, row in (("", self.totals[0]), ("m_", self.totals[1])):
            ta, fr, fa, trc, trw, ncan = (int(x) for x in row[:6])
            errs, nlab = int(row[9]), int(row[8])
            tr = trc + trw
            out.update({tag + k: v for k, v in zip(COUNT_NAMES, (ta, fr, fa, trc, trw, ncan))})
            out.update({tag + "total_wer": errs, tag + "num_word": nlab, tag + "true_rejection": tr})
            if tag == "" or nlab:
                p = float(tr) / (tr + fr)
                r = float(tr) / (tr + fa)
                out.update({tag + "PER": float(errs) / nlab * 100, tag + "precision": p * 100, tag + "recall": r * 100,
                            tag + "f1": 2 * p * r / (p + r) * 100})
        return out


SCORING_COUNTERS = ("cor_cor", "cor_nocor", "sub_sub", "sub_sub1", "sub_nosub", "ins_ins", "ins_ins1", "ins_noins",
                    "del_del", "del_del1", "del_nodel")


def scoring_rates(counters):
    """The rates egs/attention_aug/result/utils/scoring/ins_del_sub_cor_analysis.py:193-219 derives from its eleven
    counters (x_x: error of type x detected and diagnosed correctly, x_x1: detected with the wrong diagnosis, x_nox:
    accepted; cor_cor / cor_nocor: correct phones accepted / rejected).  Missing counters count as 0.
    ZeroDivisionError where the reference raises it."""
    c = {k: int(counters.get(k, 0)) for k in SCORING_COUNTERS}
    total = sum(c.values())
    tp = c["sub_sub"] + c["ins_ins"] + c["del_del"] + c["sub_sub1"] + c["ins_ins1"] + c["del_del1"]
    fp = c["cor_nocor"]
    fn = c["sub_nosub"] + c["ins_noins"] + c["del_nodel"]
    recall = tp / (tp + fn)
    precision = tp / (tp + fp)
    correct_diag = c["sub_sub"] + c["ins_ins"] + c["del_del"]
    error_diag = c["sub_sub1"] + c["ins_ins1"] + c["del_del1"]
    err_count = tp + fn
    return {"sum": total, "recall": recall, "precision": precision, "f1": 2 * precision * recall / (recall + precision),
            "TA": c["cor_cor"] / (c["cor_cor"] + c["cor_nocor"]), "n_TA": c["cor_cor"],
            "FR": c["cor_nocor"] / (c["cor_cor"] + c["cor_nocor"]), "n_FR": c["cor_nocor"],
            "FA": fn / err_count, "n_FA": fn,
            "correct_diag": correct_diag / (correct_diag + error_diag), "n_correct_diag": correct_diag,
            "error_diag": error_diag / (correct_diag + error_diag), "n_error_diag": error_diag,
            "FAR": 1 - recall, "FRR": c["cor_nocor"] / (c["cor_nocor"] + c["cor_cor"]),
            "DER": error_diag / (error_diag + correct_diag)}


def scoring_report(counters):
    """The text ins_del_sub_cor_analysis.py:194-219 prints, byte for byte."""
    r = scoring_rates(counters)
    return ("sum: %d\n" % r["sum"] + "Recall: %.4f\n" % r["recall"] + "Precision: %.4f\n" % r["precision"]
            + "f1:%.4f\n" % r["f1"] + "TA: %.4f %d\n" % (r["TA"], r["n_TA"]) + "FR: %.4f %d\n" % (r["FR"], r["n_FR"])
            + "FA: %.4f %d\n" % (r["FA"], r["n_FA"]) + "Correct Diag: %.4f %d\n" % (r["correct_diag"], r["n_correct_diag"])
            + "Error Diag: %.4f %d\n" % (r["error_diag"], r["n_error_diag"]) + "FAR: %.4f\n" % r["FAR"]
            + "FRR: %.4f\n" % r["FRR"] + "DER: %.4f\n" % r["DER"])
