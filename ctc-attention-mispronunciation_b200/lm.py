"""ARPA bigram LM -> dense fp64 table for the beam-search kernel.

Host-side table builder mirroring LanguageModel (egs/attention_aug/utils/NgramLM.py:16-78): tab-separated ARPA,
log10 -> ln scaling (:24), unigram['UNK'] aliasing (:59), get_bi_prob back-off rule (:65-78) with '' standing for <s>
(as w1) and </s> (as w2).  The dense table is a pure re-indexing of get_bi_prob; the fused score itself
(table * lm_alpha at use) is evaluated in the CUDA kernel.
"""
from __future__ import annotations

import math

import numpy as np


class LanguageModel:
    """Same constructor and get_bi_prob/get_uni_prob/score_bg behaviour as the reference class."""

    def __init__(self, arpa_file=None, n_gram=2, start="<s>", end="</s>", unk="<unk>"):
        self.n_gram = n_gram
        self.start = start
        self.end = end
        self.unk = unk
        self.scale = math.log(10)
        self.unigram = {}
        self.bigram = {}
        self._read(arpa_file)

    def _read(self, fn):
        section = 0
        with open(fn, "r") as f:          # open(None) raises TypeError exactly like the reference
            for raw in f.readlines():
                line = raw.strip("\n")
                if line == "\\1-grams:":
                    section = 1
                    continue
                if line == "\\2-grams:":
                    section = 2
                    continue
                if section not in (1, 2):
                    continue
                parts = line.split("\t")
                table = self.unigram if section == 1 else self.bigram
                if len(parts) == 3:
                    table[parts[1]] = [self.scale * float(parts[0]), self.scale * float(parts[2])]
                elif len(parts) == 2:
                    table[parts[1]] = [self.scale * float(parts[0]), 0.0]
        self.unigram["UNK"] = self.unigram[self.unk]

    def get_uni_prob(self, wid):
        return self.unigram[wid][0]

    def get_bi_prob(self, w1, w2):
        if w1 == "":
            w1 = self.start
        if w2 == "":
            w2 = self.end
        key = w1 + " " + w2
        if key not in self.bigram:
            return self.unigram[w1][1] + self.unigram[w2][0]
        return self.bigram[key][0]

    def score_bg(self, sentence):
        val = 0.0
        words = sentence.strip().split()
        val += self.get_bi_prob(self.start, words[0])
        for i in range(len(words) - 1):
            val += self.get_bi_prob(words[i], words[i + 1])
        val += self.get_bi_prob(words[-1], self.end)
        return val

    def dense_table(self, classes, n_class):
        """(n_class+1)^2 fp64: [i,j] = get_bi_prob(classes[i], classes[j]); index n_class is ''; NaN = KeyError."""
        names = [classes[i] for i in range(n_class)] + [""]
        tab = np.full((n_class + 1, n_class + 1), np.nan, dtype=np.float64)
        for i, a in enumerate(names):
            for j, b in enumerate(names):
                try:
                    tab[i, j] = self.get_bi_prob(a, b)
                except KeyError:
                    pass
        return tab
