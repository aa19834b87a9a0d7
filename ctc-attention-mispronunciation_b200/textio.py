"""Host-side conventions either side of the CTC path: length bookkeeping of the collate
function, the line formats of the evaluation logs and the CMU -> IPA rendering.  Pure host logic (lists, strings and a few
torch scalars on the CPU); the kernels take what these functions produce.

Reference:
  utils/data_loader.py:151-181 (create_input: fractional float32 input sizes), steps/train_ctc.py:68
      (input_sizes = (input_sizes * out_len).long()),
  steps/test_ctc_nosil.py:186-214 (label / canonical strings, sil filter, decode_seq.log / human_seq.log /
      transcribe.log lines),
  infer.py:328-333 (phone lists, CMU -> IPA through dict/phonetic_dict.py:21-64).
"""
from __future__ import annotations

import os
from typing import Iterable, List, Sequence

import torch

# ----------------------------------------------------------------------------------------------- lengths (rank 3)


def fractional_input_sizes(feature_lengths: Sequence[int]) -> torch.Tensor:
    """What the reference's collate stores per utterance: ``feature_length / inputs_max_length`` written into a float32
    tensor (data_loader.py:163,176,181)."""
    n = [int(v) for v in feature_lengths]
    mx = max(n)
    out = torch.zeros(len(n))
    for i, v in enumerate(n):
        out[i] = v / mx                       # Python float division, rounded to float32 on assignment -- as the reference
    return out.float()


def frame_lengths(input_sizes: torch.Tensor, out_len: int) -> torch.Tensor:
    """``(input_sizes * out_len).long()`` (train_ctc.py:68): float32 product, truncated.  Because the fraction was
    rounded to float32 first, the result is one frame short of ``feature_length * out_len // inputs_max_length`` for a
    few percent of the (length, maximum) pairs; kept bit-exact here because the loss, the greedy decode
    and the PER all consume these lengths."""
    return (input_sizes.float() * out_len).long()


def exact_frame_lengths(feature_lengths: Sequence[int], out_len: int) -> torch.Tensor:
    """The integer arithmetic the fractional detour approximates: ``feature_length * out_len // max_length``."""
    n = torch.as_tensor([int(v) for v in feature_lengths], dtype=torch.int64)
    return n * int(out_len) // int(n.max())


def collate_lengths(feature_lengths: Sequence[int], label_lengths: Sequence[int], out_len: int, device=None):
    """Lengths of one batch as the training step hands them to the loss: (input_lengths int64, target_lengths int64),
    optionally already on the device (train_ctc.py:45-47 moves them there)."""
    il = frame_lengths(fractional_input_sizes(feature_lengths), out_len)
    tl = torch.as_tensor([int(v) for v in label_lengths], dtype=torch.int64)
    if device is not None:
        il, tl = il.to(device), tl.to(device)
    return il, tl


def bucketed_batches(lengths: Sequence[int], batch_size: int, width: int = 64, shuffle_seed=None) -> List[List[int]]:
    """Length-bucketing sampler (the reference shuffles only): utterances are grouped into buckets of `width` frames and
    batches are cut inside a bucket, so a ragged batch wastes at most `width` frames per utterance.  Deterministic; with
    `shuffle_seed` the order of the batches (not their content) is permuted."""
    if batch_size <= 0 or width <= 0:
        raise ValueError("batch_size and width must be positive")
    order = sorted(range(len(lengths)), key=lambda i: (int(lengths[i]) // width, i))
    batches, cur, cur_bucket = [], [], None
    for i in order:
        b = int(lengths[i]) // width
        if cur and (b != cur_bucket or len(cur) == batch_size):
            batches.append(cur)
            cur = []
        cur_bucket = b
        cur.append(i)
    if cur:
        batches.append(cur)
    if shuffle_seed is not None:
        g = torch.Generator().manual_seed(int(shuffle_seed))
        batches = [batches[i] for i in torch.randperm(len(batches), generator=g).tolist()]
    return batches


# ----------------------------------------------------------------------------------------------- text formats (rank 4)


def ids_to_string(ids: Iterable[int], length: int, index2word) -> str:
    """``' '.join(index2word[num] for num in row[:length])`` (test_ctc_nosil.py:187-191)."""
    row = list(ids)[:int(length)]
    return " ".join(index2word[int(n)] for n in row)


def strip_sil(seq: str) -> str:
    """The 'nosil' strings of test_ctc_nosil.py:199-209: split on single blanks, drop tokens equal to 'sil', re-join.
    (Splitting on ' ' keeps empty tokens, exactly as the reference does.)"""
    return " ".join(t for t in seq.split(" ") if t != "sil")


def seq_log_lines(utt_list: Sequence[str], seqs: Sequence[str]) -> List[str]:
    """Lines of decode_seq.log / human_seq.log / transcribe.log: ``utt + ' ' + seq + '\\n'`` (test_ctc_nosil.py:212-214)."""
    if len(utt_list) != len(seqs):
        raise AssertionError("one sequence per utterance")
    return [u + " " + s + "\n" for u, s in zip(utt_list, seqs)]


class EvalLogs:
    """The three evaluation logs of test_ctc_nosil.py:140-142,212-214,343-345, written next to each other."""

    NAMES = {"decoded": "decode_seq.log", "labels": "human_seq.log", "canonicals": "transcribe.log"}

    def __init__(self, directory="."):
        self.directory = directory
        self._files = {k: open(os.path.join(directory, v), "w+") for k, v in self.NAMES.items()}

    def write_batch(self, utt_list, decoded, labels, canonicals):
        """`decoded`, `labels`, `canonicals`: the raw strings; 'sil' is stripped here as in the reference loop."""
        d = [strip_sil(s) for s in decoded]
        l = [strip_sil(s) for s in labels]
        c = [strip_sil(s) for s in canonicals]
        self._files["canonicals"].writelines(seq_log_lines(utt_list, c))
        self._files["labels"].writelines(seq_log_lines(utt_list, l))
        self._files["decoded"].writelines(seq_log_lines(utt_list, d))
        return d, l, c

    def close(self):
        for f in self._files.values():
            f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def phone_list(seq: str) -> List[str]:
    """``[c for c in seq.split(' ') if c]`` (infer.py:328-329)."""
    return [c for c in seq.split(" ") if c]


# CMU39 + AH0/ER0 -> IPA, the US-syllable table the reference renders with (dict/phonetic_dict.py:21-64)
_CMU_IPA = ("AA a|AE æ|AH0 ə|AH ʌ|AO ɔ|AW aʊ|AY aɪ|EH e|ER ɜ|ER0 ər|EY eɪ|IH ɪ|IY i|OW oʊ|OY ɔɪ|UH ʊ|UW u|B b|CH tʃ|D d|"
            "DH ð|F f|G g|HH h|JH dʒ|K k|L l|M m|N n|NG ŋ|P p|R r|S s|SH ʃ|T t|TH θ|V v|W w|Y j|Z z|ZH ʒ")
CMU_TO_IPA = dict(item.split(" ") for item in _CMU_IPA.split("|"))
IPA_TO_CMU = {v: k for k, v in CMU_TO_IPA.items()}


def cmu_to_ipa(phones: Sequence[str]) -> List[str]:
    """``[cmu_to_ipa_wiki.get(c.upper(), c) for c in phones]`` (infer.py:331-333): unknown units pass through."""
    return [CMU_TO_IPA.get(c.upper(), c) for c in phones]
