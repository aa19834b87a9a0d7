"""CPU-only tests of the host side: the C-ABI library loads and exports every symbol of include/mddctc.h, the product
fails loudly without a GPU (no CPU fallback), host-side table builders match the reference fixtures, and the N>1
sharding logic works under gloo with world_size 2."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import GOLD, ROOT, jload, npz


@pytest.fixture(scope="module")
def M():
    import mddctc
    mddctc._lib.build()
    return mddctc


def test_cabi_exports_every_declared_symbol(M):
    hdr = open(os.path.join(ROOT, "include", "mddctc.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(mdd_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 16
    L = ctypes.CDLL(M._lib.SO_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert set(M._lib.EXPORTS) <= declared
    L.mdd_error_string.restype = ctypes.c_char_p
    assert L.mdd_version() >= 100 and L.mdd_error_string(-3).startswith(b"CUDA")


def test_library_is_sm100a_and_uses_bulk_copy(M):
    out = subprocess.run(["cuobjdump", "-lelf", M._lib.SO_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(M):
    from mddctc.decoder import GreedyDecoder
    from mddctc.loss import CTCLoss
    lp = torch.log_softmax(torch.randn(5, 2, 4), -1)
    with pytest.raises(M._lib.MddCtcLibraryError):
        CTCLoss()(lp, torch.ones(2, 2, dtype=torch.long), (5, 5), (2, 2))
    with pytest.raises(M._lib.MddCtcLibraryError):
        GreedyDecoder({0: "blank", 1: "a", 2: "b", 3: "c"}, space_idx=-1).decode(lp, [5, 5])


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "ctc-attention-mispronunciation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|libmdd_oracle|orc_[a-z]|oracle[./]", src, flags=re.M), f


def test_lm_table_matches_reference(M, arpa, vocab43):
    from mddctc.lm import LanguageModel
    lm = LanguageModel(arpa)
    want = npz("lm_table.npz")["table"]
    got = lm.dense_table(vocab43, 43)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])
    assert abs(lm.get_bi_prob("", "sil") - (-0.0593)) < 1e-3
    with pytest.raises(TypeError):
        LanguageModel(None)


def test_mdd_scorer_maps(M):
    from mddctc.mdd import MDDScorer, mdd_score, print_aligned_string, stastics
    s = MDDScorer({0: "blank", 1: "UNK", 2: "sil", 3: "a", 4: "err", 5: "berr", 6: "b"})
    assert s.dec_map[2] == -1 and s.can_map[2] == -1          # sil dropped on both sides
    assert s.dec_map[4] == -1 and s.can_map[4] >= 0           # err dropped on the decoded side only
    assert s.dec_map[5] == s.can_map[6]                       # 'berr'.replace('err','') == 'b'
    assert mdd_score(2, 9, 2) == 78                           # SURVEY.md Appendix B first KAT
    assert print_aligned_string(["t", "ow"], ["I", "ow"], ["I", "-"]) == ("t  ow", "I  ow", "I  - ")
    assert stastics(["I", "-", "S", "D"], ["I", "a", "b", "c"], ["x", "a", "y", "D"]) == (["x"], ["b"], ["c"])
    with pytest.raises(ZeroDivisionError):
        mdd_score(0, 0, 0)


def test_print_align_space_canonical_origin_mirror(M):
    """Host mirror of test_ctc_nosil.py:33-82 against the d maps and printed rows the real reference produced."""
    from mddctc.corpus import print_align_space_canonical_origin as pr
    data = jload("corpus_mdd.json.gz")
    for b, ref in enumerate(data["per_utt"]):
        c = data["canonicals"][b]
        for s1, path, want in ((data["labels"][b], ref["lc"], ref["d1"]), (data["decoded"][b], ref["dc"], ref["d2"])):
            r = pr(s1, c, list(path))
            assert [[str(k), v] for k, v in r[3].items()] == want
            assert len(r[0].split()) == len(path) == len(r[1].split())
    assert pr("a bb ccc", "a dd", ["-", "S", "I"])[:3] == ("a   dd  I  ", "a   bb  ccc", "-   S   I  ")


def test_shard_helpers(M):
    from mddctc.dist import balanced_buckets, shard_bounds
    for n in (0, 1, 7, 4096):
        for ws in (1, 2, 4, 8):
            b = [shard_bounds(n, ws, r) for r in range(ws)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(ws - 1))
    rng = np.random.default_rng(0)
    lens = rng.integers(50, 1501, size=2048)
    shards = balanced_buckets(lens, 8)
    assert sorted(i for s in shards for i in s) == list(range(2048))
    loads = [int(lens[s].sum()) for s in shards]
    assert max(loads) - min(loads) <= 1500


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
import mddctc
from mddctc.dist import ShardedCTCLoss, shard_bounds
rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=ws)
g = torch.Generator().manual_seed(0)
T, B, C, L = 30, 10, 7, 5
lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
tg = torch.randint(1, C, (B, L), generator=g)
il = torch.full((B,), T); tl = torch.full((B,), L)
def fake_raw(lp, tg, il, tl, blank=0, zero_infinity=False, grad_scalar=1.0, loss_sum=None, validate=False):
    # stands in for the CUDA kernel on this CPU-only box: torch CPU loss on the local slice
    x = lp.clone().requires_grad_(True)
    nll = torch.nn.functional.ctc_loss(x, tg, il, tl, reduction="none")
    nll.sum().backward()
    loss_sum.copy_(nll.sum().detach().reshape(1))
    return nll.detach(), x.grad * grad_scalar, None
lo, hi = shard_bounds(B, ws, rank)
total, nll, grad = ShardedCTCLoss(loss_fn_raw=fake_raw)(lp[:, lo:hi], tg[lo:hi], il[lo:hi], tl[lo:hi])
ref = torch.nn.functional.ctc_loss(lp, tg, il, tl, reduction="sum")
assert abs(float(total) - float(ref)) < 1e-3 * abs(float(ref)), (float(total), float(ref))
assert grad.shape == (T, hi - lo, C)
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_sharded_loss_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs


# ------------------------------------------------------------------------------- SURVEY 8f ranks 3 and 4 (host glue)
def test_collate_lengths_match_reference_create_input():
    """data_loader.py:151-181 + train_ctc.py:68, recorded from the reference's own create_input (gen_textio): the
    fractional float32 detour and its truncation are reproduced bit for bit, including the one-frame-short cases."""
    import torch
    from conftest import jload
    from mddctc import textio
    G = jload("textio.json.gz")
    short = total = 0
    for b in G["batches"]:
        frac = textio.fractional_input_sizes(b["feature_lengths"])
        assert frac.dtype == torch.float32
        assert [float(v) for v in frac] == b["input_sizes_f32"]
        for out_len, want in b["frame_lengths"].items():
            got = textio.frame_lengths(frac, int(out_len))
            assert got.dtype == torch.int64 and got.tolist() == want
            exact = textio.exact_frame_lengths(b["feature_lengths"], int(out_len)).tolist()
            assert all(e - 1 <= g <= e for g, e in zip(want, exact))
            short += sum(g != e for g, e in zip(want, exact))
            total += len(want)
        il, tl = textio.collate_lengths(b["feature_lengths"], b["label_lengths"], max(b["feature_lengths"]) // 4 or 1)
        assert tl.tolist() == b["target_sizes"] and il.dtype == torch.int64
    assert 0 < short < 0.1 * total          # the off-by-one really occurs (a few percent)


def test_eval_log_formats_match_reference(tmp_path):
    """test_ctc_nosil.py:196-214 executed by gen_textio: nosil strings and the three log files, byte for byte."""
    from conftest import jload
    from mddctc import textio
    G = jload("textio.json.gz")
    with textio.EvalLogs(str(tmp_path)) as logs:
        for c in G["logs"]:
            d, l, k = logs.write_batch(c["utt_list"], c["decoded"], c["labels"], c["canonicals"])
            assert d == c["decoded_nosil"] and l == c["labels_nosil"] and k == c["canonicals_nosil"]
            assert "".join(textio.seq_log_lines(c["utt_list"], d)) == c["decode_seq"]
    for key, name in textio.EvalLogs.NAMES.items():
        want = "".join(c[{"decoded": "decode_seq", "labels": "human_seq", "canonicals": "transcribe"}[key]] for c in G["logs"])
        assert open(tmp_path / name).read() == want
    assert textio.ids_to_string([3, 1, 2, 0, 0], 3, {0: "blank", 1: "a", 2: "b", 3: "c"}) == "c a b"
    assert textio.phone_list(" aa  b sil ") == ["aa", "b", "sil"]


def test_cmu_to_ipa_matches_reference_table():
    from conftest import jload
    from mddctc import textio
    ipa = jload("textio.json.gz")["ipa"]
    keys = sorted(ipa)
    assert textio.cmu_to_ipa(keys) == [ipa[k] for k in keys]
    assert len(textio.CMU_TO_IPA) == 41 and textio.IPA_TO_CMU["ə"] == "AH0"


def test_bucketed_batches_cover_and_bound_waste():
    import numpy as np
    from mddctc import textio
    rng = np.random.default_rng(0)
    lens = rng.integers(50, 1501, size=1000).tolist()
    batches = textio.bucketed_batches(lens, batch_size=32, width=64)
    assert sorted(i for b in batches for i in b) == list(range(1000))
    assert all(len(b) <= 32 for b in batches)
    assert all(max(lens[i] for i in b) - min(lens[i] for i in b) < 64 for b in batches)
    again = textio.bucketed_batches(lens, batch_size=32, width=64, shuffle_seed=3)
    assert sorted(map(tuple, again)) == sorted(map(tuple, batches)) and again != batches


def test_scoring_rates_match_reference_script(M):
    """FAR / FRR / DER (and the rest) against the text ins_del_sub_cor_analysis.py:193-219 printed on the same counters."""
    from conftest import jload
    from mddctc.corpus import CorpusMDD, scoring_rates, scoring_report
    for case in jload("scoring_rates.json.gz")["cases"]:
        assert scoring_report(case["counters"]) == case["text"]
    ev = CorpusMDD({0: "blank", 1: "a", 2: "b"})
    ev.totals[0][:5] = [90, 10, 5, 30, 15]                  # TA, FR, FA, correct diagnosis, wrong diagnosis
    r = ev.scoring()
    assert abs(r["FRR"] - 0.1) < 1e-12 and abs(r["FAR"] - 5 / 50) < 1e-12 and abs(r["DER"] - 15 / 45) < 1e-12
    assert r == scoring_rates({"cor_cor": 90, "cor_nocor": 10, "sub_nosub": 5, "sub_sub": 30, "sub_sub1": 15})
