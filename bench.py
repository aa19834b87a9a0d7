#!/usr/bin/env python
"""bench.py -- CTC fwd+bwd frames/s on B200 (BASELINE.json metric) with roofline, CPU baselines, end-to-end legs and
the decode / alignment stages the metric string names.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c3|c3peaky|c2|c4|c5] [--scaling weak|strong] [--no-extras] [--no-cpu]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path over one batch.

  c3       (default, the headline) BASELINE config 3: CTC loss + gradient, T=1000, B=4096, C=42, L=120, fp32, synthetic
           log_softmax(randn); one C-ABI call (mdd_ctc_loss) per step.  --scaling weak: every rank owns 4096 utterances;
           --scaling strong: the 4096 utterances are split over the ranks (BASELINE's wording).  The scalar loss is
           all-reduced over NCCL each step; there is no other collective.
  c3peaky  the same shape on trained-model-like posteriors (a valid alignment +16 nats, 10 % confidently wrong frames):
           the regime in which the packed kernel's second tier runs.
  c2       BASELINE config 2: T=200, B=32, L=30 (the reference's training shape).
  c4       BASELINE config 4: CTC/attention joint beam decode through the CTC prefix scorer, beam 10, CTC weight 0.3,
           B=1024, T=500 (utt/s); the reference's own joint decoder (prefix beam search + LM) is timed beside it.
  c5       BASELINE config 5: ragged T 50..1500, length-bucketed over the ranks, greedy decode + Levenshtein / MDD
           alignment end to end (utt/s).

Inputs + outputs of c3 are 1.376 GB per step, > 10x the 126 MB L2: no explicit L2 flush between timed iterations (the
small workloads c2 / c4 say what they do in `config.l2`).

  value      device-resident throughput: units of all ranks / max-over-ranks CUDA-event time
  e2e        the same metric through the public host-buffer call: HOST (pinned) inputs, H2D + kernels + D2H of the
             result inside the timed region
  roofline   algorithmic bytes / CUDA-event time of the step's kernels against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference   the reference's own CPU implementation of the path (torch CPU nn.CTCLoss for the
             loss; the C restatement of its decoder / alignment code for c4 / c5) on a bounded sample
  comparator torch's CUDA ctc_loss fwd+bwd on the same tensors (the existing GPU kernel to stay ahead of)
  extras     (N=1, default on) the other quantities of BASELINE.json's metric string, each with its CPU baseline
"""
import argparse
import collections
import hashlib
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (T, B per GPU (weak) or global (strong), C, L, seed, family)
    "c3": (1000, 4096, 42, 120, 1, "randn"),
    "c3peaky": (1000, 4096, 42, 120, 7, "aligned16"),
    "c2": (200, 32, 42, 30, 0, "randn"),
}
DESCR = {"c3": "BASELINE config 3: CTC loss fwd+bwd T=%d B=%d C=%d L=%d",
         "c3peaky": "BASELINE config 3 shape on peaky posteriors (valid alignment +16 nats, 10%% wrong frames): T=%d B=%d C=%d L=%d",
         "c2": "BASELINE config 2: CTC loss fwd+bwd T=%d B=%d C=%d L=%d"}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def aligned_logits(T, B, C, L, il, tl, tg, boost, noise, g):
    """Posteriors of a trained model: a random valid alignment of each target (labels on even frames, blanks between)
    gets +boost, a fraction `noise` of the frames is confidently wrong.  Works on the generator's device."""
    dev = tg.device
    x = torch.randn(T, B, C, generator=g, device=dev)
    half = (il // 2).clamp(min=1)
    key = torch.rand(B, T // 2, generator=g, device=dev)
    key = key + (torch.arange(T // 2, device=dev)[None, :] >= half[:, None]) * 2.0
    pos = key.argsort(dim=1)[:, :L].sort(dim=1)[0] * 2
    lab = torch.zeros(B, T, dtype=torch.long, device=dev)
    valid = torch.arange(L, device=dev)[None, :] < tl[:, None]
    lab.scatter_(1, torch.where(valid, pos, torch.full_like(pos, T - 1)), torch.where(valid, tg, torch.zeros_like(tg)))
    lab = lab.t().contiguous()
    flip = torch.rand(T, B, generator=g, device=dev) < noise
    lab = torch.where(flip, torch.randint(0, C, (T, B), generator=g, device=dev), lab)
    x.scatter_add_(2, lab[..., None], torch.full((T, B, 1), float(boost), device=dev))
    return x


def make_inputs(name, rank=0, b_local=None, pinned=False, slices=1):
    """CPU-generated so that every arm sees identical numbers.  Returns (lp slices [list of (T,B_k,C)], tg, il, tl)."""
    T, B, C, L, seed, fam = WORKLOADS[name]
    B = b_local or B
    g = torch.Generator().manual_seed(seed + 1000 * rank)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.full((B,), T, dtype=torch.int64)
    tl = torch.full((B,), L, dtype=torch.int64)
    base, rem = divmod(B, slices)
    out, lo = [], 0
    for k in range(slices):
        bk = base + (1 if k < rem else 0)
        lp = torch.empty((T, bk, C), dtype=torch.float32, pin_memory=pinned)
        if fam == "randn":
            step = 125                                       # in pieces: bounds the peak host memory
            for t0 in range(0, T, step):
                lp[t0:t0 + step] = torch.log_softmax(torch.randn(min(step, T - t0), bk, C, generator=g), -1)
        else:
            x = aligned_logits(T, bk, C, L, il[lo:lo + bk], tl[lo:lo + bk], tg[lo:lo + bk], 16.0, 0.1, g)
            lp.copy_(torch.log_softmax(x, -1))
        out.append(lp)
        lo += bk
    return out, tg, il, tl


class ClockSampler:
    """Samples SM clock + throttle reasons during the timed region (pynvml)."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    INFO = {"sw_power_cap": 0x4, "hw_power_brake": 0x80, "sync_boost": 0x10}

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in {**self.BAD, **self.INFO}.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_source_hash():
    """sha1 over the CTC kernel sources: an ncu traffic capture is only valid for the sources it was taken from."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "ctc-attention-mispronunciation_b200", "csrc")
    for f in ("ctc_pair.cu", "ctc_lattice.cuh", "ctc_params.cuh", "ctc_exact.cu", "capi.cu", "common.cuh"):
        p = os.path.join(d, f)
        if os.path.exists(p):
            h.update(open(p, "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(name, algo):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed ncu --set full
    capture of this workload (profiles/ncu_traffic.json) -- None when there is no capture, or when the kernel sources
    have changed since it was taken (the entry carries the source hash of its capture)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        e = json.load(open(p)).get("%s/algo%d" % (name, algo), {})
        if e.get("src_sha") != kernel_source_hash():
            return None
        return e.get("bytes")
    except Exception:
        return None


def time_cuda(fn, steps, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


# ----------------------------------------------------------------------------------------------- CPU reference arms
def cpu_ctc_arm(name, steps, warmup, slice_b=None):
    """torch CPU ctc_loss fwd+bwd (the reference's implementation of the path, nn.CTCLoss at steps/train_ctc.py:186) on a
    bounded slice, all host threads."""
    T, B, C, L, seed, fam = WORKLOADS[name]
    sb = min(B, slice_b or (256 if T >= 1000 else B))
    lps, tg, il, tl = make_inputs(name, b_local=sb)
    lp = lps[0]
    torch.set_num_threads(max(host_threads(), 1))          # torchrun exports OMP_NUM_THREADS=1, which would hide the cores
    threads = torch.get_num_threads()
    loss_fn = torch.nn.CTCLoss(reduction="sum")

    def step():
        x = lp.clone().requires_grad_(True)          # the clone stands for the model producing `out`; not timed
        t0 = time.perf_counter()
        loss = loss_fn(x, tg, il, tl)
        loss = loss / sb
        loss.backward()
        return time.perf_counter() - t0

    for _ in range(warmup):
        step()
    dt = [step() for _ in range(steps)]
    sec = sum(dt) / len(dt)
    return {"value": float(il.sum()) / sec, "unit": "frames/s", "cores": threads, "kind": "reference",
            "sample": "torch %s CPU nn.CTCLoss(reduction='sum') fwd+bwd fp32 on a B=%d slice of the %s batch "
                      "(T=%d, C=%d, L=%d), %d threads, mean of %d steps" % (torch.__version__, sb, name, T, C, L, threads, steps),
            "ms_per_step": sec * 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    if name in ("c4", "c5"):
        r = cpu_decode_arm(name)
        metric = "joint_decode_utt_per_s" if name == "c4" else "greedy_mdd_utt_per_s"
        workload = C4_DESCR if name == "c4" else C5_DESCR
    else:
        T, B, C, L, _, _ = WORKLOADS[name]
        r = cpu_ctc_arm(name, args.steps, args.warmup)
        metric = "ctc_fwd_bwd_frames_per_s"
        workload = (DESCR[name] % (T, B, C, L)) + " per GPU"
    line = {"impl": "reference", "metric": metric, "value": r["value"], "unit": r["unit"],
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r.get("ms_per_step"),
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32" if name not in ("c4",) else "f64", "data": "synthetic",
            "config": {"workload": workload, "sample": r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": r["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- decode stages
C4_DESCR = "BASELINE config 4: CTC-attention joint beam decode (beam 10, CTC weight 0.3) prefix scoring, B=1024, T=500, C=43"
C5_DESCR = "BASELINE config 5: ragged bucketed batch T 50-1500, greedy decode + Levenshtein MDD alignment, C=43"
C4 = dict(B=1024, T=500, C=43, W=10, steps=32, lam=0.3, eos=42)
C5 = dict(B=2048, T=1500, C=43, Lc=188)


def phone_vocab():
    vocab = {0: "blank", 1: "UNK"}
    for line in open(os.path.join(ROOT, "tests", "golden", "assets", "units")):
        if line.strip():
            vocab[len(vocab)] = line.strip()
    return vocab


def c4_inputs(B, T, C, seed, dev):
    """Posteriors with a blank floor and label spikes (what a CTC model emits), ragged lengths 300..T."""
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.randn(T, B, C, device=dev, generator=g)
    u = torch.rand(T, B, device=dev, generator=g)
    k = torch.randint(1, C - 1, (T, B), device=dev, generator=g)
    x[:, :, 0] += 8.0 * (u < 0.7)
    x.scatter_add_(2, k[..., None], (6.0 * (u >= 0.7))[..., None].float())
    lens = torch.randint(300, T + 1, (B,), device=dev, generator=g)
    return torch.log_softmax(x, -1), lens


def synthetic_attention(C, periods=7, seed=0):
    """Stand-in attention decoder (the reference has none): att[b,w,:] = table[step % periods, last_label + 1]."""
    g = torch.Generator().manual_seed(seed)
    return torch.log_softmax(torch.randn(periods, C + 1, C, generator=g) * 2.0, -1)


def c5_inputs(B, T, C, Lc, seed, dev, lens=None):
    g = torch.Generator(device=dev).manual_seed(seed)
    if lens is None:
        lens = torch.randint(50, T + 1, (B,), device=dev, generator=g)
    x = torch.randn(T, B, C, device=dev, generator=g)
    x.scatter_add_(2, torch.randint(0, C, (T, B, 1), device=dev, generator=g), torch.full((T, B, 1), 4.0, device=dev))
    lp = torch.log_softmax(x, -1)
    can = torch.randint(1, C, (B, Lc), device=dev, generator=g, dtype=torch.int32)     # includes sil / err tokens
    can_len = torch.clamp((lens + 7) // 8, max=Lc).to(torch.int32)
    return lp, lens, can, can_len


def cpu_decode_arm(name):
    """The reference's decoder / alignment code on the host cores, through its C restatement (oracle/, the checker):
    c4 = ctcBeamSearch.decode (BeamSearch.py:73-153, the reference's joint decoder) on a 32-utterance slice, 1 core;
    c5 = GreedyDecoder.decode -> wer -> align_canonical_decoded (ctcDecoder.py:186-200,118-184, infer.py:162-209) on a
    512-utterance slice, 1 core."""
    import numpy as np
    from oracle import api as O
    vocab = phone_vocab()
    if name == "c4":
        lp, lens = c4_inputs(32, C4["T"], C4["C"], 4, torch.device("cpu"))
        probs = torch.exp(lp.transpose(0, 1)).contiguous().numpy()
        lm = O.lm_table(os.path.join(ROOT, "tests", "golden", "assets", "lm_phone_bg.arpa"), vocab, C4["C"])
        t0 = time.perf_counter()
        O.beam_tokens(probs, lens.numpy(), 10, 0, lm, 0.3)
        sec = time.perf_counter() - t0
        return {"value": 32 / sec, "unit": "utt/s", "cores": 1, "kind": "port", "ms_per_step": sec * 1e3,
                "sample": "C restatement of ctcBeamSearch.decode (beam 10, lm_alpha 0.3) on 32 utterances of the c4 batch, 1 core"}
    n = 512
    lp, lens, can, can_len = c5_inputs(n, C5["T"], C5["C"], C5["Lc"], 5, torch.device("cpu"))
    lpn, ln = lp.numpy(), lens.numpy()
    from mddctc.mdd import MDDScorer
    sc = MDDScorer(vocab)
    t0 = time.perf_counter()
    tokens, out_len, _ = O.greedy_tokens(lpn, ln, 0, O.word_ids(vocab))
    dec = [[int(sc.dec_map[k]) for k in tokens[b, :out_len[b]] if sc.dec_map[k] >= 0] for b in range(n)]
    cn = [[int(sc.can_map[k]) for k in can[b, :can_len[b]].tolist() if sc.can_map[k] >= 0] for b in range(n)]
    dist, paths, st = O.edit_align(dec, cn)
    for b in range(n):
        if not st[b]:
            O.mdd_post(dec[b], cn[b], paths[b])
    sec = time.perf_counter() - t0
    return {"value": n / sec, "unit": "utt/s", "cores": 1, "kind": "port", "ms_per_step": sec * 1e3,
            "sample": "C restatement of GreedyDecoder.decode + Decoder.wer + align_canonical_decoded/stastics on 512 "
                      "utterances of the ragged c5 batch (T 50-1500), 1 core"}


def run_c4(dev, steps=3, cpu=True):
    """Joint decode at the config-4 shape: (a) the CTC prefix scorer + attention fusion (north_star, a10), (b) the
    reference's own joint decoder, prefix beam search with LM fusion (a6)."""
    from mddctc.decoder import BeamDecoder
    from mddctc.prefix import joint_beam_search
    import mddctc
    L_ = mddctc._lib.lib()
    B, T, C, W, nlab = C4["B"], C4["T"], C4["C"], C4["W"], C4["steps"]
    lp, lens = c4_inputs(B, T, C, 4, dev)
    table = synthetic_attention(C).to(dev)

    def att(step, st):
        return table[step % table.shape[0]][(st["last"] + 1).long()]

    def joint():
        return joint_beam_search(lp, lens, att, beam=W, ctc_weight=C4["lam"], max_len=nlab, eos=C4["eos"])

    L_.mdd_launch_count_reset()
    ms = time_cuda(joint, steps, 1)
    launches = int(L_.mdd_launch_count()) // (steps + 1)
    out = {"prefix_scorer": {"value": B / (ms * 1e-3), "unit": "utt/s", "ms_per_batch": ms, "label_steps": nlab,
                             "gpu_launches_per_batch": launches,
                             "config": "beam 10, CTC weight 0.3, B=1024, T=500 (ragged 300-500), C=43, %d label steps, eos on, "
                                       "synthetic attention log-probs; fp64 scores" % nlab}}
    # host-buffer leg: pinned log-probs in, tokens + scores out
    lp_h = lp.cpu().pin_memory()
    lens_h = lens.cpu().pin_memory()
    lp_d, lens_d = torch.empty_like(lp), torch.empty_like(lens)

    def joint_e2e():
        lp_d.copy_(lp_h, non_blocking=True)
        lens_d.copy_(lens_h, non_blocking=True)
        tok, sc, gl, fl = joint_beam_search(lp_d, lens_d, att, beam=W, ctc_weight=C4["lam"], max_len=nlab, eos=C4["eos"])
        return tok.cpu(), sc.cpu(), gl.cpu()

    ms2 = time_cuda(joint_e2e, max(1, steps - 1), 1)
    out["prefix_scorer"]["e2e"] = {"value": B / (ms2 * 1e-3), "unit": "utt/s", "h2d_bytes_per_step": lp_h.numel() * 4 + B * 8,
                                   "d2h_bytes_per_step": B * W * (nlab * 4 + 8 + 4), "ms_per_batch": ms2}
    # the reference's own joint decoder (a6)
    arpa = os.path.join(ROOT, "tests", "golden", "assets", "lm_phone_bg.arpa")
    dec = BeamDecoder(phone_vocab(), beam_width=10, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=0.3)
    probs = torch.exp(lp.transpose(0, 1)).contiguous()
    ms3 = time_cuda(lambda: dec.decode_probs_tokens(probs, lens), 5, 2)
    out["prefix_beam_search_lm"] = {"value": B / (ms3 * 1e-3), "unit": "utt/s", "ms_per_batch": ms3,
                                    "config": "BeamDecoder beam 10, lm_alpha 0.3, same batch (device-resident probabilities)"}
    if cpu:
        out["cpu_baseline"] = cpu_decode_arm("c4")
        out["cpu_baseline_prefix_scorer"] = cpu_joint_arm()
    return out


def cpu_joint_arm():
    """fp64 C restatement of the joint CTC/attention beam step on a 16-utterance slice, up to 16 host threads."""
    from oracle import api as O
    B, T, C, W, nlab = 16, C4["T"], C4["C"], C4["W"], 8
    lp, lens = c4_inputs(B, T, C, 4, torch.device("cpu"))
    table = synthetic_attention(C)

    def att(step, st):
        return table[step % table.shape[0]][torch.from_numpy(st["last"] + 1).long()].numpy()

    t0 = time.perf_counter()
    O.joint_beam_search(lp.numpy(), lens.numpy(), att, beam=W, ctc_weight=C4["lam"], max_len=nlab, eos=C4["eos"])
    sec = time.perf_counter() - t0
    per_full = sec * C4["steps"] / nlab                     # the search is linear in the number of label steps
    return {"value": B / per_full, "unit": "utt/s", "cores": min(16, host_threads()), "kind": "port",
            "sample": "fp64 C restatement of the joint beam step, 16 utterances x %d of %d label steps (scaled), <= 16 threads"
                      % (nlab, C4["steps"])}


def run_c5(dev, world=1, rank=0, steps=5, cpu=True, b_global=None):
    """Greedy decode + MDD alignment on a ragged batch; with world > 1 the utterances are dealt to the ranks by
    mddctc.dist.balanced_buckets (length-bucketed, equal total frames)."""
    from mddctc.decoder import GreedyDecoder
    from mddctc.dist import balanced_buckets
    from mddctc.mdd import MDDScorer
    import mddctc
    L_ = mddctc._lib.lib()
    vocab = phone_vocab()
    Bg = b_global or C5["B"] * world
    T, C, Lc = C5["T"], C5["C"], C5["Lc"]
    g = torch.Generator().manual_seed(55)
    all_lens = torch.randint(50, T + 1, (Bg,), generator=g)
    mine = balanced_buckets(all_lens.tolist(), world)[rank]
    lens = all_lens[mine].to(dev)
    B = len(mine)
    lp, lens, can, can_len = c5_inputs(B, T, C, Lc, 5 + rank, dev, lens=lens)
    gd = GreedyDecoder(vocab, space_idx=-1, blank_index=0)
    sc = MDDScorer(vocab)

    def pipeline(x=lp, ln=lens):
        tokens, out_len, _ = gd.decode_tokens(x, ln)
        return sc.align_tokens(tokens, out_len, can, can_len, max_dec=T, max_can=Lc)

    L_.mdd_launch_count_reset()
    ms = time_cuda(pipeline, steps, 2)
    launches = int(L_.mdd_launch_count()) // (steps + 2)
    frames = int(lens.sum())
    # host-buffer leg: pinned log-probs in, integer results out (+ the host-side string rendering, timed separately)
    lp_h = lp.cpu().pin_memory()
    lp_d = torch.empty_like(lp)
    keys = ("dist", "status", "ops", "ops_len", "a_dec", "a_can", "a_ops", "ins", "sub", "del", "counts")

    def e2e():
        lp_d.copy_(lp_h, non_blocking=True)
        r = pipeline(lp_d)
        return {k: r[k].cpu() for k in keys if k in r}

    ms2 = time_cuda(e2e, max(2, steps - 2), 1)
    host = e2e()
    d2h = sum(v.numel() * v.element_size() for v in host.values())
    t0 = time.perf_counter()
    r = pipeline()
    rendered = sc.to_python(r)
    ms3 = (time.perf_counter() - t0) * 1e3
    out = {"value_local": B / (ms * 1e-3), "ms_per_batch": ms, "utterances_local": B, "frames_local": frames,
           "gpu_launches_per_batch": launches,
           "e2e_local": {"value": B / (ms2 * 1e-3), "unit": "utt/s", "ms_per_batch": ms2,
                         "h2d_bytes_per_step": lp_h.numel() * 4, "d2h_bytes_per_step": d2h},
           "with_host_strings_ms": ms3, "rendered": len(rendered),
           "algorithmic_bytes": 4.0 * C * T * B + B * T,       # the arg-max stream reads every padded row once
           }
    if cpu and rank == 0:
        out["cpu_baseline"] = cpu_decode_arm("c5")
    return out


# ----------------------------------------------------------------------------------------------- the CTC loss arm
def run_ours(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import mddctc
    from mddctc import _rt
    numa = _rt.bind_to_gpu_numa(local) if world > 1 else None   # before any pinned allocation
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from mddctc.loss import ctc_loss_raw, ctc_loss_streamed, last_fallback_flags
    L_ = mddctc._lib.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def allsum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t[0])

    name = args.workload
    if name in ("c4", "c5"):
        return run_decode_line(args, name, dev, world, rank, barrier, allmax, allsum)

    T, Bw, C, L, _, fam = WORKLOADS[name]
    if args.scaling == "strong":
        from mddctc.dist import shard_bounds
        lo, hi = shard_bounds(Bw, world, rank)
        B = hi - lo
    else:
        B = Bw
    K = max(1, min(args.e2e_slices, B))
    lp_hs, tg_h, il_h, tl_h = make_inputs(name, rank, b_local=B, pinned=True, slices=K)
    tg_h, il_h, tl_h = tg_h.pin_memory(), il_h.pin_memory(), tl_h.pin_memory()
    lp = torch.cat([h.to(dev) for h in lp_hs], dim=1).contiguous()
    tg, il, tl = tg_h.to(dev), il_h.to(dev), tl_h.to(dev)
    nll = torch.empty(B, dtype=torch.float32, device=dev)
    grad = torch.empty((T, B, C), dtype=torch.float32, device=dev)
    status = torch.empty(B, dtype=torch.int32, device=dev)
    loss_buf = torch.zeros(1, dtype=torch.float32, device=dev)
    frames_local = float(il_h.sum())
    frames_global = allsum(frames_local)
    inv_b = 1.0 / (Bw if args.scaling == "strong" else B)     # the reference divides the summed loss by the batch size

    # The scalar loss is the only thing exchanged.  Its all-reduce runs on NCCL's own stream and overlaps the NEXT step's
    # kernels: step k's collective is waited for (stream-side, no host block) two steps later, and the timed region ends
    # only after every pending collective has been joined.  A ring of four result buffers keeps the steps apart.
    loss_ring = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(4)]
    pending = collections.deque()
    step_no = [0]

    def join_collectives(keep=0):
        while len(pending) > keep:
            pending.popleft().wait()

    def step_resident():
        buf = loss_ring[step_no[0] % 4]
        step_no[0] += 1
        ctc_loss_raw(lp, tg, il, tl, want_grad=True, grad_scalar=inv_b, validate=False, loss_sum=buf, nll=nll,
                     grad=grad, status=status, algo=args.algo)
        if world > 1:
            pending.append(dist.all_reduce(buf, async_op=True))
            join_collectives(keep=2)

    tiers = None
    if args.algo == 0:
        ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=status, algo=2)
        n1 = int((last_fallback_flags(B, dev) != 0).sum())
        step_resident()
        n3 = int((last_fallback_flags(B, dev) != 0).sum())
        tiers = {"flagged_by_first_pass": n1, "left_for_fp64_kernel": n3, "utterances": B}
    W = max(args.warmup, 3)
    for _ in range(W):
        step_resident()
    join_collectives()
    barrier()
    L_.mdd_launch_count_reset()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        barrier()
        t_wall = time.perf_counter()
        for a, b in ev:
            a.record()
            step_resident()
            if b is ev[-1][1]:
                join_collectives()                       # the last step's collective ends inside the timed region
            b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall
    launches = int(L_.mdd_launch_count())
    kern_ms = allmax(sum(a.elapsed_time(b) for a, b in ev))
    total_ms = allmax(ev[0][0].elapsed_time(ev[-1][1]))
    ms_per_step = total_ms / args.steps
    value = frames_global / (ms_per_step * 1e-3)
    loss_val = float(loss_ring[(step_no[0] - 1) % 4])

    # ---- end to end through the public host-buffer call: slice-major pinned log-probs (one contiguous H2D per slice on
    # a side stream, slice k+1 crossing PCIe while slice k computes), kernels, D2H of the loss
    lp_ds = [torch.empty(h.shape, dtype=torch.float32, device=dev) for h in lp_hs]
    tg_d, il_d, tl_d = torch.empty_like(tg), torch.empty_like(il), torch.empty_like(tl)
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    def step_e2e():
        tg_d.copy_(tg_h, non_blocking=True)
        il_d.copy_(il_h, non_blocking=True)
        tl_d.copy_(tl_h, non_blocking=True)
        total, _, _ = ctc_loss_streamed(lp_hs, tg_d, il_d, tl_d, grad_scalar=inv_b, staging=lp_ds, algo=args.algo)
        loss_buf.copy_(total)
        if world > 1:
            dist.all_reduce(loss_buf)
        loss_host.copy_(loss_buf, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(loss_host)

    n_e2e = max(2, min(args.steps, 10))
    step_e2e()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_e2e):
        step_e2e()
    e1.record()
    barrier()
    e2e_ms = allmax(e0.elapsed_time(e1) / n_e2e)
    e2e_value = frames_global / (e2e_ms * 1e-3)
    h2d = sum(h.numel() for h in lp_hs) * 4 + tg_h.numel() * 8 + il_h.numel() * 8 + tl_h.numel() * 8

    # the ceiling of that leg on this box: the same pinned bytes as plain contiguous copies, all ranks at once
    def h2d_only():
        for h, d in zip(lp_hs, lp_ds):
            d.copy_(h, non_blocking=True)
    h2d_only()
    barrier()
    e0.record()
    for _ in range(3):
        h2d_only()
    e1.record()
    barrier()
    h2d_ms = allmax(e0.elapsed_time(e1) / 3)
    h2d_gbs_rank = sum(h.numel() for h in lp_hs) * 4 / (h2d_ms * 1e-3) / 1e9

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg_bytes = 4.0 * C * frames_local + 4.0 * C * T * B            # per launch (one rank)
        kernel_ms = kern_ms / args.steps
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
        line = {"metric": "ctc_fwd_bwd_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world,
                "steps": args.steps, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": (DESCR[name] % (T, Bw, C, L)) + (" per GPU" if args.scaling == "weak" else " in total, split over the GPUs"),
                           "global_batch": int(B * world if args.scaling == "weak" else Bw), "batch_per_gpu": B,
                           "sharding": "dp%d, batch-sharded, scalar loss all-reduce (NCCL, on its own stream, overlapping the next step; "
                                       "all joined inside the timed region)" % world,
                           "l2": "inputs+outputs %.2f GB per step per GPU vs 126 MB L2, no explicit flush" % (alg_bytes / 1e9),
                           "algo": args.algo, "loss": loss_val, "tiers": tiers, "numa_binding": numa},
                "clocks": clk.summary(),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic(name, args.algo), "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": kernel_ms,
                             "note": "CUDA-event time of one mdd_ctc_loss call (all tiers + loss sum); traffic = dram bytes "
                                     "of the ncu capture whose source hash matches this build, else null"},
                "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                        "ms_per_step": e2e_ms, "slices": len(lp_hs),
                        "h2d_ceiling_ms": h2d_ms, "h2d_ceiling_gbs_per_rank": h2d_gbs_rank,
                        "frac_of_h2d_ceiling": h2d_ms / e2e_ms,
                        "note": "host layout slice-major: %d contiguous pinned (T,B/%d,C) tensors, one cudaMemcpyAsync each"
                                % (len(lp_hs), len(lp_hs))},
                "gpu_launches": launches, "wall_s_timed_region": t_wall}
        if world == 1 and not args.no_comparator:
            line["comparator"] = torch_cuda_comparator(lp, tg, il, tl, frames_local)
        if not args.no_cpu and world == 1:
            r = cpu_ctc_arm(name, 3, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        elif not args.no_cpu:
            line["cpu_baseline"] = None
        if world == 1 and not args.no_extras and name == "c3":
            del lp, grad, lp_ds
            torch.cuda.empty_cache()
            line["extras"] = run_extras(dev, cpu=not args.no_cpu)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def torch_cuda_comparator(lp, tg, il, tl, frames):
    """torch's own CUDA ctc_loss fwd+bwd on the same device tensors (not ours; the kernel to stay ahead of)."""
    x = lp.detach().clone().requires_grad_(True)

    def step():
        x.grad = None
        loss = torch.nn.functional.ctc_loss(x, tg, il, tl, reduction="sum")
        loss.backward()

    ms = time_cuda(step, 3, 1)
    return {"impl": "torch %s CUDA F.ctc_loss fwd+bwd (aten native kernel)" % torch.__version__, "ms_per_step": ms,
            "value": frames / (ms * 1e-3), "unit": "frames/s"}


def run_small_ctc(dev, name, cpu=True, calls=200):
    """A small CTC workload (c2) device-resident: mean time of one mdd_ctc_loss call, next to torch CUDA and torch CPU.
    The batch (0.5 MB) lives in L2 between calls -- as it does in training, where the model's output layer has just
    written it."""
    from mddctc.loss import ctc_loss_raw
    T, B, C, L, _, _ = WORKLOADS[name]
    lps, tg, il, tl = make_inputs(name)
    lp, tg, il, tl = lps[0].to(dev), tg.to(dev), il.to(dev), tl.to(dev)
    nll = torch.empty(B, device=dev)
    grad = torch.empty((T, B, C), device=dev)
    st = torch.empty(B, dtype=torch.int32, device=dev)
    ms = time_cuda(lambda: ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st), calls, 10)
    frames = float(il.sum())
    out = {"ms_per_call": ms, "value": frames / (ms * 1e-3), "unit": "frames/s",
           "config": (DESCR[name] % (T, B, C, L)) + "; batch resident in L2 between calls",
           "comparator": torch_cuda_comparator(lp, tg, il, tl, frames)}
    if cpu:
        r = cpu_ctc_arm(name, 5, 2)
        out["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    return out


def run_peaky(dev, steps=5):
    """c3peaky device-resident: one GPU, per-tier counts (inputs generated on the device)."""
    from mddctc.loss import ctc_loss_raw, last_fallback_flags
    T, B, C, L, seed, _ = WORKLOADS["c3peaky"]
    g = torch.Generator(device=dev).manual_seed(seed)
    tg = torch.randint(1, C, (B, L), generator=g, device=dev)
    il = torch.full((B,), T, dtype=torch.int64, device=dev)
    tl = torch.full((B,), L, dtype=torch.int64, device=dev)
    lp = torch.empty((T, B, C), device=dev)
    for lo in range(0, B, 1024):
        x = aligned_logits(T, 1024, C, L, il[lo:lo + 1024], tl[lo:lo + 1024], tg[lo:lo + 1024], 16.0, 0.1, g)
        lp[:, lo:lo + 1024] = torch.log_softmax(x, -1)
        del x
    nll = torch.empty(B, device=dev)
    grad = torch.empty((T, B, C), device=dev)
    st = torch.empty(B, dtype=torch.int32, device=dev)
    ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st, algo=2)
    n1 = int((last_fallback_flags(B, dev) != 0).sum())
    ms = time_cuda(lambda: ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st), steps, 2)
    n3 = int((last_fallback_flags(B, dev) != 0).sum())
    return {"ms_per_step": ms, "value": T * B / (ms * 1e-3), "unit": "frames/s",
            "tiers": {"flagged_by_first_pass": n1, "left_for_fp64_kernel": n3, "utterances": B},
            "config": DESCR["c3peaky"] % (T, B, C, L)}


def run_extras(dev, cpu=True):
    """The other quantities of BASELINE.json's metric string, each next to its CPU baseline (cores stated)."""
    out = {}
    for key, fn in (("c2", lambda: run_small_ctc(dev, "c2", cpu)), ("c3peaky", lambda: run_peaky(dev)),
                    ("joint_decode", lambda: run_c4(dev, cpu=cpu)), ("greedy_mdd", lambda: extras_c5(dev, cpu)),
                    ("ctc_from_logits", lambda: run_logits_extra(dev))):
        try:
            out[key] = fn()
        except Exception as e:          # noqa
            out[key] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        torch.cuda.empty_cache()
    return out


def extras_c5(dev, cpu):
    r = run_c5(dev, cpu=cpu)
    return {"value": r["value_local"], "unit": "utt/s", "ms_per_batch": r["ms_per_batch"], "e2e": r["e2e_local"],
            "with_host_strings_ms": r["with_host_strings_ms"], "cpu_baseline": r.get("cpu_baseline"),
            "config": "ragged T 50-1500, B=%d (one GPU's share), C=43, greedy decode + Levenshtein/MDD alignment" % r["utterances_local"]}


def run_decode_line(args, name, dev, world, rank, barrier, allmax, allsum):
    """--workload c4 / c5 as stand-alone bench lines (utt/s)."""
    import torch.distributed as dist
    peak, peak_src = measured_peak_gbs()
    if name == "c4":
        r = run_c4(dev, steps=max(2, min(args.steps, 5)), cpu=(rank == 0 and not args.no_cpu))
        ps = r["prefix_scorer"]
        ms = allmax(ps["ms_per_batch"])
        B, T, C, W, nl = C4["B"], C4["T"], C4["C"], C4["W"], C4["steps"]
        # per label step: score reads the parents' state and the rows once, advance reads + writes a state
        alg = nl * B * (3.0 * W * T * 16 + T * C * 4)
        line = {"metric": "joint_decode_utt_per_s", "value": B * world / (ms * 1e-3), "unit": "utt/s", "ms_per_step": ms,
                "dtype": "f64", "config": {"workload": C4_DESCR, "detail": ps["config"], "utterances_per_gpu": B,
                                           "l2": "state + rows 0.25 GB per label step > 126 MB L2"},
                "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": alg / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                             "note": "latency-bound scan over T in fp64 log space (3 log-sum-exp per frame and candidate)"},
                "e2e": ps["e2e"], "gpu_launches": ps["gpu_launches_per_batch"],
                "also": {"prefix_beam_search_lm": r["prefix_beam_search_lm"]},
                "cpu_baseline": r.get("cpu_baseline_prefix_scorer"), "cpu_baseline_reference_decoder": r.get("cpu_baseline")}
    else:
        r = run_c5(dev, world, rank, steps=max(3, min(args.steps, 10)), cpu=not args.no_cpu)
        ms = allmax(r["ms_per_batch"])
        utts = allsum(r["utterances_local"])
        e2e_ms = allmax(r["e2e_local"]["ms_per_batch"])
        alg = r["algorithmic_bytes"]
        line = {"metric": "greedy_mdd_utt_per_s", "value": utts / (ms * 1e-3), "unit": "utt/s", "ms_per_step": ms,
                "dtype": "u8/i32", "config": {"workload": C5_DESCR, "utterances": int(utts), "utterances_rank0": r["utterances_local"],
                                              "frames_rank0": r["frames_local"],
                                              "sharding": "mddctc.dist.balanced_buckets over %d ranks, no collective" % world,
                                              "l2": "log-probs %.2f GB per rank vs 126 MB L2" % (alg / 1e9)},
                "roofline": {"bound": "hbm", "achieved": alg / (r["ms_per_batch"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": alg / (r["ms_per_batch"] * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                             "note": "whole pipeline (arg-max stream + collapse + alignment) against the arg-max stream's bytes"},
                "e2e": {"value": utts / (e2e_ms * 1e-3), "unit": "utt/s", "ms_per_step": e2e_ms,
                        "h2d_bytes_per_step": r["e2e_local"]["h2d_bytes_per_step"],
                        "d2h_bytes_per_step": r["e2e_local"]["d2h_bytes_per_step"]},
                "gpu_launches": r["gpu_launches_per_batch"], "with_host_strings_ms": r["with_host_strings_ms"],
                "cpu_baseline": r.get("cpu_baseline")}
    if rank == 0:
        line.update({"n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "higher_is_better": True,
                     "scaling": "weak", "vs_baseline": None, "data": "synthetic"})
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_logits_extra(dev, steps=10):
    """SURVEY 8f rank 1 at the config-3 shape: logits in, logit gradient out.  Fused = mdd_ctc_loss_logits (row-LSE
    pre-pass + CTC kernel); unfused = torch log_softmax -> mdd_ctc_loss -> torch log_softmax backward on the same GPU."""
    from mddctc.loss import ctc_loss_raw, row_lse
    T, B, C, L = 1000, 4096, 42, 120
    g = torch.Generator(device=dev).manual_seed(5)
    x = torch.randn(T, B, C, device=dev, generator=g) * 2
    tg = torch.randint(1, C, (B, L), device=dev, generator=g)
    il = torch.full((B,), T, dtype=torch.int64, device=dev)
    tl = torch.full((B,), L, dtype=torch.int64, device=dev)
    nll = torch.empty(B, device=dev)
    grad = torch.empty((T, B, C), device=dev)
    st = torch.empty(B, dtype=torch.int32, device=dev)
    lse = torch.empty((T, B), device=dev)
    am = torch.empty((T, B), dtype=torch.uint8, device=dev)

    def fused():
        ctc_loss_raw(x, tg, il, tl, validate=False, nll=nll, grad=grad, status=st, logits=True, row_lse=lse, argmax=am)

    def unfused():
        lp = torch.log_softmax(x, -1)
        ctc_loss_raw(lp, tg, il, tl, validate=False, nll=nll, grad=grad, status=st)
        return torch.ops.aten._log_softmax_backward_data(grad, lp, -1, torch.float32)

    def prepass():
        row_lse(x, want_argmax=True)

    res = {}
    for name, fn in (("fused_ms", fused), ("unfused_ms", unfused), ("row_lse_prepass_ms", prepass)):
        res[name] = time_cuda(fn, steps, 3)
    res["fused_frames_per_s"] = T * B / (res["fused_ms"] * 1e-3)
    res["row_lse_prepass_gbs"] = (T * B * (4 * C + 5)) / (res["row_lse_prepass_ms"] * 1e-3) / 1e9
    res["config"] = "config-3 shape, logits (T,B,C) in, logit gradient out; unfused = torch log_softmax + mdd_ctc_loss + torch log_softmax backward"
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS) + ["c4", "c5"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank owns the workload's batch; strong: the batch is split over the ranks")
    ap.add_argument("--algo", type=int, default=0)
    ap.add_argument("--e2e-slices", type=int, default=4,
                    help="batch slices of the end-to-end leg (H2D of slice k+1 overlaps the kernels of slice k)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-extras", action="store_true", help="skip the decode / small-batch / peaky extras (N=1 only)")
    ap.add_argument("--no-comparator", action="store_true", help="skip torch's CUDA ctc_loss")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
