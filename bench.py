#!/usr/bin/env python
"""bench.py -- CTC fwd+bwd frames/s on B200 (BASELINE.json metric), with roofline, CPU baseline and end-to-end arms.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2] [--extras]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path over one batch: loss + gradient of the whole batch in one C-ABI call
(mdd_ctc_loss).  Workload at every N: BASELINE config 3 per GPU (T=1000, B=4096, C=42, L=120, full lengths, fp32,
synthetic log_softmax(randn)); N>1 is weak scaling (each rank owns a 4096-utterance shard, the scalar loss is
all-reduced over NCCL each step; no other collective).  The 1.376 GB of log-probs + gradients per step is > 10x the
126 MB L2, so no explicit L2 flush is needed between timed iterations.

  value      device-resident throughput: frames of all ranks / max-over-ranks CUDA-event time
  e2e        same metric through the Python CTCLoss drop-in path with HOST (pinned) inputs: H2D of log-probs, targets and
             lengths + kernel + D2H of the loss inside the timed region
  roofline   algorithmic bytes (4*C*sum(T_b) read + 4*C*T*B written) / CUDA-event time of the step's kernels, against
             MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference   torch CPU ctc_loss fwd+bwd (the reference's own implementation of this path:
             nn.CTCLoss at steps/train_ctc.py:186) with all host threads on a bounded slice of the same batch
"""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (T, B per GPU, C, L, seed)
    "c3": (1000, 4096, 42, 120, 1),
    "c2": (200, 32, 42, 30, 0),
}


def make_inputs(name, rank=0, device=None, pinned=False):
    T, B, C, L, seed = WORKLOADS[name]
    g = torch.Generator().manual_seed(seed + 1000 * rank)
    # generated in pieces to bound the peak host memory, on the CPU so that every arm sees identical numbers
    lp = torch.empty((T, B, C), dtype=torch.float32, pin_memory=pinned)
    step = 125
    for t0 in range(0, T, step):
        lp[t0:t0 + step] = torch.log_softmax(torch.randn(min(step, T - t0), B, C, generator=g), -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.full((B,), T, dtype=torch.int64)
    tl = torch.full((B,), L, dtype=torch.int64)
    return lp, tg, il, tl


class ClockSampler:
    """Samples SM clock + throttle reasons during the timed region (pynvml; nvidia-smi fallback)."""

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


def ncu_traffic(name, algo):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed ncu --set full
    capture of this workload (profiles/ncu_traffic.json); None when no capture exists for it."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        return json.load(open(p)).get("%s/algo%d" % (name, algo), {}).get("bytes")
    except Exception:
        return None


def cpu_reference_arm(name, steps, warmup, slice_b=None):
    """torch CPU ctc_loss fwd+bwd (the reference's implementation of the path) on a bounded slice, all host threads."""
    T, B, C, L, seed = WORKLOADS[name]
    sb = min(B, slice_b or (256 if name == "c3" else B))
    lp, tg, il, tl = make_inputs(name)
    lp, tg, il, tl = lp[:, :sb].contiguous(), tg[:sb], il[:sb], tl[:sb]
    # all the host threads the process may use (torchrun exports OMP_NUM_THREADS=1, which would hide the cores)
    try:
        ncores = len(os.sched_getaffinity(0))
    except AttributeError:
        ncores = os.cpu_count() or 1
    torch.set_num_threads(max(ncores, 1))
    threads = torch.get_num_threads()
    loss_fn = torch.nn.CTCLoss(reduction="sum")

    def step():
        x = lp.clone().requires_grad_(True)          # the clone stands for the model producing `out`; not timed
        t0 = time.perf_counter()
        loss = loss_fn(x, tg, il, tl)
        loss = loss / sb
        loss.backward()
        return time.perf_counter() - t0, float(loss.detach())

    for _ in range(warmup):
        step()
    dt = [step()[0] for _ in range(steps)]
    sec = sum(dt) / len(dt)
    frames = float(il.sum())
    return {"value": frames / sec, "unit": "frames/s", "cores": threads, "kind": "reference",
            "sample": "torch %s CPU nn.CTCLoss(reduction='sum') fwd+bwd fp32 on a B=%d slice of the %s batch "
                      "(T=%d, C=%d, L=%d), %d threads, mean of %d steps" % (torch.__version__, sb, name, T, C, L, threads, steps),
            "ms_per_step": sec * 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    T, B, C, L, _ = WORKLOADS[args.workload]
    r = cpu_reference_arm(args.workload, args.steps, args.warmup)
    line = {"impl": "reference", "metric": "ctc_fwd_bwd_frames_per_s", "value": r["value"], "unit": "frames/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE config 3: CTC loss fwd+bwd T=%d B=%d C=%d L=%d per GPU" % (T, B, C, L)
                       if args.workload == "c3" else "BASELINE config 2: T=%d B=%d C=%d L=%d" % (T, B, C, L),
                       "sample": r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import mddctc
    from mddctc.loss import ctc_loss_raw
    L_ = mddctc._lib.lib()

    name = args.workload
    T, B, C, L, _ = WORKLOADS[name]
    lp_h, tg_h, il_h, tl_h = make_inputs(name, rank, pinned=True)
    tg_h, il_h, tl_h = tg_h.pin_memory(), il_h.pin_memory(), tl_h.pin_memory()
    lp, tg, il, tl = lp_h.to(dev), tg_h.to(dev), il_h.to(dev), tl_h.to(dev)
    nll = torch.empty(B, dtype=torch.float32, device=dev)
    grad = torch.empty((T, B, C), dtype=torch.float32, device=dev)
    status = torch.empty(B, dtype=torch.int32, device=dev)
    loss_buf = torch.zeros(1, dtype=torch.float32, device=dev)
    frames_local = float(il_h.sum())
    inv_b = 1.0 / B                                       # the reference divides the summed loss by the batch size

    def step_resident():
        ctc_loss_raw(lp, tg, il, tl, want_grad=True, grad_scalar=inv_b, validate=False, loss_sum=loss_buf, nll=nll,
                     grad=grad, status=status, algo=args.algo)
        if world > 1:
            dist.all_reduce(loss_buf)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    L_.mdd_launch_count_reset()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        barrier()
        t_wall = time.perf_counter()
        for a, b in ev:
            a.record()
            step_resident()
            b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall
    launches = int(L_.mdd_launch_count())
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = ev[0][0].elapsed_time(ev[-1][1])
    tot = torch.tensor([total_ms, sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms, kern_ms = float(tot[0]), float(tot[1])
    ms_per_step = total_ms / args.steps
    value = frames_local * world / (ms_per_step * 1e-3)
    loss_val = float(loss_buf)

    # ---- end to end through the host-buffer path (H2D inputs, kernel, D2H loss), same K steps
    lp_d = torch.empty_like(lp)
    tg_d, il_d, tl_d = torch.empty_like(tg), torch.empty_like(il), torch.empty_like(tl)
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    from mddctc.loss import ctc_loss_streamed

    def step_e2e():
        tg_d.copy_(tg_h, non_blocking=True)
        il_d.copy_(il_h, non_blocking=True)
        tl_d.copy_(tl_h, non_blocking=True)
        if args.e2e_slices > 1:
            # the public host-buffer call: batch slices cross PCIe on a side stream while the previous slice computes
            total, _, _ = ctc_loss_streamed(lp_h, tg_d, il_d, tl_d, slices=args.e2e_slices, grad_scalar=inv_b,
                                            staging=lp_d, algo=args.algo)
            loss_buf.copy_(total)
        else:
            lp_d.copy_(lp_h, non_blocking=True)
            ctc_loss_raw(lp_d, tg_d, il_d, tl_d, want_grad=True, grad_scalar=inv_b, validate=False, loss_sum=loss_buf,
                         nll=nll, grad=grad, status=status, algo=args.algo)
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
    e2e_ms = torch.tensor([e0.elapsed_time(e1) / n_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = frames_local * world / (float(e2e_ms) * 1e-3)
    h2d = lp_h.numel() * 4 + tg_h.numel() * 8 + il_h.numel() * 8 + tl_h.numel() * 8

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg_bytes = 4.0 * C * frames_local + 4.0 * C * T * B            # per launch (one rank)
        kernel_ms = kern_ms / args.steps
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
        line = {"metric": "ctc_fwd_bwd_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": ("BASELINE config 3: CTC loss fwd+bwd T=%d B=%d C=%d L=%d per GPU" if name == "c3"
                                        else "BASELINE config 2: CTC loss fwd+bwd T=%d B=%d C=%d L=%d") % (T, B, C, L),
                           "global_batch": B * world, "sharding": "dp%d, batch-sharded, scalar loss all-reduce (NCCL)" % world,
                           "l2": "inputs+outputs %.2f GB per step >> 126 MB L2, no explicit flush" % (alg_bytes / 1e9),
                           "algo": args.algo, "loss": loss_val},
                "clocks": clk.summary(),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic(name, args.algo), "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                             "kernel_ms": kernel_ms,
                             "note": "CUDA-event time of one mdd_ctc_loss call (first packed pass + compaction and second tier of the flagged utterances + log-space sweep, both empty here + loss sum)"},
                "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                        "ms_per_step": float(e2e_ms)},
                "gpu_launches": launches, "wall_s_timed_region": t_wall}
        if not args.no_cpu and world == 1:
            r = cpu_reference_arm(name, 3, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        elif not args.no_cpu:
            line["cpu_baseline"] = None
        if args.extras:
            line["extras"] = run_extras(dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_extras(dev):
    """Secondary numbers named by BASELINE.json's metric string: joint-decode utt/s (config 4 shape) and greedy + MDD
    alignment utt/s (config 5 shape, one GPU's share)."""
    import numpy as np
    from mddctc.decoder import BeamDecoder, GreedyDecoder
    from mddctc.mdd import MDDScorer
    out = {}
    arpa = os.path.join(ROOT, "tests", "golden", "assets", "lm_phone_bg.arpa")
    vocab = {0: "blank", 1: "UNK"}
    for line in open(os.path.join(ROOT, "tests", "golden", "assets", "units")):
        if line.strip():
            vocab[len(vocab)] = line.strip()
    g = torch.Generator(device=dev).manual_seed(4)
    B, T, C = 1024, 500, 43
    x = torch.randn(B, T, C, device=dev, generator=g)
    u = torch.rand(B, T, device=dev, generator=g)
    k = torch.randint(1, C, (B, T), device=dev, generator=g)
    x[:, :, 0] += 8.0 * (u < 0.7)
    x.scatter_add_(2, k[..., None], (6.0 * (u >= 0.7))[..., None].float())
    probs = torch.softmax(x, -1)
    dec = BeamDecoder(vocab, beam_width=10, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=0.3)
    lens = torch.full((B,), T, dtype=torch.int64, device=dev)
    for _ in range(2):
        dec.decode_probs_tokens(probs, lens)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        dec.decode_probs_tokens(probs, lens)
    e1.record()
    torch.cuda.synchronize()
    out["joint_decode_utt_per_s"] = {"value": B / (e0.elapsed_time(e1) / 5 * 1e-3), "config": "beam 10, lm_alpha 0.3, B=1024, T=500, C=43 (device-resident probabilities)"}
    # greedy + alignment, ragged
    B2, T2 = 2048, 1500
    lens2 = torch.randint(50, T2 + 1, (B2,), device=dev, generator=g)
    x2 = torch.randn(T2, B2, C, device=dev, generator=g)
    x2.scatter_add_(2, torch.randint(0, C, (T2, B2, 1), device=dev, generator=g), torch.full((T2, B2, 1), 4.0, device=dev))
    lp2 = torch.log_softmax(x2, -1)
    del x2
    gd = GreedyDecoder(vocab, space_idx=-1, blank_index=0)
    sc = MDDScorer(vocab)
    can = torch.randint(2, C, (B2, 188), device=dev, generator=g, dtype=torch.int32)
    can_len = torch.clamp((lens2 + 7) // 8, max=188).to(torch.int32)

    def pipeline():
        tokens, out_len, _ = gd.decode_tokens(lp2, lens2)
        return sc.align_tokens(tokens, out_len, can, can_len, max_dec=T2, max_can=188)

    try:
        for _ in range(2):
            pipeline()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            pipeline()
        e1.record()
        torch.cuda.synchronize()
        out["greedy_mdd_utt_per_s"] = {"value": B2 / (e0.elapsed_time(e1) / 5 * 1e-3),
                                       "config": "ragged T 50-1500, B=2048, C=43, greedy decode + Levenshtein/MDD alignment"}
    except Exception as e:          # noqa
        out["greedy_mdd_utt_per_s"] = {"error": str(e)[:200]}
    del lp2
    try:
        out["ctc_from_logits"] = run_logits_extra(dev)
    except Exception as e:          # noqa
        out["ctc_from_logits"] = {"error": str(e)[:200]}
    return out


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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, fn in (("fused_ms", fused), ("unfused_ms", unfused), ("row_lse_prepass_ms", prepass)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / steps
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
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--algo", type=int, default=0)
    ap.add_argument("--e2e-slices", type=int, default=4,
                    help="batch slices of the end-to-end leg (H2D of slice k+1 overlaps the kernels of slice k); 1 = one copy")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--extras", action="store_true", help="also time joint decode and greedy+MDD alignment")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
