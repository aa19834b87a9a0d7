"""GPU parity tests: the CUDA path (through the C ABI / the Python drop-ins) against the CPU oracle, the committed
golden fixtures recorded from the real reference, and torch CPU ctc_loss.  Run on the B200 box: pytest -m gpu.

Tolerances (BASELINE.json north_star): loss 1e-5 relative, gradient 1e-4 absolute in fp32 against torch CPU CTCLoss;
decoded strings, alignments and MDD labels bit-exact."""
import math

import numpy as np
import pytest
import torch

from conftest import jload, npz
from oracle import api as O

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5
GRAD_ATOL = 1e-4


@pytest.fixture(scope="module")
def M():
    import mddctc
    assert torch.cuda.is_available()
    mddctc._lib.lib()          # fails loudly if libmddctc.so is missing
    return mddctc


def _torch_cpu_ctc(lp, tg, il, tl, zero_inf=False, dtype=torch.float32):
    x = lp.detach().cpu().to(dtype).clone().requires_grad_(True)
    nll = torch.nn.functional.ctc_loss(x, tg.cpu(), il.cpu(), tl.cpu(), blank=0, reduction="none", zero_infinity=zero_inf)
    nll.sum().backward()
    return nll.detach(), x.grad


def _check_ctc(M, lp, tg, il, tl, zero_inf=False, algo=0, max_fallback=None):
    from mddctc.loss import ctc_loss_raw, last_fallback_flags
    nll, grad, st = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), zero_infinity=zero_inf, algo=algo)
    if max_fallback is not None and algo == 0:
        # the packed kernel itself must have produced these results, not the per-utterance log-space fallback
        flags = last_fallback_flags(lp.shape[1])
        assert int((flags != 0).sum()) <= max_fallback, flags.nonzero().flatten()[:8].tolist()
    ref_nll, ref_grad = _torch_cpu_ctc(lp, tg, il, tl, zero_inf)
    ref64_nll, ref64_grad = _torch_cpu_ctc(lp, tg, il, tl, zero_inf, torch.float64)
    nll, grad = nll.cpu(), grad.cpu()
    fin = torch.isfinite(ref_nll)
    assert torch.equal(torch.isfinite(nll), fin)
    assert torch.equal(nll[~fin], ref_nll[~fin])
    # fp32 torch itself carries rounding noise of a few 1e-7 relative on long sequences: compare against both
    np.testing.assert_allclose(nll[fin].numpy(), ref64_nll[fin].numpy(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(nll[fin].numpy(), ref_nll[fin].numpy(), rtol=LOSS_RTOL)
    assert torch.equal(torch.isnan(grad), torch.isnan(ref_grad)), (torch.isnan(grad).sum(), torch.isnan(ref_grad).sum())
    ok = ~torch.isnan(ref_grad)
    # truth = torch CPU fp64.  torch's own fp32 run carries rounding noise that grows with the magnitude of nll
    # (its alpha/beta are unnormalised), so against it the bound is 1e-4 plus that run's own deviation from fp64.
    dev32 = float((ref_grad[ok].double() - ref64_grad[ok]).abs().max()) if ok.any() else 0.0
    assert float((grad[ok].double() - ref64_grad[ok]).abs().max()) <= GRAD_ATOL
    assert float((grad[ok] - ref_grad[ok]).abs().max()) <= GRAD_ATOL + dev32
    return nll, grad


@pytest.mark.parametrize("algo", [1, 0])
@pytest.mark.parametrize("name", ["basic", "repeat_ragged", "one_d", "infeasible", "infeasible_zi", "c42", "tiny"])
def test_ctc_golden(M, name, algo):
    z = npz("ctc_small.npz")
    g = lambda k: z[name + "/" + k]
    from mddctc.loss import ctc_loss_raw
    zi = bool(g("zero_inf"))
    nll, grad, st = ctc_loss_raw(torch.from_numpy(g("lp")).cuda(), torch.from_numpy(g("targets")).cuda(),
                                 torch.from_numpy(g("in_len")).cuda(), torch.from_numpy(g("tgt_len")).cuda(),
                                 zero_infinity=zi, algo=algo)
    nll, grad = nll.cpu().numpy(), grad.cpu().numpy()
    want, wg = g("nll_f32"), g("grad_f32")
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(nll), fin) and np.array_equal(nll[~fin], want[~fin])
    np.testing.assert_allclose(nll[fin], want[fin], rtol=LOSS_RTOL)
    np.testing.assert_allclose(nll[fin], g("nll_f64")[fin], rtol=LOSS_RTOL)
    assert np.array_equal(np.isnan(grad), np.isnan(wg))
    np.testing.assert_allclose(grad, wg, rtol=0, atol=GRAD_ATOL, equal_nan=True)
    np.testing.assert_allclose(grad, g("grad_f64"), rtol=0, atol=GRAD_ATOL, equal_nan=True)


@pytest.mark.parametrize("algo", [1, 0])
def test_ctc_config2_vs_torch_cpu(M, algo):
    """BASELINE config 2: T=200, B=32, C=42, L=30, plus the ragged / repeat / infeasible variant."""
    g = torch.Generator().manual_seed(0)
    lp = torch.log_softmax(torch.randn(200, 32, 42, generator=g), -1)
    tg = torch.randint(1, 42, (32, 30), generator=g)
    full_i, full_t = torch.full((32,), 200), torch.full((32,), 30)
    _check_ctc(M, lp, tg, full_i, full_t, algo=algo, max_fallback=0)
    il = torch.randint(100, 201, (32,), generator=g)
    tl = torch.randint(10, 31, (32,), generator=g)
    tg2 = tg.clone()
    tg2[:, 1::2] = tg2[:, 0::2]                     # forced repeats l_i == l_{i+1}
    il[5], tl[5] = 20, 30                           # infeasible: T_b < L_b + repeats
    tl[7] = 0
    _check_ctc(M, lp, tg2, il, tl, algo=algo)
    _check_ctc(M, lp, tg2, il, tl, zero_inf=True, algo=algo)


@pytest.mark.parametrize("algo", [1, 0])
def test_ctc_config3_slice_vs_torch_cpu(M, algo):
    """BASELINE config 3 shape on a B=48 slice (torch CPU finishes in seconds): T=1000, C=42, L=120."""
    g = torch.Generator().manual_seed(1)
    lp = torch.log_softmax(torch.randn(1000, 48, 42, generator=g), -1)
    tg = torch.randint(1, 42, (48, 120), generator=g)
    il = torch.full((48,), 1000)
    tl = torch.full((48,), 120)
    il[40:] = torch.randint(300, 1000, (8,), generator=g)
    tl[44:] = torch.randint(1, 120, (4,), generator=g)
    _check_ctc(M, lp, tg, il, tl, algo=algo, max_fallback=0)


@pytest.mark.parametrize("T,B,C,L", [(1, 1, 42, 1), (7, 3, 42, 3), (8, 2, 42, 4), (9, 5, 42, 4), (65, 33, 42, 20),
                                     (200, 9, 43, 30), (333, 17, 42, 60), (1000, 40, 42, 120)])
def test_ctc_few_utterances_concurrent_sweeps(M, T, B, C, L):
    """The path of small batches (ctc_dual.cu + the segment instantiation of the backward kernel): alpha and beta are
    swept concurrently with separate exponents, the gradient is computed on independent segments of every pair.  Shapes
    around the chunk / segment boundaries, odd batch sizes (a pair with one half empty), ragged partners (one utterance
    a fraction of the other's length), empty targets; every utterance against torch CPU fp64 (loss 1e-5 relative,
    gradient 1e-4 absolute), nothing handed to the fp64 kernel, bitwise reproducible."""
    from mddctc.loss import ctc_loss_raw
    g = torch.Generator().manual_seed(100 + T + B)
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.randint(max(1, T // 4), T + 1, (B,), generator=g)
    il[0] = T
    tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g), (il - 1).clamp(min=0) // 2 + 1).clamp(max=L)
    tl[0] = min(L, (T + 1) // 2)
    if B > 2:
        tl[2] = 0                                                 # an empty target next to a full one
    _check_ctc(M, lp, tg, il, tl, algo=0, max_fallback=0)
    a = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda())
    n0, g0 = a[0].clone(), a[1].clone()
    b = ctc_loss_raw(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda())
    assert torch.equal(n0, b[0]) and torch.equal(g0, b[1])
    # an infeasible utterance (more labels than frames allow) beside feasible ones: inf / NaN rows exactly as torch
    if T >= 8 and B >= 2:
        il2 = il.clone()
        il2[1] = max(1, int(tl[1]) - 1) if int(tl[1]) > 1 else il2[1]
        _check_ctc(M, lp, tg, il2, tl, algo=0)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_ctc_random_shapes_vs_fp64_kernel(M, seed):
    """Randomised sweep over shapes, class counts (up to the packed path's limit and beyond), ragged lengths, input
    scales and batch sizes on both sides of every dispatch threshold: the product path against the fp64 kernel (itself
    pinned to torch CPU fp64 elsewhere) -- loss 1e-5 relative, gradient 1e-4 absolute, same infinities, bitwise
    reproducible.  (This sweep is what found C = 63 corrupting its rows in round 2.)"""
    from mddctc.loss import ctc_loss_raw
    g = torch.Generator(device="cuda").manual_seed(seed)
    ri = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g, device="cuda"))
    worst = [0.0, 0.0]
    for it in range(100):
        C = [5, 20, 42, 43, 62, 63][ri(0, 6)]
        T = ri(1, 420)
        B = ri(1, 40) if it % 3 == 0 else (ri(1, 600) if it % 3 == 1 else ri(600, 1500))
        L = ri(0, min(127, max(1, T // 2)) + 1)
        x = torch.randn(T, B, C, generator=g, device="cuda") * [1.0, 1.0, 2.0, 0.3][it % 4]
        lp = torch.log_softmax(x, -1)
        tg = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
        il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
        tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
        n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
        n0, g0 = n0.clone(), g0.clone()
        n0b, g0b, _ = ctc_loss_raw(lp, tg, il, tl, algo=0, validate=False)
        assert torch.equal(n0, n0b) and torch.equal(g0.nan_to_num(7.0), g0b.nan_to_num(7.0)), (it, T, B, C, L)
        n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1, validate=False)
        fin = torch.isfinite(n1)
        assert torch.equal(torch.isfinite(n0), fin), (it, T, B, C, L)
        if bool(fin.any()):
            ne = float(((n0[fin] - n1[fin]).abs() / n1[fin].abs().clamp(min=1e-3)).max())
            ge = float((g0[:, fin] - g1[:, fin]).abs().max())
            worst = [max(worst[0], ne), max(worst[1], ge)]
            assert ne < LOSS_RTOL and ge <= GRAD_ATOL, (it, T, B, C, L, ne, ge)
    print("random shapes seed %d: worst loss rel %.2e, grad abs %.2e" % (seed, worst[0], worst[1]))


@pytest.mark.parametrize("seed", [11, 12])
def test_ctc_random_modes_vs_fp64_kernel(M, seed):
    """The same sweep over the call's modes: logits in, concatenated 1-D targets, per-utterance upstream weights,
    zero_infinity, peaky rows, long sequences (T up to 1300), long targets (L up to 255: every lattice width of the
    packed path), class counts on both sides of its limit.  Loss within 1e-5 relative (+ 1e-6 absolute: a likelihood
    within 1e-3 of 1 has a loss below fp32's resolution of P, in torch's fp32 path as much as here), gradient 1e-4,
    identical inf / NaN pattern, bitwise reproducible."""
    from mddctc.loss import ctc_loss_raw
    g = torch.Generator(device="cuda").manual_seed(seed)
    ri = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g, device="cuda"))
    for it in range(90):
        C = [5, 20, 42, 43, 62, 63, 64, 80][ri(0, 8)]
        T = ri(1, 1300) if it % 5 == 0 else ri(1, 420)
        B = ri(1, 40) if it % 3 == 0 else (ri(1, 600) if it % 3 == 1 else ri(600, 1300))
        if T > 600:
            B = min(B, 700)
        L = ri(0, min([31, 63, 127, 255][ri(0, 4)], max(1, T // 2)) + 1)
        mode = it % 6
        x = torch.randn(T, B, C, generator=g, device="cuda") * [1.0, 1.0, 2.0, 0.3, 3.0][it % 5]
        if mode == 4:                                             # peaky rows
            x = x + 12.0 * torch.nn.functional.one_hot(torch.randint(0, C, (T, B), generator=g, device="cuda"), C)
        logits = mode == 1
        inp = x if logits else torch.log_softmax(x, -1)
        il = torch.randint(1, T + 1, (B,), generator=g, device="cuda")
        tl = torch.minimum(torch.randint(0, L + 1, (B,), generator=g, device="cuda"), (il + 1) // 2).clamp(max=L)
        tg = torch.randint(1, C, (B, max(L, 1)), generator=g, device="cuda")
        if mode == 3 and int(tl.sum()) > 0:                       # concatenated targets
            tg = torch.cat([tg[b, :int(tl[b])] for b in range(B)])
        go = torch.rand(B, generator=g, device="cuda") + 0.5 if mode == 5 else None
        kw = dict(validate=False, zero_infinity=bool(it % 2), grad_out=go, logits=logits)
        n0, g0, _ = ctc_loss_raw(inp, tg, il, tl, algo=0, **kw)
        n0, g0 = n0.clone(), g0.clone()
        n0b, g0b, _ = ctc_loss_raw(inp, tg, il, tl, algo=0, **kw)
        assert torch.equal(n0, n0b) and torch.equal(g0.nan_to_num(7.0), g0b.nan_to_num(7.0)), (it, T, B, C, L, mode)
        n1, g1, _ = ctc_loss_raw(inp, tg, il, tl, algo=1, **kw)
        fin = torch.isfinite(n1)
        assert torch.equal(torch.isfinite(n0), fin) and torch.equal(torch.isnan(g0), torch.isnan(g1)), (it, T, B, C, L, mode)
        if bool(fin.any()):
            assert bool(((n0[fin] - n1[fin]).abs() <= LOSS_RTOL * n1[fin].abs() + 1e-6).all()), (it, T, B, C, L, mode)
            ok = ~torch.isnan(g1)
            assert float((g0[ok] - g1[ok]).abs().max()) <= GRAD_ATOL, (it, T, B, C, L, mode)


# ------------------------------------------------------------------------------- robustness envelope of the packed kernel
def _aligned_logits(T, B, C, L, il, tl, tg, boost, noise, g):
    """Posteriors of a trained model: a random valid alignment of each target (labels on even frames, blanks between)
    gets +boost, a fraction `noise` of the frames is confidently wrong (a random class gets the boost instead)."""
    dev = tg.device
    x = torch.randn(T, B, C, generator=g, device=dev)
    half = (il // 2).clamp(min=1)
    key = torch.rand(B, T // 2, generator=g, device=dev)
    key = key + (torch.arange(T // 2, device=dev)[None, :] >= half[:, None]) * 2.0      # positions beyond T_b/2 sort last
    pos = key.argsort(dim=1)[:, :L].sort(dim=1)[0] * 2                                   # (B, L) label frames, increasing
    lab = torch.zeros(B, T, dtype=torch.long, device=dev)
    valid = torch.arange(L, device=dev)[None, :] < tl[:, None]
    lab.scatter_(1, torch.where(valid, pos, torch.full_like(pos, T - 1)), torch.where(valid, tg, torch.zeros_like(tg)))
    lab = lab.t().contiguous()
    flip = torch.rand(T, B, generator=g, device=dev) < noise
    lab = torch.where(flip, torch.randint(0, C, (T, B), generator=g, device=dev), lab)
    x.scatter_add_(2, lab[..., None], torch.full((T, B, 1), float(boost), device=dev))
    return x


ENVELOPE = [  # name, boost (None = unstructured randn * scale), noise / scale, the first pass keeps everything, the
    # second tier (2-frame chunks over the flagged list) leaves nothing for the fp64 kernel
    ("randn*2", None, 2.0, True, True), ("randn*3", None, 3.0, False, True), ("randn*5", None, 5.0, False, False),
    # (aligned+10: at this batch size the beta sweep of ctc_dual.cu carries its own exponents and its decay guard hands
    # one utterance of 256 to the second tier; the one-warp-per-pair kernel of larger batches keeps all of them)
    ("aligned+6", 6, 0.1, True, True), ("aligned+10", 10, 0.1, False, True), ("aligned+10 clean", 10, 0.0, True, True),
    ("aligned+16", 16, 0.1, False, True), ("aligned+25", 25, 0.1, False, False),
    ("aligned+16 noise .3", 16, 0.3, False, True),
]


@pytest.mark.parametrize("name,boost,par,all_fast,no_log", ENVELOPE, ids=[e[0] for e in ENVELOPE])
def test_ctc_packed_kernel_envelope(M, name, boost, par, all_fast, no_log):
    """The packed fp32 kernel has a finite dynamic range (block floating point, one exponent per lane and chunk).
    Whatever it does not hand on must be RIGHT, and whatever it hands on is redone by the fp64 kernel: on peaky
    'trained model' posteriors, with and without confidently wrong frames, EVERY utterance of the product path --
    whichever tier produced it -- holds north_star's tolerance against torch CPU fp64; on the benign families the first
    pass keeps all of them."""
    from mddctc.loss import ctc_loss_raw, last_fallback_flags
    g = torch.Generator(device="cuda").manual_seed(21)
    T, B, C, L = 600, 256, 42, 80
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.randint(400, T + 1, (B,), generator=g, device="cuda")
    tl = torch.randint(20, L + 1, (B,), generator=g, device="cuda")
    if boost is None:
        x = torch.randn(T, B, C, generator=g, device="cuda") * par
    else:
        x = _aligned_logits(T, B, C, L, il, tl, tg, boost, par, g)
    lp = torch.log_softmax(x, -1)
    ref_nll, ref_grad = _torch_cpu_ctc(lp, tg, il, tl, dtype=torch.float64)   # the truth, all B utterances
    ref_nll, ref_grad = ref_nll.cuda(), ref_grad.cuda()

    def err(n, gr, mask=None):
        ne = ((n.double() - ref_nll) / ref_nll).abs()
        ge = (gr.double() - ref_grad).abs().amax(dim=(0, 2))
        if mask is not None:
            ne, ge = ne[mask], ge[mask]
        return (float(ne.max()), float(ge.max())) if ne.numel() else (0.0, 0.0)

    n1, g1, _ = ctc_loss_raw(lp, tg, il, tl, algo=1)                          # the fp64 kernel: exact to fp32 rounding
    e1 = err(n1, g1)
    assert e1[0] < 1e-6 and e1[1] < 2e-6, (name, e1)
    n2, g2, _ = ctc_loss_raw(lp, tg, il, tl, algo=2, validate=False)          # first packed pass alone, no fallback
    kept = last_fallback_flags(B) == 0
    if all_fast:
        assert bool(kept.all()), (name, (~kept).nonzero().flatten()[:8].tolist())
    e2 = err(n2, g2, kept)                                                    # what it keeps is right
    assert e2[0] < LOSS_RTOL and e2[1] <= GRAD_ATOL, (name, e2)
    n0, g0, _ = ctc_loss_raw(lp, tg, il, tl, algo=0)                          # the product path
    n0, g0 = n0.clone(), g0.clone()
    n0b, g0b, _ = ctc_loss_raw(lp, tg, il, tl, algo=0)                        # ... is bitwise reproducible,
    assert torch.equal(n0, n0b) and torch.equal(g0, g0b)                      # whichever tiers ran
    left = last_fallback_flags(B) != 0                                        # what the two packed tiers did not finish
    if no_log:                                                                # a stray utterance (< 2 %) may still go on
        assert int(left.sum()) <= (0 if all_fast else B // 50), (name, left.nonzero().flatten()[:8].tolist())
    e0 = err(n0, g0)
    print("envelope %-20s first pass kept %d/%d, left for fp64 %d | loss rel %.2e grad abs %.2e (fp64 kernel %.1e)"
          % (name, int(kept.sum()), B, int(left.sum()), e0[0], e0[1], e1[1]))
    assert e0[0] < LOSS_RTOL and e0[1] <= GRAD_ATOL, (name, e0)


@pytest.mark.parametrize("seed", [3, 4, 5])
def test_ctc_long_ragged_t1500(M, seed):
    """BASELINE config 5's length range: T ragged 50..1500, L up to 180, log_softmax(randn) and a peaky family;
    loss 1e-5 relative, gradient 1e-4 absolute against torch CPU fp64 on every utterance, margin reported."""
    from mddctc.loss import ctc_loss_raw, last_fallback_flags
    g = torch.Generator(device="cuda").manual_seed(seed)
    T, B, C, L = 1500, 96, 42, 180
    il = torch.randint(50, T + 1, (B,), generator=g, device="cuda")
    il[:8] = T                                                                # the longest case is always present
    tl = torch.minimum(torch.randint(1, L + 1, (B,), generator=g, device="cuda"), (il // 3).clamp(min=1))
    tl[:8] = L
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    worst = 0.0
    for fam in ("randn", "aligned+12"):
        if fam == "randn":
            x = torch.randn(T, B, C, generator=g, device="cuda")
        else:
            x = _aligned_logits(T, B, C, L, il, tl, tg, 12, 0.1, g)
        lp = torch.log_softmax(x, -1)
        ref_nll, ref_grad = _torch_cpu_ctc(lp, tg, il, tl, dtype=torch.float64)
        for algo in (0, 1):
            nll, grad, _ = ctc_loss_raw(lp, tg, il, tl, algo=algo)
            np.testing.assert_allclose(nll.cpu().numpy(), ref_nll.numpy(), rtol=LOSS_RTOL)
            ge = float((grad.cpu().double() - ref_grad).abs().max())
            if algo == 0:
                worst = max(worst, ge)
                left = int((last_fallback_flags(B) != 0).sum())
                assert left <= B // 32, (fam, left)                           # the packed tiers produce (nearly) all of them
            assert ge <= (GRAD_ATOL if algo == 0 else 2e-6), (fam, algo, ge)
    print("T=1500 ragged seed %d: worst gradient error of the packed path %.2e (bound %.0e)" % (seed, worst, GRAD_ATOL))


# ------------------------------------------------------------------------------- CTC from logits
def _torch_cpu_ctc_logits(x, tg, il, tl, zero_inf=False, dtype=torch.float32, go=None):
    """The reference's own sequence: out = F.log_softmax(out, -1) (model_ctc.py:214-218) -> CTCLoss -> backward()."""
    xx = x.detach().cpu().to(dtype).clone().requires_grad_(True)
    nll = torch.nn.functional.ctc_loss(torch.log_softmax(xx, -1), tg.cpu(), il.cpu(), tl.cpu(), blank=0,
                                       reduction="none", zero_infinity=zero_inf)
    (nll if go is None else nll * go.to(dtype)).sum().backward()
    return nll.detach(), xx.grad


def _check_ctc_logits(M, x, tg, il, tl, zero_inf=False, algo=0, max_fallback=None, go=None):
    from mddctc.loss import ctc_loss_raw, last_fallback_flags
    T, B, C = x.shape
    lse = torch.empty((T, B), dtype=torch.float32, device="cuda")
    am = torch.empty((T, B), dtype=torch.uint8, device="cuda")
    nll, grad, st = ctc_loss_raw(x.cuda(), tg.cuda(), il.cuda(), tl.cuda(), zero_infinity=zero_inf, algo=algo,
                                 logits=True, row_lse=lse, argmax=am, grad_out=None if go is None else go.cuda())
    if max_fallback is not None and algo == 0:
        flags = last_fallback_flags(B)
        assert int((flags != 0).sum()) <= max_fallback, flags.nonzero().flatten()[:8].tolist()
    ref_nll, ref_grad = _torch_cpu_ctc_logits(x, tg, il, tl, zero_inf, go=go)
    ref64_nll, ref64_grad = _torch_cpu_ctc_logits(x, tg, il, tl, zero_inf, torch.float64, go=go)
    nll, grad = nll.cpu(), grad.cpu()
    np.testing.assert_allclose(lse.cpu().numpy(), torch.logsumexp(x.double(), -1).numpy(), rtol=2e-6, atol=2e-6)
    assert torch.equal(am.cpu().long(), torch.max(x, dim=-1)[1])
    fin = torch.isfinite(ref_nll)
    assert torch.equal(torch.isfinite(nll), fin)
    assert torch.equal(nll[~fin], ref_nll[~fin])
    np.testing.assert_allclose(nll[fin].numpy(), ref64_nll[fin].numpy(), rtol=LOSS_RTOL)
    np.testing.assert_allclose(nll[fin].numpy(), ref_nll[fin].numpy(), rtol=LOSS_RTOL)
    assert torch.equal(torch.isnan(grad), torch.isnan(ref_grad)), (torch.isnan(grad).sum(), torch.isnan(ref_grad).sum())
    ok = ~torch.isnan(ref_grad)
    dev32 = float((ref_grad[ok].double() - ref64_grad[ok]).abs().max()) if ok.any() else 0.0
    assert float((grad[ok].double() - ref64_grad[ok]).abs().max()) <= GRAD_ATOL
    assert float((grad[ok] - ref_grad[ok]).abs().max()) <= GRAD_ATOL + dev32
    return nll, grad


@pytest.mark.parametrize("algo", [1, 0])
@pytest.mark.parametrize("name", ["basic", "repeat_ragged", "infeasible", "infeasible_zi", "c42", "c45_wide"])
def test_ctc_logits_golden(M, name, algo):
    z = npz("ctc_logits_small.npz")
    g = lambda k: z[name + "/" + k]
    from mddctc.loss import ctc_loss_raw
    x = torch.from_numpy(g("logits")).cuda()
    T, B, C = x.shape
    lse = torch.empty((T, B), dtype=torch.float32, device="cuda")
    am = torch.empty((T, B), dtype=torch.uint8, device="cuda")
    nll, grad, st = ctc_loss_raw(x, torch.from_numpy(g("targets")).cuda(), torch.from_numpy(g("in_len")).cuda(),
                                 torch.from_numpy(g("tgt_len")).cuda(), zero_infinity=bool(g("zero_inf")), algo=algo,
                                 logits=True, row_lse=lse, argmax=am, grad_out=torch.from_numpy(g("grad_out")).cuda())
    nll, grad = nll.cpu().numpy(), grad.cpu().numpy()
    want, wg = g("nll_f32"), g("grad_f32")
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(nll), fin) and np.array_equal(nll[~fin], want[~fin])
    np.testing.assert_allclose(nll[fin], want[fin], rtol=LOSS_RTOL)
    np.testing.assert_allclose(nll[fin], g("nll_f64")[fin], rtol=LOSS_RTOL)
    assert np.array_equal(np.isnan(grad), np.isnan(wg))
    # truth = torch CPU fp64; torch's own fp32 run deviates from it on wide logits, so against fp32 the bound is 1e-4
    # plus that deviation (as in _check_ctc)
    dev32 = float(np.nanmax(np.abs(wg.astype(np.float64) - g("grad_f64")))) if np.isfinite(wg).any() else 0.0
    np.testing.assert_allclose(grad, g("grad_f64"), rtol=0, atol=GRAD_ATOL, equal_nan=True)
    np.testing.assert_allclose(grad, wg, rtol=0, atol=GRAD_ATOL + dev32, equal_nan=True)
    np.testing.assert_allclose(lse.cpu().numpy(), g("lse_f64"), rtol=2e-6, atol=2e-6)
    assert np.array_equal(am.cpu().numpy(), g("argmax"))
    # and against the oracle's restatement of the same chain
    r = O.ctc_loss_logits(g("logits"), g("targets"), g("in_len"), g("tgt_len"), zero_infinity=bool(g("zero_inf")),
                          grad_out=g("grad_out"))
    np.testing.assert_allclose(grad, r["grad"], rtol=0, atol=GRAD_ATOL, equal_nan=True)


@pytest.mark.parametrize("algo", [1, 0])
def test_ctc_logits_config2_vs_torch_cpu(M, algo):
    """BASELINE config 2 shape on logits (scaled randn), plus the ragged / repeat / infeasible variant."""
    g = torch.Generator().manual_seed(10)
    x = torch.randn(200, 32, 42, generator=g) * 2.5
    tg = torch.randint(1, 42, (32, 30), generator=g)
    _check_ctc_logits(M, x, tg, torch.full((32,), 200), torch.full((32,), 30), algo=algo, max_fallback=0)
    il = torch.randint(100, 201, (32,), generator=g)
    tl = torch.randint(10, 31, (32,), generator=g)
    tg2 = tg.clone()
    tg2[:, 1::2] = tg2[:, 0::2]
    il[5], tl[5] = 20, 30
    tl[7] = 0
    go = torch.rand(32, generator=g) + 0.5
    _check_ctc_logits(M, x, tg2, il, tl, algo=algo, go=go)
    _check_ctc_logits(M, x, tg2, il, tl, zero_inf=True, algo=algo)


def test_ctc_logits_config3_slice_vs_torch_cpu(M):
    """BASELINE config 3 shape on a B=48 slice, logits in; odd batch tail and ragged lengths."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(1000, 47, 42, generator=g) * 2 + torch.randn(1000, 47, 1, generator=g) * 5   # per-frame offsets
    tg = torch.randint(1, 42, (47, 120), generator=g)
    il = torch.full((47,), 1000)
    tl = torch.full((47,), 120)
    il[40:] = torch.randint(300, 1000, (7,), generator=g)
    tl[44:] = torch.randint(1, 120, (3,), generator=g)
    _check_ctc_logits(M, x, tg, il, tl, algo=0, max_fallback=0)


@pytest.mark.parametrize("C", [43, 45, 5, 62, 63, 64])
def test_ctc_logits_other_class_counts_and_strides(M, C):
    g = torch.Generator().manual_seed(100 + C)
    T, B, L = 150, 9, 20
    big = torch.randn(T, 2 * B, C + 3, generator=g) * 1.5
    x = big[:, ::2, :C]                                # strided view: stride_b = 2*(C+3), stride_t = 2B*(C+3)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.randint(60, T + 1, (B,), generator=g)
    tl = torch.randint(0, L + 1, (B,), generator=g)
    from mddctc.loss import ctc_loss_raw
    for algo in (1, 0):
        _check_ctc_logits(M, x.contiguous(), tg, il, tl, algo=algo, max_fallback=0)
    # the strided view itself (no copy on the way in).  The strided pre-pass sums a row's exponentials in a different
    # order (warp per row) than the tiled one (thread per row): equal up to the last bits of the row log-sum-exp.
    xs = big.cuda()[:, ::2, :C]
    a = ctc_loss_raw(xs, tg.cuda(), il.cuda(), tl.cuda(), logits=True)
    a = (a[0].clone(), a[1].clone())
    b = ctc_loss_raw(xs.contiguous(), tg.cuda(), il.cuda(), tl.cuda(), logits=True)
    np.testing.assert_allclose(a[0].cpu().numpy(), b[0].cpu().numpy(), rtol=1e-6)
    assert float((a[1] - b[1]).abs().max()) < 1e-5


def test_ctc_from_logits_module_train_step(M):
    """The training-step call pattern of steps/train_ctc.py:72-96 with the model tail of model_ctc.py:214-218 folded in:
    loss = CTCLoss('sum')(logits...); loss /= batch_size; loss.backward(); plus the argmax for the train-time PER."""
    from mddctc.loss import CTCLossFromLogits
    g = torch.Generator().manual_seed(12)
    T, B, C, L = 120, 6, 42, 15
    x0 = torch.randn(T, B, C, generator=g) * 3
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.tensor([120, 100, 90, 120, 61, 77])
    tl = torch.tensor([15, 12, 9, 15, 3, 8])
    x = x0.cuda().requires_grad_(True)
    am = torch.empty((T, B), dtype=torch.uint8, device="cuda")
    loss = CTCLossFromLogits(reduction="sum")(x, tg.cuda(), il.cuda(), tl.cuda(), argmax=am)
    loss = loss / B
    loss.backward()
    xr = x0.clone().requires_grad_(True)
    out = torch.log_softmax(xr, -1)
    ref = torch.nn.CTCLoss(reduction="sum")(out, tg, il, tl) / B
    ref.backward()
    np.testing.assert_allclose(float(loss), float(ref), rtol=LOSS_RTOL)
    assert float((x.grad.cpu() - xr.grad).abs().max()) <= GRAD_ATOL
    index = torch.max(out, dim=-1)[1].transpose(0, 1)
    assert torch.equal(am.cpu().long().transpose(0, 1), index)
    for red in ("mean", "none"):
        got = CTCLossFromLogits(reduction=red)(x0.cuda(), tg.cuda(), il.cuda(), tl.cuda())
        want = torch.nn.CTCLoss(reduction=red)(out.detach(), tg, il, tl)
        np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=LOSS_RTOL)


def test_ctc_loss_streamed_matches_whole_batch(M):
    """Host-resident log-probs in batch slices over a side stream: per-utterance results are those of the one-shot call."""
    from mddctc.loss import ctc_loss_raw, ctc_loss_streamed
    g = torch.Generator().manual_seed(15)
    T, B, C, L = 300, 37, 42, 40
    lp_h = torch.log_softmax(torch.randn(T, B, C, generator=g), -1).pin_memory()
    tg = torch.randint(1, C, (B, L), generator=g).cuda()
    il = torch.randint(150, T + 1, (B,), generator=g).cuda()
    tl = torch.randint(0, L + 1, (B,), generator=g).cuda()
    nll, grad, _ = ctc_loss_raw(lp_h.cuda(), tg, il, tl, grad_scalar=0.5)
    for slices in (1, 4, 5):
        total, nll_s, grads = ctc_loss_streamed(lp_h, tg, il, tl, slices=slices, grad_scalar=0.5)
        gs = torch.cat(grads, dim=1)
        if slices == 1:
            assert torch.equal(nll_s, nll) and torch.equal(gs, grad)
        else:
            # an utterance's chunk boundaries (hence its rounding) follow its partner in the warp when the two differ in
            # length, and slicing changes the partners: equal to the last bits, not bitwise
            np.testing.assert_allclose(nll_s.cpu().numpy(), nll.cpu().numpy(), rtol=1e-6)
            assert float((gs - grad).abs().max()) < 2e-6
        np.testing.assert_allclose(float(total), float(nll.sum()), rtol=1e-6)
    # slice-major host layout: a list of contiguous (T, B_k, C) pinned tensors, one contiguous copy each
    cuts = [0, 10, 20, 29, 37]
    parts = [lp_h[:, a:b].contiguous().pin_memory() for a, b in zip(cuts[:-1], cuts[1:])]
    total, nll_s, grads = ctc_loss_streamed(parts, tg, il, tl, grad_scalar=0.5)
    np.testing.assert_allclose(nll_s.cpu().numpy(), nll.cpu().numpy(), rtol=1e-6)
    assert float((torch.cat(grads, dim=1) - grad).abs().max()) < 2e-6
    np.testing.assert_allclose(float(total), float(nll.sum()), rtol=1e-6)
    with pytest.raises(RuntimeError):
        ctc_loss_streamed(lp_h.cuda(), tg, il, tl)


def test_ctc_from_logits_api_edges(M):
    """(T,C) input, fp64 logits, lengths as tuples, CPU logits in, zero-length targets under 'mean', zero_infinity."""
    from mddctc.loss import ctc_loss_from_logits
    F = torch.nn.functional
    g = torch.Generator().manual_seed(14)
    x2 = (torch.randn(50, 9, generator=g) * 2).requires_grad_(True)                   # (T, C): one utterance
    tg1 = torch.randint(1, 9, (6,), generator=g)
    got = ctc_loss_from_logits(x2, tg1, (50,), (6,), reduction="sum")
    got.backward()
    xr = x2.detach().clone().requires_grad_(True)
    want = F.ctc_loss(torch.log_softmax(xr, -1), tg1, (50,), (6,), reduction="sum")
    want.backward()
    np.testing.assert_allclose(float(got.detach()), float(want.detach()), rtol=LOSS_RTOL)
    assert x2.grad.shape == xr.grad.shape and float((x2.grad - xr.grad).abs().max()) <= GRAD_ATOL
    T, B, C, L = 80, 5, 11, 7
    x = torch.randn(T, B, C, generator=g, dtype=torch.float64) * 2
    tg = torch.randint(1, C, (B, L), generator=g)
    il, tl = torch.tensor([80, 64, 80, 12, 5]), torch.tensor([7, 0, 3, 7, 7])                # row 4 infeasible, row 1 empty
    for red in ("mean", "sum", "none"):
        for zi in (False, True):
            got = ctc_loss_from_logits(x.cuda(), tg.cuda(), il, tl, reduction=red, zero_infinity=zi)
            want = F.ctc_loss(torch.log_softmax(x, -1), tg, il, tl, reduction=red, zero_infinity=zi)
            assert got.dtype == torch.float64 and got.shape == want.shape
            np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=LOSS_RTOL)
    xc = x.float().requires_grad_(True)                                                   # CPU tensor in, CPU grad out
    loss = ctc_loss_from_logits(xc, tg, il, tl, reduction="sum", zero_infinity=True)
    loss.backward()
    assert not loss.is_cuda and not xc.grad.is_cuda
    xr = x.float().clone().requires_grad_(True)
    F.ctc_loss(torch.log_softmax(xr, -1), tg, il, tl, reduction="sum", zero_infinity=True).backward()
    assert float((xc.grad - xr.grad).abs().max()) <= GRAD_ATOL
    with pytest.raises(RuntimeError):
        ctc_loss_from_logits(x.cuda(), tg.cuda(), torch.tensor([81, 1, 1, 1, 1]), tl)


def test_ctc_logits_full_config3_properties(M):
    """Config 3 at full size (T=1000, B=4096, C=42, L=120) through size-independent properties: the logit gradient's
    rows sum to zero, the loss equals the log-prob path's on log_softmax of the same logits, two launches agree bitwise,
    a 16-utterance sample agrees with torch CPU."""
    from mddctc.loss import ctc_loss_raw
    g = torch.Generator(device="cuda").manual_seed(13)
    T, B, C, L = 1000, 4096, 42, 120
    x = torch.randn(T, B, C, generator=g, device="cuda") * 2
    tg = torch.randint(1, C, (B, L), generator=g, device="cuda")
    il = torch.full((B,), T, device="cuda")
    tl = torch.full((B,), L, device="cuda")
    nll, grad, _ = ctc_loss_raw(x, tg, il, tl, logits=True)
    nll2, grad2, _ = ctc_loss_raw(x, tg, il, tl, logits=True)
    assert torch.equal(nll, nll2) and torch.equal(grad, grad2)
    assert float(grad.sum(-1).abs().max()) < 2e-5
    nll_lp, grad_lp, _ = ctc_loss_raw(torch.log_softmax(x, -1), tg, il, tl)
    np.testing.assert_allclose(nll.cpu().numpy(), nll_lp.cpu().numpy(), rtol=2e-6)
    assert float((grad - grad_lp).abs().max()) < 2e-5
    sel = torch.arange(0, B, B // 16)[:16]
    ref_nll, ref_grad = _torch_cpu_ctc_logits(x[:, sel].cpu(), tg[sel].cpu(), il[sel].cpu(), tl[sel].cpu(),
                                              dtype=torch.float64)
    np.testing.assert_allclose(nll[sel].cpu().numpy(), ref_nll.numpy(), rtol=LOSS_RTOL)
    assert float((grad[:, sel].cpu().double() - ref_grad).abs().max()) <= GRAD_ATOL


@pytest.mark.parametrize("C", [43, 45, 5, 62, 63, 64])
def test_ctc_other_class_counts(M, C):
    g = torch.Generator().manual_seed(C)
    T, B, L = 150, 9, 20
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.randint(60, T + 1, (B,), generator=g)
    tl = torch.randint(0, L + 1, (B,), generator=g)
    for algo in (1, 0):
        _check_ctc(M, lp, tg, il, tl, algo=algo, max_fallback=0)


def test_ctc_peaky_and_extreme_inputs(M):
    """Peaky posteriors (trained-model like) and rows with log-probs down to -80: the fast path must either stay within
    tolerance or hand the utterance to the log-space kernel."""
    g = torch.Generator().manual_seed(77)
    T, B, C, L = 300, 16, 42, 40
    x = torch.randn(T, B, C, generator=g)
    x[:, :, 0] += 6 * (torch.rand(T, B, generator=g) < 0.7)
    idx = torch.randint(1, C, (T, B), generator=g)
    x.scatter_add_(2, idx[..., None], 9 * torch.ones(T, B, 1))
    x[::7] *= 8                                     # some frames with huge dynamic range
    lp = torch.log_softmax(x, -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.full((B,), T)
    tl = torch.randint(5, L + 1, (B,), generator=g)
    _check_ctc(M, lp, tg, il, tl, algo=0)
    _check_ctc(M, lp, tg, il, tl, algo=1)


def test_ctc_full_config3_properties(M):
    """Full BASELINE config 3 (T=1000, B=4096, C=42, L=120): size-independent properties + a 16-utterance slice vs torch."""
    from mddctc.loss import ctc_loss_raw
    g = torch.Generator().manual_seed(1)
    T, B, C, L = 1000, 4096, 42, 120
    lp = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.full((B,), T)
    il[::5] = torch.randint(500, T, (len(il[::5]),), generator=g)
    tl = torch.full((B,), L)
    d = lp.cuda()
    nll, grad, st = ctc_loss_raw(d, tg.cuda(), il.cuda(), tl.cuda())
    assert int(st.abs().sum()) == 0
    assert bool(torch.isfinite(nll).all())
    from mddctc.loss import last_fallback_flags
    assert int((last_fallback_flags(B) != 0).sum()) == 0       # no utterance needed the log-space fallback
    # rows of the gradient sum to ~0 (softmax prob - occupancy, both sum to 1) for valid frames, exactly 0 beyond T_b
    rs = grad.sum(-1)
    assert float(rs.abs().max()) < 5e-4
    tmask = torch.arange(T, device="cuda")[:, None] >= il.cuda()[None, :]
    assert float(grad.abs().sum(-1)[tmask].max()) == 0.0
    # determinism: bitwise identical on a second run
    nll2, grad2, _ = ctc_loss_raw(d, tg.cuda(), il.cuda(), tl.cuda())
    assert torch.equal(nll, nll2) and torch.equal(grad, grad2)
    sel = torch.arange(0, B, 256)
    ref_nll, ref_grad = _torch_cpu_ctc(lp[:, sel], tg[sel], il[sel], tl[sel], dtype=torch.float64)
    np.testing.assert_allclose(nll[sel.cuda()].cpu().numpy(), ref_nll.numpy(), rtol=LOSS_RTOL)
    assert float((grad[:, sel.cuda()].cpu().double() - ref_grad).abs().max()) <= GRAD_ATOL


def test_ctcloss_module_autograd(M):
    """The call pattern of train_ctc.py:72-96: reduction='sum', in-place division by the batch size, backward."""
    from mddctc.loss import CTCLoss
    g = torch.Generator().manual_seed(5)
    T, B, C, L = 120, 8, 42, 25
    logits = torch.randn(T, B, C, generator=g)
    tg = torch.randint(1, C, (B, L), generator=g)
    il = torch.randint(80, T + 1, (B,), generator=g)
    tl = torch.randint(5, L + 1, (B,), generator=g)
    for red in ("sum", "mean", "none"):
        a = logits.clone().cuda().requires_grad_(True)
        out = CTCLoss(reduction=red)(torch.log_softmax(a, -1), tg.cuda(), il.cuda(), tl.cuda())
        if red == "sum":
            out /= B
        (out.sum()).backward()
        refs = {}
        for dt in (torch.float32, torch.float64):
            b = logits.clone().to(dt).requires_grad_(True)
            ref = torch.nn.CTCLoss(reduction=red)(torch.log_softmax(b, -1), tg, il, tl)
            if red == "sum":
                ref /= B
            ref.sum().backward()
            refs[dt] = (ref.detach(), b.grad)
        np.testing.assert_allclose(out.detach().cpu().numpy(), refs[torch.float32][0].numpy(), rtol=LOSS_RTOL)
        dev32 = float((refs[torch.float32][1].double() - refs[torch.float64][1]).abs().max())
        assert float((a.grad.cpu().double() - refs[torch.float64][1]).abs().max()) <= GRAD_ATOL
        assert float((a.grad.cpu() - refs[torch.float32][1]).abs().max()) <= GRAD_ATOL + dev32
    with pytest.raises(RuntimeError):
        CTCLoss()(torch.log_softmax(logits, -1).cuda(), tg.cuda(), (il + 1000).cuda(), tl.cuda())
    # tuple lengths + 1-D targets + (T,C) input
    lp1 = torch.log_softmax(logits[:, 0], -1)
    out = CTCLoss(reduction="sum")(lp1.cuda(), tg[0, :7].cuda(), (T,), (7,))
    ref = torch.nn.CTCLoss(reduction="sum")(lp1, tg[0, :7], (T,), (7,))
    np.testing.assert_allclose(float(out), float(ref), rtol=LOSS_RTOL)


# ----------------------------------------------------------------------------------------------------- greedy
def test_greedy_golden(M):
    from mddctc.decoder import GreedyDecoder
    z, js = npz("greedy.npz"), jload("greedy.json.gz")
    for i, case in js.items():
        voc = {int(k): v for k, v in case["vocab"].items()}
        dec = GreedyDecoder(voc, space_idx=-1, blank_index=0)
        assert dec.decode(torch.from_numpy(z[i + "/lp"]), z[i + "/lens"].tolist()) == case["out"], i
        assert dec.decode(torch.from_numpy(z[i + "/lp"]).cuda(), z[i + "/lens"].tolist()) == case["out"], i


def test_greedy_random_vs_oracle(M, vocab43):
    from mddctc.decoder import GreedyDecoder
    rng = np.random.default_rng(3)
    dec = GreedyDecoder(vocab43, space_idx=-1, blank_index=0)
    for T, B in ((500, 64), (37, 5), (129, 131), (1, 1)):
        x = rng.standard_normal((T, B, 43)).astype(np.float32)
        x = np.round(x * 2) / 2                       # exact ties
        run = rng.integers(0, 43, size=(T, B))
        for t in range(1, T):
            keep = rng.random(B) < 0.7
            run[t, keep] = run[t - 1, keep]
        x[np.arange(T)[:, None], np.arange(B)[None, :], run] += 2.5
        lens = rng.integers(0, T + 1, size=B)
        want = O.greedy_decode(x, lens, vocab43)
        assert dec.decode(torch.from_numpy(x), lens.tolist()) == want
        tokens, out_len, amax = dec.decode_tokens(torch.from_numpy(x).cuda(), lens.tolist())
        assert torch.equal(amax[:T].long().cpu(), torch.from_numpy(x).argmax(-1))


def test_greedy_full_size_properties(M, vocab43):
    """Config-3-sized stream (T=1000, B=4096): argmax equals torch's, collapse is idempotent and length-consistent."""
    from mddctc.decoder import GreedyDecoder
    g = torch.Generator(device="cuda").manual_seed(4)
    T, B, C = 1000, 4096, 43
    lp = torch.randn(T, B, C, device="cuda", generator=g)
    lens = torch.randint(50, T + 1, (B,), device="cuda", generator=g)
    dec = GreedyDecoder(vocab43, space_idx=-1, blank_index=0)
    tokens, out_len, amax = dec.decode_tokens(lp, lens)
    assert torch.equal(amax.long(), lp.argmax(-1))
    assert int(out_len.max()) <= T and int(out_len.min()) >= 0
    sel = [0, 17, 4095]
    want_tok, want_len, _ = O.greedy_tokens(lp[:, sel].cpu().numpy(), lens[sel].cpu().numpy())
    for j, b in enumerate(sel):
        assert int(out_len[b]) == int(want_len[j])
        assert tokens[b, :out_len[b]].cpu().tolist() == want_tok[j, :want_len[j]].tolist()


# ----------------------------------------------------------------------------------------------------- Levenshtein
def test_wer_corpus_totals(M):
    """Known-answer totals recorded by the reference's own runs."""
    from mddctc.decoder import Decoder
    dec = Decoder({0: "blank"}, space_idx=-1, blank_index=0)
    data = jload("wer_corpus.json.gz")
    for name, c in data.items():
        res = dec.wer_batch(c["hyp"], c["ref"])
        assert sum(r[0] for r in res) == c["total"], name
        assert [r[0] for r in res] == c["dist"]
        if c["paths"]:
            assert ["".join(r[1]) for r in res] == c["paths"]
    d2 = Decoder({0: "blank"}, space_idx=-1, blank_index=0, variant="cnn-rnn-ctc")
    c = data["cnn_rnn_hyp_human"]
    assert sum(d2.wer_batch(c["hyp"], c["ref"])) == 7934


def test_wer_single_call_semantics(M):
    from mddctc.decoder import Decoder
    dec = Decoder({0: "blank"}, space_idx=-1, blank_index=0)
    assert dec.wer("a b c d", "a x c e f") == (3, ["-", "S", "-", "D", "S"])
    assert dec.wer("a b c", "b c d") == (2, ["I", "-", "-", "D"])
    with pytest.raises(TypeError):
        dec.wer("", "a b")
    assert dec.cer("kitten", "sitting") == 3


def test_edit_align_random_vs_oracle(M):
    from mddctc.decoder import edit_align_tokens
    rng = np.random.default_rng(8)
    for (B, Lh, Lr, V) in ((300, 40, 40, 6), (64, 200, 190, 40), (40, 33, 500, 3), (16, 700, 650, 12), (50, 1, 1, 2)):
        hl = rng.integers(0, Lh + 1, size=B).astype(np.int32)
        rl = rng.integers(0, Lr + 1, size=B).astype(np.int32)
        H = rng.integers(0, V, size=(B, Lh)).astype(np.int32)
        R = rng.integers(0, V, size=(B, Lr)).astype(np.int32)
        for b in range(0, B, 3):                     # correlated pairs
            n = min(hl[b], rl[b])
            H[b, :n] = R[b, :n]
            flip = rng.random(n) < 0.2
            H[b, :n][flip] = rng.integers(0, V, size=int(flip.sum()))
        dist, paths, st = O.edit_align([H[b, :hl[b]].tolist() for b in range(B)], [R[b, :rl[b]].tolist() for b in range(B)])
        r = edit_align_tokens(torch.from_numpy(H).cuda(), torch.from_numpy(hl).cuda(), torch.from_numpy(R).cuda(),
                              torch.from_numpy(rl).cuda(), Lh, Lr, want_ops=True)
        assert r["dist"].cpu().numpy().tolist() == dist.tolist()
        assert r["status"].cpu().numpy().tolist() == st.tolist()
        ops, ol = r["ops"].cpu().numpy(), r["ops_len"].cpu().numpy()
        for b in range(B):
            assert "".join(chr(c) for c in ops[b, :ol[b]]) == "".join(paths[b]), (B, Lh, Lr, b)


def test_compute_wer_vs_oracle(M):
    from mddctc.mdd import compute_wer
    rng = np.random.default_rng(12)
    B, T, Lmax = 32, 200, 30
    idx = rng.integers(0, 42, size=(B, T))
    for t in range(1, T):
        keep = rng.random(B) < 0.6
        idx[keep, t] = idx[keep, t - 1]
    idx[rng.random((B, T)) < 0.4] = 0
    il = rng.integers(100, T + 1, size=B)
    tg = rng.integers(1, 42, size=(B, Lmax))
    tl = rng.integers(1, Lmax + 1, size=B)
    assert compute_wer(idx, il, tg, tl) == O.compute_wer(idx, il, tg, tl)


# ----------------------------------------------------------------------------------------------------- beam search
def test_beam_golden_small(M, arpa):
    from mddctc.decoder import BeamDecoder
    z, meta = npz("beam_small.npz"), jload("beam_small.json.gz")
    exc = {"IndexError": IndexError, "ValueError": ValueError, "KeyError": KeyError}
    for i, m in enumerate(meta):
        voc = {k: w for k, w in enumerate(m["vocab"])}
        dec = BeamDecoder(voc, beam_width=m["W"], blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=m["alpha"])
        probs = torch.from_numpy(z[str(i)][None])
        if isinstance(m["out"], dict):
            with pytest.raises(exc[m["out"]["exc"]]):
                dec.decode_probs(probs, [m["len"]])
        else:
            assert dec.decode_probs(probs, [m["len"]]) == [m["out"]], i


def test_beam_golden_c4(M, arpa, vocab43):
    from mddctc.decoder import BeamDecoder
    z, meta = npz("beam_c4.npz"), jload("beam_c4.json.gz")
    for alpha in (0.0, 0.3):
        dec = BeamDecoder(vocab43, beam_width=10, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=alpha)
        want = meta["out_alpha_%g" % alpha]
        assert all(isinstance(w, str) for w in want)
        assert dec.decode_probs(torch.from_numpy(z["probs"]), meta["lens"]) == want
        # through the log-prob entry point with CPU tensors (torch CPU exp, the reference's own boundary)
        lp = torch.log(torch.from_numpy(z["probs"])).transpose(0, 1).contiguous()
        if torch.equal(torch.exp(lp.transpose(0, 1)), torch.from_numpy(z["probs"])):
            assert dec.decode(lp, meta["lens"]) == want


def _peaky(rng, B, T, C):
    x = rng.standard_normal((B, T, C)).astype(np.float32)
    u = rng.random((B, T))
    k = rng.integers(1, C, size=(B, T))
    bi, ti = np.nonzero(u < 0.7)
    x[bi, ti, 0] += 8
    bi, ti = np.nonzero(u >= 0.7)
    x[bi, ti, k[bi, ti]] += 6
    return torch.exp(torch.log_softmax(torch.from_numpy(x), -1)).numpy()


@pytest.mark.parametrize("alpha", [0.0, 0.3])
def test_beam_random_vs_oracle(M, arpa, vocab43, alpha):
    """C4-shaped inputs (peaky posteriors, beam 10) + dense random + quantised ties, against the C oracle."""
    from mddctc.decoder import BeamDecoder
    rng = np.random.default_rng(31)
    lm = O.lm_table(arpa, vocab43, 43)
    dec = BeamDecoder(vocab43, beam_width=10, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=alpha)
    probs = _peaky(rng, 96, 500, 43)
    dense = torch.softmax(torch.from_numpy(rng.standard_normal((32, 120, 43)).astype(np.float32)), -1).numpy()
    ties = torch.softmax(torch.from_numpy(np.round(rng.standard_normal((32, 60, 43))).astype(np.float32)), -1).numpy()
    for P in (probs, dense, ties):
        B, T, _ = P.shape
        lens = rng.integers(T // 2, T + 1, size=B)
        toks, olen, st, score = O.beam_tokens(P, lens, 10, 0, lm, alpha)
        t2, l2, s2, sc2 = dec.decode_probs_tokens(torch.from_numpy(P), lens.tolist())
        t2, l2, s2, sc2 = t2.cpu().numpy(), l2.cpu().numpy(), s2.cpu().numpy(), sc2.cpu().numpy()
        assert s2.tolist() == st.tolist()
        assert l2.tolist() == olen.tolist()
        for b in range(B):
            assert t2[b, :l2[b]].tolist() == toks[b, :olen[b]].tolist(), b
        ok = st == 0
        np.testing.assert_allclose(sc2[ok], score[ok], rtol=1e-12)


def test_beam_wide_beam(M, arpa, vocab43):
    """The reference's default beam width (200) on a short utterance."""
    from mddctc.decoder import BeamDecoder
    rng = np.random.default_rng(32)
    lm = O.lm_table(arpa, vocab43, 43)
    P = torch.softmax(torch.from_numpy(rng.standard_normal((3, 25, 43)).astype(np.float32)) * 2, -1).numpy()
    dec = BeamDecoder(vocab43, beam_width=200, blank_index=0, space_idx=-1, lm_path=arpa, lm_alpha=0.01)
    toks, olen, st, _ = O.beam_tokens(P, [25, 25, 25], 200, 0, lm, 0.01)
    t2, l2, s2, _ = dec.decode_probs_tokens(torch.from_numpy(P), [25, 25, 25])
    assert s2.cpu().tolist() == st.tolist() and l2.cpu().tolist() == olen.tolist()
    for b in range(3):
        assert t2[b, :olen[b]].cpu().tolist() == toks[b, :olen[b]].tolist()


# ----------------------------------------------------------------------------------------------------- MDD
def _tok(vocab_inv, s):
    return [vocab_inv[w] for w in s.split()]


def test_mdd_post_golden(M, vocab43):
    """wer -> align_canonical_decoded -> stastics -> score on 600 noisy pairs + the KATs of SURVEY Appendix B."""
    from mddctc.mdd import MDDScorer, align_canonical_decoded, print_aligned_string, stastics
    cases = jload("mdd_post.json.gz")
    inv = {w: k for k, w in vocab43.items()}
    sc = MDDScorer(vocab43)
    dec = [_tok(inv, c["decoded"]) for c in cases]
    can = [_tok(inv, c["canonical"]) for c in cases]
    B = len(cases)
    Ld, Lc = max(max(map(len, dec)), 1), max(max(map(len, can)), 1)
    D = np.zeros((B, Ld), np.int32); Cn = np.zeros((B, Lc), np.int32)
    for b in range(B):
        D[b, :len(dec[b])] = dec[b]; Cn[b, :len(can[b])] = can[b]
    r = sc.align_tokens(torch.from_numpy(D).cuda(), torch.tensor([len(x) for x in dec], dtype=torch.int32).cuda(),
                        torch.from_numpy(Cn).cuda(), torch.tensor([len(x) for x in can], dtype=torch.int32).cuda())
    got = sc.to_python(r)
    for b, c in enumerate(cases):
        want = c["out"]
        if "exc" in want:
            assert isinstance(got[b], TypeError), b
            continue
        for k in ("path0", "dec", "can", "path", "ins", "sub", "del", "correct", "del_sub", "score"):
            assert got[b][k] == want[k], (b, k, c["decoded"], c["canonical"])
        assert list(print_aligned_string(got[b]["dec"], got[b]["can"], got[b]["path"])) == want["printed"]
    # the literal single-utterance call surface of infer.py:335-340
    for c in cases[:40]:
        want = c["out"]
        if "exc" in want:
            continue
        dn = " ".join(w for w in c["decoded"].split(" ") if w != "sil").replace("err", "").replace("  ", " ")
        cn = " ".join(w for w in c["canonical"].split(" ") if w != "sil")
        s1, s2, l = align_canonical_decoded([w for w in dn.split(" ") if w], [w for w in cn.split(" ") if w],
                                            list(want["path0"]))
        assert (s1, s2, l) == (want["dec"], want["can"], want["path"])
        assert stastics(l, s2, s1) == (want["ins"], want["sub"], want["del"])


def test_config1_words_end_to_end(M):
    """BASELINE config 1: greedy decode + MDD alignment of the vocabulary/single word list against the reference run."""
    from mddctc.decoder import GreedyDecoder
    from mddctc.mdd import MDDScorer
    z, js = npz("c1_words.npz"), jload("c1_words.json.gz")
    voc = {k: w for k, w in enumerate(js["vocab"])}
    inv = {w: k for k, w in voc.items()}
    dec = GreedyDecoder(voc, space_idx=-1, blank_index=0)
    lp = torch.from_numpy(z["lp"]).cuda()
    lens = torch.from_numpy(z["lens"])
    assert dec.decode(lp, lens.tolist()) == js["decoded"]
    tokens, out_len, _ = dec.decode_tokens(lp, lens.tolist())
    B = len(js["words"])
    Lc = max(len(c) for c in js["canonical"])
    Cn = np.zeros((B, Lc), np.int32)
    for b, c in enumerate(js["canonical"]):
        Cn[b, :len(c)] = [inv[w] for w in c]
    sc = MDDScorer(voc)
    r = sc.align_tokens(tokens, out_len, torch.from_numpy(Cn).cuda(),
                        torch.tensor([len(c) for c in js["canonical"]], dtype=torch.int32).cuda())
    got = sc.to_python(r)
    for b, want in enumerate(js["mdd"]):
        for k in ("path0", "dec", "can", "path", "ins", "sub", "del", "correct", "del_sub", "score"):
            assert got[b][k] == want[k], (js["words"][b], k)


def test_config5_ragged_greedy_mdd_pipeline_vs_oracle(M, vocab43):
    """BASELINE config 5 as a config: a ragged batch (T 50..1500, 2048 utterances, C = 43 with 'sil' and 'err' among the
    classes), dealt to 8 ranks by mddctc.dist.balanced_buckets; every rank's share goes through greedy decode ->
    sil/err filtering -> Levenshtein alignment -> MDD post-processing on the device and must equal the CPU oracle
    (decoded tokens, distance, op path, aligned rows, fault lists, counts) for every utterance."""
    from mddctc.decoder import GreedyDecoder
    from mddctc.dist import balanced_buckets
    from mddctc.mdd import MDDScorer
    g = torch.Generator().manual_seed(505)
    B, T, C, Lc, world = 2048, 1500, 43, 188, 8
    lens = torch.randint(50, T + 1, (B,), generator=g)
    shards = balanced_buckets(lens.tolist(), world)
    assert sorted(i for s in shards for i in s) == list(range(B))
    frames = [int(lens[s].sum()) for s in shards]
    assert max(frames) - min(frames) <= T                         # equal work to within one utterance
    gd = GreedyDecoder(vocab43, space_idx=-1, blank_index=0)
    sc = MDDScorer(vocab43)
    wid = O.word_ids(vocab43)
    checked = 0
    for rank in (0, 5):                                           # two of the eight shares, every utterance of each
        idx = torch.tensor(shards[rank])
        ln = lens[idx]
        Bs = len(idx)
        gg = torch.Generator().manual_seed(1000 + rank)
        x = torch.randn(T, Bs, C, generator=gg)
        x.scatter_add_(2, torch.randint(0, C, (T, Bs, 1), generator=gg), torch.full((T, Bs, 1), 4.0))
        lp = torch.log_softmax(x, -1)
        can = torch.randint(1, C, (Bs, Lc), generator=gg, dtype=torch.int32)        # 'sil' (2) and 'err' (42) occur
        can_len = torch.clamp((ln + 7) // 8, max=Lc).to(torch.int32)
        tokens, out_len, _ = gd.decode_tokens(lp.cuda(), ln.cuda())
        r = sc.align_tokens(tokens, out_len, can.cuda(), can_len.cuda(), max_dec=T, max_can=Lc)
        got = sc.to_python(r)
        wt, wl, _ = O.greedy_tokens(lp.numpy(), ln.numpy(), 0, wid)
        tk, ol = tokens.cpu().numpy(), out_len.cpu().numpy()
        assert ol.tolist() == wl.tolist()
        dec = [[int(sc.dec_map[k]) for k in wt[b, :wl[b]] if sc.dec_map[k] >= 0] for b in range(Bs)]
        cn = [[int(sc.can_map[k]) for k in can[b, :can_len[b]].tolist() if sc.can_map[k] >= 0] for b in range(Bs)]
        dist, paths, st = O.edit_align(dec, cn)
        for b in range(Bs):
            assert tk[b, :ol[b]].tolist() == wt[b, :wl[b]].tolist()
            if st[b]:
                assert isinstance(got[b], TypeError)
                continue
            w = O.mdd_post(dec[b], cn[b], paths[b])
            gb = got[b]
            assert gb["dist"] == int(dist[b]) and gb["path0"] == paths[b]
            name = sc.name
            assert gb["dec"] == [name(k) for k in w["dec"]] and gb["can"] == [name(k) for k in w["can"]]
            assert gb["path"] == w["ops"]
            assert gb["ins"] == [name(k) for k in w["ins"]] and gb["sub"] == [name(k) for k in w["sub"]]
            assert gb["del"] == [name(k) for k in w["del"]]
            assert (gb["correct"], gb["del_sub"]) == (w["correct"], w["del_sub"])
            checked += 1
    assert checked > 400


# ----------------------------------------------------------------------------------------------------- prefix scorer
def test_prefix_scorer_vs_oracle_and_invariant(M):
    from mddctc.prefix import CTCPrefixScorer
    g = torch.Generator().manual_seed(9)
    T, B, C, W = 60, 3, 12, 4
    x = torch.log_softmax(torch.randn(T, B, C, generator=g), -1)
    lens = torch.tensor([60, 45, 31])
    scorer = CTCPrefixScorer(x.cuda(), lens.cuda(), blank=0, beam=W)
    labels = [[1, 2, 2, 5, 7], [3, 3, 3], [11, 1]]
    # hypothesis w of utterance b follows labels[b] (rotated by w so that the W hypotheses differ)
    state = scorer.initial_state(all_slots_live=True)
    seqs = [[list(np.roll(labels[b], w)) for w in range(W)] for b in range(B)]
    nsteps = max(len(s) for s in labels)
    r_or = [[O.prefix_init(x[:lens[b], b].double().numpy()) for w in range(W)] for b in range(B)]
    last = [[-1] * W for _ in range(B)]
    for n in range(nsteps):
        psi, eos = scorer.score(state)
        psi, eos = psi.cpu().numpy(), eos.cpu().numpy()
        sel_l = torch.zeros((B, W), dtype=torch.int32)
        for b in range(B):
            for w in range(W):
                glen = min(n, len(seqs[b][w]))
                want_psi, want_r, want_eos = O.prefix_score(x[:lens[b], b].double().numpy(), r_or[b][w], last[b][w], glen)
                fin = np.isfinite(want_psi)
                assert np.array_equal(np.isfinite(psi[b, w]), fin)
                np.testing.assert_allclose(psi[b, w][fin], want_psi[fin], rtol=1e-12, atol=1e-12)   # fp64 on both sides
                np.testing.assert_allclose(eos[b, w], want_eos, rtol=1e-12, atol=1e-12)
                c = seqs[b][w][n] if n < len(seqs[b][w]) else None
                sel_l[b, w] = c if c is not None else -1
                if c is not None:
                    r_or[b][w], last[b][w] = want_r[c], c
        state = scorer.extend(state, sel_l.cuda())         # label -1: the slot keeps its state
    psi, eos = scorer.score(state)
    for b in range(B):
        nll = torch.nn.functional.ctc_loss(x[:, b:b + 1].double(), torch.tensor([labels[b]]), lens[b:b + 1],
                                           torch.tensor([len(labels[b])]), reduction="none")
        assert abs(float(eos[b, 0]) + float(nll)) < 1e-10 * max(1.0, abs(float(nll)))


def _synthetic_attention(C, periods=7, seed=0):
    """A stand-in attention decoder (the reference has none): log-probs that depend on the decoding step and on the
    hypothesis' last label, att[b,w,:] = table[step % periods, last + 1]; identical on both sides of a parity test."""
    g = torch.Generator().manual_seed(seed)
    return torch.log_softmax(torch.randn(periods, C + 1, C, generator=g) * 2.0, -1)


@pytest.mark.parametrize("eos", [None, 42])
def test_joint_beam_search_config4_vs_oracle(M, eos):
    """BASELINE config 4 shape (beam 10, CTC weight 0.3, T = 500, C = 43) on 64 utterances: the whole joint beam search
    (prefix scorer + in-library top-W selection + state advance) against the fp64 CPU restatement of the same loop:
    tokens, lengths and flags identical, scores to 1e-9."""
    from mddctc.prefix import joint_beam_search
    g = torch.Generator().manual_seed(44)
    T, B, C, W, steps = 500, 64, 43, 10, 24
    xl = torch.randn(T, B, C, generator=g)
    xl[:, :, 0] += 3.0
    pk = torch.randint(1, C - 1, (T, B), generator=g)
    xl.scatter_add_(2, pk[..., None], 4.0 * (torch.rand(T, B, 1, generator=g) < 0.3).float())
    x = torch.log_softmax(xl, -1)
    lens = torch.randint(300, T + 1, (B,), generator=g)
    lens[0], lens[1] = T, 14                                     # a full-length one and one shorter than the search
    table = _synthetic_attention(C)
    tab_d = table.cuda()

    def att_dev(step, st):
        return tab_d[step % tab_d.shape[0]][(st["last"] + 1).long()]

    def att_cpu(step, st):
        return table[step % table.shape[0]][torch.from_numpy(st["last"] + 1).long()].numpy()

    tok, sc, gl, fl = joint_beam_search(x.cuda(), lens.cuda(), att_dev, beam=W, ctc_weight=0.3, max_len=steps, eos=eos)
    wt, ws, wg, wf = O.joint_beam_search(x.numpy(), lens.numpy(), att_cpu, beam=W, ctc_weight=0.3, max_len=steps, eos=eos)
    assert np.array_equal(fl.cpu().numpy(), wf)
    assert np.array_equal(gl.cpu().numpy(), wg)
    assert np.array_equal(tok.cpu().numpy(), wt)
    got = sc.cpu().numpy()
    fin = np.isfinite(ws)
    assert np.array_equal(np.isfinite(got), fin)
    np.testing.assert_allclose(got[fin], ws[fin], rtol=1e-9, atol=1e-9)
    assert bool((np.diff(np.where(fin, ws, -1e300), axis=1) <= 0).all())          # best first
    if eos is not None:
        assert int((wf == 1).sum()) > 0                                          # some hypotheses did end


# ----------------------------------------------------------------------------------------------------- corpus MDD evaluation
CORPUS_KEYS = ["true_accept", "false_rejection", "false_accept", "true_rejection_correct_diagnose",
               "true_rejection_wrong_diagnose", "total_phonemes_in_canonical"]


def test_corpus_mdd_vs_reference_golden(M):
    """test_ctc_nosil.py:218-343 executed by the real reference on 500 triples (tests/golden/corpus_mdd.json.gz):
    per-utterance counters, m_speaker block, PER numerators and corpus totals, bit-exact."""
    from mddctc.corpus import CorpusMDD
    data = jload("corpus_mdd.json.gz")
    ev = CorpusMDD({0: "blank"}, m_speaker=data["m_speaker"])
    r = ev.evaluate_strings(data["utt"], data["labels"], data["canonicals"], data["decoded"])
    counts = r["counts"].cpu().numpy()
    for b, ref in enumerate(data["per_utt"]):
        want, c = ref["counts"], counts[b].tolist()
        assert c[:6] == [want[k] for k in CORPUS_KEYS], (b, c, want)
        assert c[9] == want["total_wer"] and c[8] == want["num_word"]
        if data["utt"][b][:4] in data["m_speaker"]:
            assert [c[0], c[6], c[7], c[3], c[4], c[5]] == [want["m_" + k] for k in CORPUS_KEYS], (b, c, want)
        n = int(r["lc"]["ops_len"][b])
        assert bytes(r["lc"]["ops"][b, :n].cpu().numpy()).decode() == ref["lc"]
    s = ev.summary()
    tot = data["total"]
    for k in CORPUS_KEYS + ["total_wer", "num_word"]:
        assert s[k] == tot[k], k
    for k in CORPUS_KEYS:
        assert s["m_" + k] == tot["m_" + k], k
    assert s["m_total_wer"] == tot["m_wer"] and s["m_num_word"] == tot["m_decoder_num"]
    tr = tot["true_rejection_correct_diagnose"] + tot["true_rejection_wrong_diagnose"]
    assert s["precision"] == float(tr) / (tr + tot["false_rejection"]) * 100
    assert s["recall"] == float(tr) / (tr + tot["false_accept"]) * 100
    with pytest.raises(TypeError):
        ev.evaluate_strings(["X"], ["sil"], ["a b"], ["a"])


def test_corpus_mdd_random_vs_oracle(M):
    """Class-id rows with 'sil' inside, long rows, bursts of insertions at one canonical position."""
    from mddctc.corpus import CorpusMDD
    rng = np.random.default_rng(21)
    vocab = {i: ("sil" if i == 2 else "p%d" % i) for i in range(12)}
    ev = CorpusMDD(vocab)
    for (B, Lc, V) in ((257, 12, 5), (64, 150, 11), (33, 400, 4)):
        def noisy(rows, p):
            out = []
            for r in rows:
                o = []
                for w in r:
                    u = rng.random()
                    if u < p:
                        continue
                    o.append(int(rng.integers(1, V + 1)) if u < 2 * p else w)
                    while rng.random() < p:
                        o.append(int(rng.integers(1, V + 1)))
                out.append(o or [3])
            return out
        can = [[int(x) for x in rng.integers(1, V + 1, size=int(rng.integers(1, Lc + 1)))] for _ in range(B)]
        for r in can:
            if all(w == 2 for w in r):
                r[0] = 3
        lab = noisy(can, 0.12)
        dec = noisy(lab, 0.15)
        for rows in (lab, dec):
            for r in rows:
                if all(w == 2 for w in r):
                    r.append(4)
        grp = (rng.random(B) < 0.4).astype(np.uint8)

        def pack(rows):
            n = max(len(r) for r in rows)
            A = np.zeros((B, n), dtype=np.int32)
            for i, r in enumerate(rows):
                A[i, :len(r)] = r
            return torch.from_numpy(A).cuda(), torch.tensor([len(r) for r in rows], dtype=torch.int32).cuda()
        (D, dl), (Lb, ll), (Cn, cl) = pack(dec), pack(lab), pack(can)
        ev.reset()
        r = ev.evaluate_tokens(D, dl, Lb, ll, Cn, cl, group=torch.from_numpy(grp).cuda())
        got = r["counts"].cpu().numpy()
        assert int(r["status"].abs().sum()) == 0
        f = lambda rows: [[w for w in q if w != 2] for q in rows]
        fl, fc, fd = f(lab), f(can), f(dec)
        _, lc, _ = O.edit_align(fl, fc)
        dd, _, _ = O.edit_align(fd, fl)
        _, dc, _ = O.edit_align(fd, fc)
        tot = np.zeros((2, 12), dtype=np.int64)
        for b in range(B):
            c = O.corpus_counts(fl[b], lc[b], fd[b], dc[b]) + [len(fl[b]), int(dd[b]), 0, 0]
            assert got[b].tolist() == c, (b, got[b].tolist(), c)
            tot[0] += c
            if grp[b]:
                tot[1] += [c[0], c[6], c[7]] + c[3:]
        assert np.array_equal(ev.totals, tot)
