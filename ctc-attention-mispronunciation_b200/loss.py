"""nn.CTCLoss-compatible module on the sm_100a kernels.

Drop-in for the call sites egs/attention_aug/steps/train_ctc.py:186,72,96 and egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67:

    loss_fn = CTCLoss(reduction='sum')
    loss = loss_fn(out, targets, input_sizes, target_sizes); loss /= batch_size; loss.backward()

Semantics follow torch.nn.functional.ctc_loss: blank/reduction/zero_infinity, (T,B,C) or
(T,C) log-probs, 2-D padded or 1-D concatenated targets, lengths as tensors (any device) or sequences, 'mean' dividing
by clamp_min(target_length, 1), inf loss / NaN gradient for infeasible alignments unless zero_infinity.

The gradient is produced in the same launch as the loss (one pass over the log-probs); autograd's backward only scales
it by the upstream gradient.  Compute dtype is fp32 (fp64 inputs are accepted and computed in fp32).
"""
from __future__ import annotations

import torch

from . import _lib, _rt


def _as_len_tensor(x, dev):
    if isinstance(x, torch.Tensor):
        return x.to(device=dev, dtype=torch.int64).contiguous(), (x if not x.is_cuda else None)
    t = torch.as_tensor(list(x), dtype=torch.int64)
    return t.to(dev), t


def ctc_loss_raw(log_probs, targets, input_lengths, target_lengths, blank=0, zero_infinity=False, want_grad=True,
                 grad_out=None, grad_scalar=1.0, algo=0, validate=True, loss_sum=None, nll=None, grad=None,
                 status=None, logits=False, row_lse=None, argmax=None):
    """One call into mdd_ctc_loss.  Returns (nll[B], grad[T,B,C] or None, status[B]); everything stays on the device.

    With ``logits=True`` the first argument holds the output layer's logits and the call goes to mdd_ctc_loss_logits:
    ``grad`` is then the gradient with respect to the logits, ``row_lse`` (T,B) fp32 receives the per-frame
    log-sum-exp and ``argmax`` (T,B) uint8, when given, the frame arg-max."""
    L = _lib.lib()
    if log_probs.dim() == 2:
        log_probs = log_probs.unsqueeze(1)
    if log_probs.dim() != 3:
        raise RuntimeError("log_probs must be (T,B,C) or (T,C)")
    dev = _rt.device(log_probs)
    lp = log_probs.to(device=dev, dtype=torch.float32)
    if lp.stride(2) != 1:
        lp = lp.contiguous()
    T, B, C = lp.shape
    if not (0 <= blank < C):
        raise RuntimeError("blank must be in label range")
    il, il_host = _as_len_tensor(input_lengths, dev)
    tl, tl_host = _as_len_tensor(target_lengths, dev)
    if il.numel() != B or tl.numel() != B:
        raise RuntimeError("input_lengths and target_lengths must have size batch_size")
    tg = targets.to(device=dev, dtype=torch.int64).contiguous()
    if tg.dim() == 2:
        if tg.size(0) != B:
            raise RuntimeError("targets must be (B, Lmax) or 1-D")
        lmax, tstride = tg.size(1), tg.size(1)
    elif tg.dim() == 1:
        # concatenated targets: the padding width is not known on the host without reading the lengths
        src = tl_host if tl_host is not None else tl.cpu()
        lmax, tstride = (int(src.max()) if B else 0), 0
        tl_host = src
    else:
        raise RuntimeError("targets must be 1-D or 2-D")
    if validate:
        # torch validates on the host (aten reads the lengths); do the same when they are host-visible for free,
        # otherwise after the launch through the status vector.
        if il_host is not None and B and (int(il_host.max()) > T or int(il_host.min()) < 0):
            raise RuntimeError("Expected input_lengths to have value at most %d and be non-negative" % T)
        if tl_host is not None and B:
            if int(tl_host.min()) < 0 or (tg.dim() == 2 and int(tl_host.max()) > lmax):
                raise RuntimeError("Expected target_lengths to be non-negative and at most the padded target width")
            if tg.dim() == 1 and int(tl_host.sum()) > tg.numel():
                raise RuntimeError("Expected the concatenated targets to hold sum(target_lengths) labels")
    nll = torch.empty(B, dtype=torch.float32, device=dev) if nll is None else nll
    if want_grad and grad is None:
        grad = torch.empty((T, B, C), dtype=torch.float32, device=dev)
    status = torch.empty(B, dtype=torch.int32, device=dev) if status is None else status
    need = _lib.C.c_size_t(0)
    _lib.check(L.mdd_ctc_loss_workspace_bytes(T, B, C, lmax, algo, _lib.C.byref(need)), "mdd_ctc_loss_workspace_bytes")
    ws = _rt.workspace(need.value, dev, "ctc")
    go = None if grad_out is None else grad_out.to(device=dev, dtype=torch.float32).contiguous()
    if logits:
        if row_lse is None:
            row_lse = torch.empty((T, B), dtype=torch.float32, device=dev)
        rc = L.mdd_ctc_loss_logits(lp.data_ptr(), T, B, C, lp.stride(0), lp.stride(1), _rt.ptr(tg), tstride,
                                   il.data_ptr(), tl.data_ptr(), lmax, blank, int(bool(zero_infinity)), _rt.ptr(go),
                                   float(grad_scalar), nll.data_ptr(), _rt.ptr(grad if want_grad else None),
                                   _rt.ptr(loss_sum), status.data_ptr(), row_lse.data_ptr(), _rt.ptr(argmax), algo,
                                   ws.data_ptr(), ws.numel(), _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_ctc_loss_logits")
    else:
        rc = L.mdd_ctc_loss(lp.data_ptr(), T, B, C, lp.stride(0), lp.stride(1), _rt.ptr(tg), tstride, il.data_ptr(),
                            tl.data_ptr(), lmax, blank, int(bool(zero_infinity)), _rt.ptr(go), float(grad_scalar),
                            nll.data_ptr(), _rt.ptr(grad if want_grad else None), _rt.ptr(loss_sum),
                            status.data_ptr(), algo, ws.data_ptr(), ws.numel(), _rt.stream_ptr(dev))
        _lib.check(rc, "mdd_ctc_loss")
    if validate and (il_host is None or tl_host is None):
        if bool((status != 0).any()):
            raise RuntimeError("ctc_loss: input_lengths/target_lengths/targets out of range (status %s)"
                               % status.nonzero().flatten()[:8].tolist())
    return nll, (grad if want_grad else None), status


def ctc_loss_streamed(log_probs_host, targets, input_lengths, target_lengths, slices=4, blank=0, zero_infinity=False,
                      grad_scalar=1.0, staging=None, algo=0):
    """CTC loss + gradient of a batch whose log-probs live in (page-locked) HOST memory, as the reference keeps them for
    decoding (infer.py:300): slice k+1 of the batch crosses PCIe on a side stream while the kernels of slice k run.
    Targets and lengths are device tensors.

    ``log_probs_host`` is either
      * a LIST of contiguous fp32 (T, B_k, C) host tensors -- slice-major, the layout a producer that emits the batch in
        micro-batches has anyway: every transfer is ONE contiguous copy; or
      * one contiguous fp32 (T, B, C) host tensor, cut into `slices` batch slices (each transfer is then a strided 2-D
        copy of T rows, mdd_copy_h2d_2d).
    Returns (loss_sum [1] device, nll [B] device, [grad slices (T, B_k, C)]); `staging` optionally holds the device
    buffers that receive the log-probs between steps (a (T,B,C) tensor for the single-tensor form, a list of
    (T,B_k,C) tensors for the list form)."""
    L = _lib.lib()
    dev = _rt.device(targets)
    cur, side = torch.cuda.current_stream(dev), _rt.copy_stream(dev)
    side.wait_stream(cur)                                   # the staging buffers may still be read by earlier kernels
    events, views, bounds = [], [], []
    if isinstance(log_probs_host, (list, tuple)):
        lo = 0
        for k, h in enumerate(log_probs_host):
            if h.is_cuda or h.dim() != 3 or h.dtype != torch.float32 or not h.is_contiguous():
                raise RuntimeError("ctc_loss_streamed takes contiguous fp32 (T,B_k,C) host tensors")
            d = staging[k] if staging is not None else torch.empty(h.shape, dtype=torch.float32, device=dev)
            with torch.cuda.stream(side):
                d.copy_(h, non_blocking=True)               # one contiguous cudaMemcpyAsync
            ev = torch.cuda.Event()
            ev.record(side)
            events.append(ev)
            views.append(d)
            bounds.append((lo, lo + h.shape[1]))
            lo += h.shape[1]
        B = lo
    else:
        if log_probs_host.is_cuda or log_probs_host.dim() != 3 or log_probs_host.dtype != torch.float32 \
                or not log_probs_host.is_contiguous():
            raise RuntimeError("ctc_loss_streamed takes a contiguous fp32 (T,B,C) host tensor")
        T, B, C = log_probs_host.shape
        lp_d = torch.empty((T, B, C), dtype=torch.float32, device=dev) if staging is None else staging
        slices = max(1, min(int(slices), B))
        base, rem = divmod(B, slices)
        lo = 0
        for k in range(slices):
            hi = lo + base + (1 if k < rem else 0)
            bounds.append((lo, hi))
            lo = hi
        pitch = B * C * 4
        for lo, hi in bounds:
            _lib.check(L.mdd_copy_h2d_2d(lp_d.data_ptr() + lo * C * 4, pitch, log_probs_host.data_ptr() + lo * C * 4,
                                         pitch, (hi - lo) * C * 4, T, side.cuda_stream), "mdd_copy_h2d_2d")
            ev = torch.cuda.Event()
            ev.record(side)
            events.append(ev)
            views.append(lp_d[:, lo:hi])
    parts = torch.zeros(len(bounds), dtype=torch.float32, device=dev)
    nll = torch.empty(B, dtype=torch.float32, device=dev)
    grads = []
    for k, (lo, hi) in enumerate(bounds):
        cur.wait_event(events[k])
        _, g, _ = ctc_loss_raw(views[k], targets[lo:hi], input_lengths[lo:hi], target_lengths[lo:hi], blank=blank,
                               zero_infinity=zero_infinity, grad_scalar=grad_scalar, algo=algo, validate=False,
                               loss_sum=parts[k:k + 1], nll=nll[lo:hi])
        grads.append(g)
    return parts.sum(dim=0, keepdim=True), nll, grads


def last_fallback_flags(B, dev=None):
    """Diagnostics: the per-utterance flags the last mdd_ctc_loss call on this device left in its workspace
    (0 = the packed linear-domain kernel produced the result -- in its first pass or, with algo=0, in its second tier at
    2 frames per chunk -- otherwise the fp64 kernel redid the utterance; after an algo=2 call: what the first pass
    alone handed on)."""
    dev = _rt.device() if dev is None else dev
    ws = _rt._ws_cache[(dev.index, "ctc")]
    off = (B * 8 + 255) // 256 * 256
    return ws[off:off + 4 * B].view(torch.int32).clone()


class _CTCLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, log_probs, targets, input_lengths, target_lengths, blank, zero_infinity, algo, validate,
                logits=False, row_lse=None, argmax=None):
        want = ctx.needs_input_grad[0]
        nll, grad, _ = ctc_loss_raw(log_probs.detach(), targets, input_lengths, target_lengths, blank, zero_infinity,
                                    want_grad=want, algo=algo, validate=validate, logits=logits, row_lse=row_lse,
                                    argmax=argmax)
        ctx.grad_buf = grad
        ctx.in_shape = tuple(log_probs.shape)
        ctx.in_dtype = log_probs.dtype
        ctx.in_device = log_probs.device
        return nll

    @staticmethod
    def backward(ctx, g):
        grad = ctx.grad_buf
        if grad is None:
            raise RuntimeError("CTCLoss: the gradient buffer of this forward was already consumed "
                               "(backward through the same graph twice is not supported)")
        ctx.grad_buf = None
        T, B, C = grad.shape
        dev = grad.device
        scale = g.to(device=dev, dtype=torch.float32).expand(B).contiguous()
        L = _lib.lib()
        _lib.check(L.mdd_ctc_scale_grad(grad.data_ptr(), T, B, C, scale.data_ptr(), 1.0, _rt.stream_ptr(dev)),
                   "mdd_ctc_scale_grad")
        out = grad
        if len(ctx.in_shape) == 2:
            out = out.squeeze(1)
        out = out.to(device=ctx.in_device, dtype=ctx.in_dtype)
        return out, None, None, None, None, None, None, None, None, None, None


def ctc_loss(log_probs, targets, input_lengths, target_lengths, blank=0, reduction="mean", zero_infinity=False,
             algo=0, validate=True, _logits=False, _row_lse=None, _argmax=None):
    """torch.nn.functional.ctc_loss-compatible functional form."""
    if reduction not in ("none", "mean", "sum"):
        raise ValueError("%s is not a valid value for reduction" % reduction)
    nll = _CTCLossFn.apply(log_probs, targets, input_lengths, target_lengths, int(blank), bool(zero_infinity),
                           int(algo), bool(validate), bool(_logits), _row_lse, _argmax)
    if log_probs.dim() == 2 and reduction == "none":
        nll_out = nll.squeeze(0)
    else:
        nll_out = nll
    if reduction == "none":
        out = nll_out
    elif reduction == "sum":
        out = nll.sum()
    else:
        tl = target_lengths if isinstance(target_lengths, torch.Tensor) else torch.as_tensor(list(target_lengths))
        tl = tl.to(device=nll.device, dtype=nll.dtype).clamp_min(1)
        out = (nll / tl).mean()
    if log_probs.dtype != torch.float32:
        out = out.to(log_probs.dtype)
    if not log_probs.is_cuda:
        out = out.to(log_probs.device)
    return out


class CTCLoss(torch.nn.Module):
    """Same constructor and call signature as torch.nn.CTCLoss (train_ctc.py:186 builds it with reduction='sum')."""

    def __init__(self, blank: int = 0, reduction: str = "mean", zero_infinity: bool = False, algo: int = 0,
                 validate: bool = True):
        super().__init__()
        self.blank = blank
        self.reduction = reduction
        self.zero_infinity = zero_infinity
        self.algo = algo
        self.validate = validate

    def forward(self, log_probs, targets, input_lengths, target_lengths):
        return ctc_loss(log_probs, targets, input_lengths, target_lengths, self.blank, self.reduction,
                        self.zero_infinity, self.algo, self.validate)


def ctc_loss_from_logits(logits, targets, input_lengths, target_lengths, blank=0, reduction="mean",
                         zero_infinity=False, algo=0, validate=True, row_lse=None, argmax=None):
    """``F.ctc_loss(F.log_softmax(logits, -1), ...)`` in one step.

    Replaces the tail of the reference models' forward (``out = F.log_softmax(out, dim=-1)``,
    models/model_ctc.py:214-218, cnn_rnn.py:168-172) together with the loss call at steps/train_ctc.py:72: the
    log-probability array is never materialised and autograd receives the gradient with respect to the logits
    directly, (softmax - occupancy) * upstream -- the value torch reaches by back-propagating its CTC gradient through
    log_softmax.  ``row_lse`` (T,B) fp32 / ``argmax`` (T,B) uint8, when given, are filled by the same pre-pass
    (log_probs = logits - row_lse[..., None]; argmax feeds the train-time PER of train_ctc.py:77-81)."""
    return ctc_loss(logits, targets, input_lengths, target_lengths, blank, reduction, zero_infinity, algo, validate,
                    _logits=True, _row_lse=row_lse, _argmax=argmax)


class CTCLossFromLogits(CTCLoss):
    """CTCLoss taking the output layer's logits instead of log-probs (see ctc_loss_from_logits)."""

    def forward(self, logits, targets, input_lengths, target_lengths, row_lse=None, argmax=None):
        return ctc_loss_from_logits(logits, targets, input_lengths, target_lengths, self.blank, self.reduction,
                                    self.zero_infinity, self.algo, self.validate, row_lse, argmax)


def row_lse(logits, want_argmax=False):
    """Per-frame log-sum-exp (and optionally arg-max) of (T,B,C) logits: the pre-pass of ctc_loss_from_logits."""
    L = _lib.lib()
    dev = _rt.device(logits)
    x = logits.to(device=dev, dtype=torch.float32)
    if x.dim() == 2:
        x = x.unsqueeze(1)
    if x.stride(2) != 1:
        x = x.contiguous()
    T, B, C = x.shape
    lse = torch.empty((T, B), dtype=torch.float32, device=dev)
    am = torch.empty((T, B), dtype=torch.uint8, device=dev) if want_argmax else None
    _lib.check(L.mdd_row_lse(x.data_ptr(), T, B, C, x.stride(0), x.stride(1), lse.data_ptr(), _rt.ptr(am),
                             _rt.stream_ptr(dev)), "mdd_row_lse")
    return (lse, am) if want_argmax else lse
