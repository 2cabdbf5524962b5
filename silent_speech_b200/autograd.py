"""Backward orchestration: loss.backward() of train_model_official.py:437 over the C-ABI kernels.

``backward_impl`` walks the stages of ``functional.forward_impl`` in reverse and ACCUMULATES
(+=) every parameter gradient into caller-provided tensors (a dict keyed by the reference's
state-dict names -- in training these are views of one flat bucket that is then all-reduced
and fed to the fused clip+Adam kernel).  The serial parts (BPTT, pooled head, fused CNN
backward) are dedicated kernels; everything of GEMM shape (dW = dY^T X, dX = dY W) goes
through the strided GEMM entry point.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import ops
from .functional import G3, H, _passes, forward_impl


def _split_k(K: int) -> int:
    return max(1, min(64, K // 512))


def _split_k_tc(rows_out: int, cols_out: int, K: int) -> int:
    """K splits of a tensor-core weight-gradient GEMM: about two 128x192-tile CTAs per SM in flight, >= 4 chunks each."""
    tiles = -(-rows_out // 128) * -(-cols_out // 192)
    return max(1, min(K // 128, (2 * 148 + tiles - 1) // tiles))


def _wgrad(dY2d, y_col0, y_cols, ldy, X2d, x_col0, x_cols, ldx, gW, gw_off, ldw, rows, amax=None):
    """gW[row, :] += sum_m dY[m, y_col0+row] * X[m, x_col0 + :]   (rows = y_cols)."""
    if amax is not None:          # tensor-core modes; amax = max|dY| (device scalar) scales dY into fp16's range
        ops.gemm_tcx(dY2d, X2d, gW, y_cols, x_cols, rows, sa=(1, ldy), sb=(ldx, 1), ldc=ldw, accumulate=True,
                     split_k=_split_k_tc(y_cols, x_cols, rows), a_off=y_col0, b_off=x_col0, c_off=gw_off, a_amax=amax)
        return
    ops.sgemm(dY2d, X2d, gW, y_cols, x_cols, rows, sa=(1, ldy), sb=(ldx, 1), ldc=ldw, accumulate=True,
              split_k=_split_k(rows), a_off=y_col0, b_off=x_col0, c_off=gw_off)


_SIDE_STREAMS: Dict[int, "torch.cuda.Stream"] = {}


def _side_stream(device) -> "torch.cuda.Stream":
    """One auxiliary stream per device for work that is off the backward pass's critical path."""
    idx = torch.device(device).index
    idx = torch.cuda.current_device() if idx is None else idx
    if idx not in _SIDE_STREAMS:
        _SIDE_STREAMS[idx] = torch.cuda.Stream(idx)
    return _SIDE_STREAMS[idx]


def _gru_layer_backward(layer: int, dseq, seq_out, gates, lengths, B, T, inputs, model, grads, want_dinput, side=None):
    """inputs: list of (tensor2d [B*T, width], column offset inside weight_ih) feeding this layer.

    Order: BPTT, then the data gradient (the next stage waits for it), then the weight / bias gradients.  With ``side``
    (a CUDA stream) the latter are enqueued there: nothing downstream reads them before the optimizer, and the next
    stage (the other layer's BPTT: 128 latency-bound CTAs) leaves room on every SM for the GEMM CTAs to run beside it."""
    p = model.gru.layer(layer)
    tc = _passes(model)[1] != 0          # tensor-core modes: GEMM-shaped gradients on tcgen05 (split-fp16 x3)
    whhT = [ops.transpose(p["weight_hh"][d]) for d in (0, 1)]
    dgi, dghn, hprev = ops.gru_layer_bwd(dseq, seq_out, gates, whhT[0], whhT[1], lengths)
    M = B * T
    dgi2d, dghn2d, hprev2d = dgi.view(M, 2 * G3), dghn.view(M, 2 * H), hprev.view(M, 2 * H)
    amax = ops.absmax(dgi) if tc else None          # |dghn| = |dgi_n * r| <= |dgi_n|: one scale serves both tensors
    dinput = None
    if want_dinput is not None:
        col0, ncols = want_dinput
        dinput = torch.empty(M, ncols, device=dgi.device, dtype=torch.float32)
        for d in (0, 1):
            W = p["weight_ih"][d]
            kw = dict(sa=(2 * G3, 1), sb=(W.shape[1], 1), ldc=ncols, accumulate=(d == 1), lengths=lengths, T=T,
                      a_off=d * G3, b_off=col0)
            if tc:
                ops.gemm_tcx(dgi2d, W, dinput, M, ncols, G3, a_amax=amax, **kw)
            else:
                ops.sgemm(dgi2d, W, dinput, M, ncols, G3, **kw)

    def weight_grads():
        for d, sfx in ((0, ""), (1, "_reverse")):
            g_wih, g_whh = grads[f"gru.weight_ih_l{layer}{sfx}"], grads[f"gru.weight_hh_l{layer}{sfx}"]
            g_bih, g_bhh = grads[f"gru.bias_ih_l{layer}{sfx}"], grads[f"gru.bias_hh_l{layer}{sfx}"]
            width = g_wih.shape[1]
            for x2d, col0 in inputs:
                _wgrad(dgi2d, d * G3, G3, 2 * G3, x2d, 0, x2d.shape[1], x2d.stride(0), g_wih, col0, width, M, amax)
            ops.colsum(dgi2d[:, d * G3:(d + 1) * G3], g_bih)
            # hidden projection: r,z rows share dgi; the n rows use dghn = dgi_n * r
            _wgrad(dgi2d, d * G3, 2 * H, 2 * G3, hprev2d, d * H, H, 2 * H, g_whh, 0, H, M, amax)
            _wgrad(dghn2d, d * H, H, 2 * H, hprev2d, d * H, H, 2 * H, g_whh, 2 * H * H, H, M, amax)
            ops.colsum(dgi2d[:, d * G3:d * G3 + 2 * H], g_bhh[:2 * H])
            ops.colsum(dghn2d[:, d * H:(d + 1) * H], g_bhh[2 * H:])

    if side is None:
        weight_grads()
    else:
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(dgi.device))   # BPTT outputs, amax and everything enqueued before them
        with torch.cuda.stream(side):
            side.wait_event(ready)
            weight_grads()
        for t in (dgi, dghn, hprev, amax):           # freed on return, still being read on the side stream
            if t is not None:
                t.record_stream(side)
    return dinput


def backward_impl(model, ctx: Dict, grads: Dict[str, torch.Tensor], dlogits: Optional[torch.Tensor] = None,
                  y: Optional[torch.Tensor] = None, label_smoothing: float = 0.0, grad_scale: float = 1.0):
    """Either ``dlogits`` (autograd path) or ``y`` (+ label_smoothing, grad_scale: fused CE) must be given."""
    B, T, lengths = ctx["B"], ctx["T"], ctx["lengths"]
    attn, pooled, xhat, rstd, ln, hid = ctx["head_saved"]
    hw = model.head_tensors()
    C = model.num_classes
    dlog, dhid = ops.head_bwd_logits(ctx["logits"] if dlogits is None else None, dlogits, y, label_smoothing,
                                     grad_scale, hw[6], hid, ctx["head_p"])
    # head.4 / head.1 weight gradients and dln as GEMMs
    ops.sgemm(dlog, hid, grads["head.4.weight"], C, 128, B, sa=(1, C), sb=(128, 1), ldc=128, accumulate=True)
    ops.colsum(dlog, grads["head.4.bias"])
    ops.sgemm(dhid, ln, grads["head.1.weight"], 128, 2 * H, B, sa=(1, 128), sb=(2 * H, 1), ldc=2 * H, accumulate=True)
    ops.colsum(dhid, grads["head.1.bias"])
    dln = torch.empty(B, 2 * H, device=dhid.device, dtype=torch.float32)
    ops.sgemm(dhid, hw[4], dln, B, 2 * H, 128, sa=(128, 1), sb=(2 * H, 1), ldc=2 * H)
    dseq1 = ops.head_bwd_pool(dln, xhat, rstd, attn, ctx["seq1"], lengths, hw[2], hw[0],
                              grads["head.0.weight"], grads["head.0.bias"], grads["pool.score.weight"],
                              grads["pool.score.bias"])
    # layer 1
    M = B * T
    side = _side_stream(dhid.device) if _passes(model)[1] != 0 else None
    dseq0 = _gru_layer_backward(1, dseq1, ctx["seq1"], ctx["gates1"], lengths, B, T,
                                [(ctx["seq0_in"].view(M, 2 * H), 0)], model, grads, want_dinput=(0, 2 * H), side=side)
    if ctx["gru_p"] > 0:
        ops.dropout(dseq0, ctx["gru_p"], ctx["drop_seed"], out=dseq0)      # same mask, same 1/(1-p)
    # layer 0
    X2d = ctx["X"].view(M, -1)
    x_dim = X2d.shape[1]
    inputs = [(X2d, 0)]
    emb = ctx.get("emb")
    if emb is not None:
        inputs.append((emb.view(M, -1), x_dim))
    d_emb = _gru_layer_backward(0, dseq0.view(B, T, 2 * H), ctx["seq0"], ctx["gates0"], lengths, B, T, inputs, model,
                                grads, want_dinput=(x_dim, emb.shape[-1]) if emb is not None else None, side=side)
    if emb is not None:
        names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
                 "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
        ops.roi_cnn_bwd(ctx["R"], lengths, model.roi_cnn.tensors(), model.standardize_roi, d_emb.view(B, T, -1), 0,
                        [grads[n] for n in names], saved=ctx.get("cnn_saved"))
    if side is not None:
        torch.cuda.current_stream(dhid.device).wait_stream(side)      # all gradients complete before anything that follows


class ClassifierFn(torch.autograd.Function):
    """model(X, lengths, R) as one autograd node, so that ``loss_fn(logits, y); loss.backward()``
    (train_model_official.py:433-437) fills every parameter's .grad."""

    @staticmethod
    def forward(fctx, model, names, X, lengths, R, *params):
        out, ctx = forward_impl(model, X, lengths, R, training=model.training, save=True)
        fctx.model, fctx.names, fctx.ctx = model, names, ctx
        return out["logits"]

    @staticmethod
    def backward(fctx, dlogits):
        model, names = fctx.model, fctx.names
        with torch.cuda.device(dlogits.device):
            grads = {n: torch.zeros_like(p) for n, p in model.named_parameters()}
            backward_impl(model, fctx.ctx, grads, dlogits=dlogits.contiguous().float())
        fctx.ctx = None
        return (None, None, None, None, None) + tuple(grads[n] for n in names)
