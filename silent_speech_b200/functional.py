"""Forward / backward orchestration of the classifier over the C-ABI kernels.

Stage order follows BiGRUClassifier.forward (train_model_official.py:279-310):
  K2 ROI CNN -> K3a input projections (GEMM, both directions stacked into one [B,T,1152]
  buffer; the concat of train:297 is never materialised -- X and the ROI embedding are two
  K-slices of the same projection) -> K3b recurrence, layer 0 -> (dropout 0.1, train only)
  -> K3a/K3b layer 1 -> K4 pooled head.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops

H = ops.HIDDEN
G3 = 3 * H


def _passes(model):
    """(CNN passes, GRU passes) of a precision mode: 0 = fp32 CUDA cores, 1 = single fp16 pass, 3 = hi/lo split."""
    return {"fp32": (0, 0), "tc3": (3, 3), "fp16": (1, 1), "cnn16": (1, 3), "gru16": (3, 1)}[getattr(model, "precision", "fp32")]


# Inference projections: splitting the activations too is the default.  Rounding them to fp16 once (ssb_proj_gemm_wsplit:
# 128-row tiles, two MMAs per k-step; model.proj_full_split = False or a tuple of the layers that keep the full split) makes
# the two projections 0.7 ms faster per 4096 clips but costs about 1e-4 of max |logit| PER LAYER (tests/probes/proj_probe.py:
# 1.6e-4 .. 2.5e-4 with both) -- the recurrence carries a gi error through 90 steps, unlike the ROI CNN, whose activations
# can be rounded for free (cnn_passes).  That misses the 1e-4 bar of the tc3 mode, so it stays opt-in.
PROJ_FULL_SPLIT_DEFAULT = True


def cnn_passes(model, save: bool = False) -> int:
    """`passes` of the ROI-CNN kernel for this call.  In the fp32-parity mode (tc3) the training forward (save) runs the full
    split (3: weights and activations as fp16 hi + lo; the saved activations are fp32); inference runs 2: split weights,
    activations rounded to fp16 once -- K2 12.0 ms instead of 16.0 ms per 4096 clips, logits still within 1e-4 of the CPU
    reference (tests/test_gpu_forward.py, bench.py's parity block).  model.cnn_full_split = True forces 3 everywhere."""
    p = _passes(model)[0]
    if p == 3 and not save and not getattr(model, "cnn_full_split", False):
        return 2
    return p


class Workspace:
    """Per-call buffers (allocated by PyTorch's caching allocator; the library owns nothing)."""

    def __init__(self):
        self.saved = {}


def _project(A2d, w_pair, b_pair, gi2d, lengths, T, extra=None, x_dim=None):
    """gi[:, d*576:(d+1)*576] = A @ W_ih_d^T + b_ih_d (+ E @ W_ih_d[:, x_dim:]^T) for d = fwd, rev."""
    for d in (0, 1):
        W, b = w_pair[d], b_pair[d]
        k = A2d.shape[1]
        ops.linear_nt(A2d, W, b, gi2d, lengths=lengths, T=T, k=k, c_col0=d * G3)
        if extra is not None:
            ops.linear_nt(extra, W, None, gi2d, accumulate=True, lengths=lengths, T=T,
                          w_col0=k, k=extra.shape[1], c_col0=d * G3)


def forward_impl(model, X, lengths, R, training: bool, save: bool, k_top: int = 0, y=None,
                 label_smoothing: float = 0.0, emb=None):
    """emb (optional, inference): precomputed ROI-CNN embeddings f32 [B,T,32] used instead of running K2 on R --
    the streaming scheduler caches them per frame."""
    dev = X.device
    if not X.is_cuda:
        raise ops._lib.SsbError("inputs must be CUDA tensors: this package has no CPU path")
    X = X.contiguous().float()
    lengths = lengths.to(device=dev, dtype=torch.int64).contiguous()
    B, T, x_dim = X.shape
    ctx = {"B": B, "T": T, "lengths": lengths, "X": X}
    if emb is not None:
        emb = emb.contiguous().float()
        ctx["emb"] = emb
    elif R is not None:
        R = R.to(dev).contiguous()
        cnn_saved = None
        if save and _passes(model)[0] == 3:      # training in split mode: the forward keeps what the backward would recompute
            cnn_saved = ops.roi_cnn_saved(B, T, dev)
        emb = ops.roi_cnn_fwd(R, lengths, model.roi_cnn.tensors(), standardize=model.standardize_roi,
                              passes=cnn_passes(model, save), saved=cnn_saved)
        ctx["R"] = R
        ctx["cnn_saved"] = cnn_saved
        ctx["emb"] = emb
    gi = torch.empty(B, T, 2 * G3, device=dev, dtype=torch.float32)
    gi2d = gi.view(B * T, 2 * G3)
    p0, p1 = model.gru.layer(0), model.gru.layer(1)
    passes = _passes(model)[1]
    order = None
    if passes and B > 128:      # homogeneous 128-clip tiles: each cluster runs max(len in tile) steps
        order = torch.argsort(lengths, descending=True).to(torch.int32)

    def layer(pl, A2d, extra, gates, idx):
        if not passes:
            _project(A2d, pl["weight_ih"], pl["bias_ih"], gi2d, lengths, T, extra=extra)
            return ops.gru_layer_fwd(gi, lengths, pl["weight_hh"][0], pl["bias_hh"][0], pl["weight_hh"][1],
                                     pl["bias_hh"][1], gates=gates)
        # single-pass inference: the GEMM writes gi in the GRU-blocked column order (a CTA's 144 values of a row contiguous)
        blk = passes == 1 and gates is None
        fold = ops.gru_fold_bias(pl["bias_ih"][0], pl["bias_hh"][0], pl["bias_ih"][1], pl["bias_hh"][1], gru_blocked=blk)
        if passes == 1:                  # single-pass fp16: one persistent GEMM for both directions, packed weights
            packed = ops.proj_prepack(pl["weight_ih"][0], pl["weight_ih"][1], A2d.shape[1], 0 if extra is None else extra.shape[1],
                                      gru_blocked=blk)
            g = gi
            if blk and getattr(model, "gi_half", False):
                # opt-in (model.gi_half = True): gi stored as fp16 -- projection GEMM -13 %, recurrence -2 %, logit error
                # 2.6e-4 -> 4.4e-4 of max|logit| (tests/probes/gi16_probe.py); off by default, the step gains only 2 %
                g = torch.empty(B, T, 2 * G3, device=dev, dtype=torch.float16)
            ops.proj_gemm(A2d, packed, fold, g.view(B * T, 2 * G3), 2 * G3, A2=extra, lengths=lengths, T=T)
            return ops.gru_layer_fwd_tc(g, lengths, pl["weight_hh"][0], pl["bias_hh"][0], pl["weight_hh"][1],
                                        pl["bias_hh"][1], passes=passes, order=order, gates=gates, gi_blocked=blk)
        # split mode (fp32-parity grade): the same persistent GEMM with hi/lo operands, both directions in one launch
        packed = ops.proj_prepack(pl["weight_ih"][0], pl["weight_ih"][1], A2d.shape[1], 0 if extra is None else extra.shape[1],
                                  split=True)
        # training (gates saved) always splits the activations; inference follows model.proj_full_split (see above)
        full = getattr(model, "proj_full_split", None)       # None: default policy; True / False / tuple of layer indices
        if full is None:
            full = PROJ_FULL_SPLIT_DEFAULT
        full = (idx in full) if isinstance(full, (tuple, list, set)) else bool(full)
        ops.proj_gemm(A2d, packed, fold, gi2d, 2 * G3, A2=extra, lengths=lengths, T=T, split=True,
                      act_split=gates is not None or full)
        return ops.gru_layer_fwd_tc(gi, lengths, pl["weight_hh"][0], pl["bias_hh"][0], pl["weight_hh"][1],
                                    pl["bias_hh"][1], passes=passes, order=order, gates=gates)

    gates0 = torch.empty(B, T, 2, 4, H, device=dev, dtype=torch.float32) if save else None
    seq0 = layer(p0, X.view(B * T, x_dim), None if emb is None else emb.view(B * T, -1), gates0, 0)
    seq0_in = seq0
    drop_seed = 0
    if training and model.gru_dropout > 0:
        drop_seed = model._next_seed()
        seq0_in = ops.dropout(seq0, model.gru_dropout, drop_seed)
    gates1 = torch.empty(B, T, 2, 4, H, device=dev, dtype=torch.float32) if save else None
    seq1 = layer(p1, seq0_in.view(B * T, 2 * H), None, gates1, 1)
    saved = None
    if save:
        saved = [torch.empty(B, T, device=dev), torch.empty(B, 2 * H, device=dev), torch.empty(B, 2 * H, device=dev),
                 torch.empty(B, device=dev), torch.empty(B, 2 * H, device=dev), torch.empty(B, 128, device=dev)]
    head_p = model.head_dropout if training else 0.0
    head_seed = model._next_seed() if head_p > 0 else 0
    out = ops.head_fwd(seq1, lengths, model.head_tensors(), model.num_classes, saved=saved,
                       want_probs=k_top > 0, k_top=k_top, y=y, label_smoothing=label_smoothing,
                       dropout_p=head_p, seed=head_seed)
    ctx.update(seq0=seq0, seq0_in=seq0_in, seq1=seq1, gates0=gates0, gates1=gates1, head_saved=saved,
               drop_seed=drop_seed, gru_p=model.gru_dropout if training else 0.0, head_p=head_p,
               logits=out["logits"])
    return out, ctx


def classifier_forward(model, X, lengths, R=None, k_top: int = 0):
    if not X.is_cuda:
        raise ops._lib.SsbError("inputs must be CUDA tensors: this package has no CPU path")
    with torch.cuda.device(X.device):          # kernels and streams follow the tensors, not the caller's current device
        return _classifier_forward(model, X, lengths, R, k_top)


def _classifier_forward(model, X, lengths, R, k_top):
    needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in model.parameters())
    if X.shape[0] == 0 and not k_top:          # empty batch: nothing to launch (the reference returns a [0, C] tensor too)
        z = torch.zeros(0, model.num_classes, device=X.device, dtype=torch.float32)
        return z + sum(p.sum() * 0 for p in model.parameters()) if needs_grad else z
    if k_top or not needs_grad:
        with torch.no_grad():
            out, _ = forward_impl(model, X, lengths, R, training=model.training and needs_grad, save=False, k_top=k_top)
        if k_top:
            return out["topk_ids"], out["topk_probs"], out["logits"]
        return out["logits"]
    from .autograd import ClassifierFn
    names = [n for n, _ in model.named_parameters()]
    params = [p for _, p in model.named_parameters()]
    return ClassifierFn.apply(model, names, X, lengths, R, *params)
