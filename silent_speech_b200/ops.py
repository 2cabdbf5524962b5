"""Thin tensor-level wrappers over the C ABI (one Python function per extern "C" entry).

PyTorch is used for device memory and streams only: every function below checks
shapes/dtypes, hands raw device pointers to libssb200.so and returns the output
tensors it allocated.  All compute happens in the CUDA library; nothing here has an
eager fallback (calls raise if the library is missing or the tensors are not CUDA)."""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import check, f32, i32, i64, ptr, ptr_array, stream_ptr, u64

HIDDEN = 192
ROI_H, ROI_W, ROI_EMB = 48, 96, 32


def _need(t: torch.Tensor, dtype, name: str):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise _lib.SsbError(f"{name} must be a CUDA tensor (this package has no CPU path)")
    if t.dtype != dtype:
        raise _lib.SsbError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise _lib.SsbError(f"{name} must be contiguous")
    return t


def device_info():
    sm, major, minor = i32(), i32(), i32()
    import ctypes
    check(_lib.lib().ssb_device_info(ctypes.byref(sm), ctypes.byref(major), ctypes.byref(minor)))
    _lib._launches -= 1          # a probe, not a kernel launch
    return sm.value, major.value, minor.value


# ----------------------------------------------------------------------------- K1
def features_fwd(landmarks: torch.Tensor, ids: torch.Tensor, anchor_ids: Sequence[int],
                 frame_w: float, frame_h: float, present: Optional[torch.Tensor] = None,
                 gate=(60.0, 150.0), compact: bool = False, want_xy_px: bool = False):
    """landmarks f32 [B,T,N,S>=2] -> dict(X [B,T,2K+4], center, mouth_w, valid, lengths[, xy_px [B,T,K,2]])."""
    _need(landmarks, torch.float32, "landmarks")
    _need(ids, torch.int32, "ids")
    B, T, N, S = landmarks.shape
    K = ids.numel()
    if present is not None:
        _need(present, torch.uint8, "present")
    dev = landmarks.device
    X = torch.empty(B, T, 2 * K + 4, device=dev, dtype=torch.float32)
    center = torch.empty(B, T, 2, device=dev, dtype=torch.float32)
    mouth_w = torch.empty(B, T, device=dev, dtype=torch.float32)
    valid = torch.empty(B, T, device=dev, dtype=torch.uint8)
    lengths = torch.full((B,), T, device=dev, dtype=torch.int64)
    xy_px = torch.empty(B, T, K, 2, device=dev, dtype=torch.float32) if want_xy_px else None
    lo, hi = (-float("inf"), float("inf")) if gate is None else gate
    anchors = (i32 * 4)(*[int(a) for a in anchor_ids])
    check(_lib.lib().ssb_features_fwd(ptr(landmarks), i64(B), i64(T), i64(N), i64(S), ptr(ids), i32(K), anchors,
                                      ptr(present), f32(frame_w), f32(frame_h), f32(lo), f32(hi),
                                      i32(1 if compact else 0), ptr(X), ptr(center), ptr(mouth_w), ptr(valid),
                                      ptr(lengths), ptr(xy_px), stream_ptr()))
    if compact:
        _lib._launches += 1          # compact mode enqueues two kernels (the usable-flag pass, then the features)
    else:
        lengths = valid.sum(dim=1, dtype=torch.int64)
    out = {"X": X, "center": center, "mouth_w": mouth_w, "valid": valid, "lengths": lengths}
    if want_xy_px:
        out["xy_px"] = xy_px
    return out


# ----------------------------------------------------------------------------- K2
def roi_cnn_fwd(R: torch.Tensor, lengths: Optional[torch.Tensor], weights: Sequence[torch.Tensor],
                standardize: bool = True, out: Optional[torch.Tensor] = None, out_col: int = 0,
                passes: int = 0, saved: Optional[torch.Tensor] = None) -> torch.Tensor:
    """R u8 [B,T,48,96] -> embedding f32 [B,T,32] (or written into out[..., out_col:out_col+32]).
    passes: 0 fp32 CUDA cores; 1 single fp16 pass; 2 split weights x fp16 activations; 3 both operands split (hi + lo).
    saved (passes = 3): uint8 [B*T, roi_cnn_saved_bytes()] receiving the activations roi_cnn_bwd needs."""
    _need(R, torch.uint8, "R")
    B, T, Hh, Ww = R.shape
    for w in weights:
        _need(w, torch.float32, "roi weight")
    if lengths is not None:
        _need(lengths, torch.int64, "lengths")
    if out is None:
        out = torch.empty(B, T, ROI_EMB, device=R.device, dtype=torch.float32)
        out_col = 0
    _need(out, torch.float32, "out")
    stride = out.shape[-1]
    base = out.data_ptr() + 4 * out_col
    import ctypes
    if passes:
        check(_lib.lib().ssb_roi_cnn_fwd_tc3(ptr(R), ptr(lengths), i64(B), i64(T), i32(Hh), i32(Ww), ptr_array(weights),
                                            i32(1 if standardize else 0), ctypes.c_void_p(base), i64(stride),
                                            i32(passes), ptr(saved), stream_ptr()))
        return out
    check(_lib.lib().ssb_roi_cnn_fwd(ptr(R), ptr(lengths), i64(B), i64(T), i32(Hh), i32(Ww), ptr_array(weights),
                                     i32(1 if standardize else 0), ctypes.c_void_p(base), i64(stride), stream_ptr()))
    return out


# ----------------------------------------------------------------------------- GEMM
def sgemm(A: torch.Tensor, Bm: torch.Tensor, C: torch.Tensor, M: int, N: int, K: int,
          sa=(None, None), sb=(None, None), ldc: Optional[int] = None, bias: Optional[torch.Tensor] = None,
          accumulate: bool = False, split_k: int = 1, lengths: Optional[torch.Tensor] = None, T: int = 0,
          a_off: int = 0, b_off: int = 0, c_off: int = 0):
    """C[i,j] (+)= sum_k A[a_off + i*sai + k*sak] * B[b_off + k*sbk + j*sbj] (+ bias[j]); element offsets/strides."""
    import ctypes
    check(_lib.lib().ssb_sgemm(ctypes.c_void_p(A.data_ptr() + 4 * a_off), i64(sa[0]), i64(sa[1]),
                               ctypes.c_void_p(Bm.data_ptr() + 4 * b_off), i64(sb[0]), i64(sb[1]), ptr(bias),
                               ctypes.c_void_p(C.data_ptr() + 4 * c_off), i64(ldc), i64(M), i64(N), i64(K),
                               i32(1 if accumulate else 0), i32(split_k), ptr(lengths), i64(T), stream_ptr()))
    return C


def gemm_tcx(A: torch.Tensor, Bm: torch.Tensor, C: torch.Tensor, M: int, N: int, K: int,
             sa=(None, None), sb=(None, None), ldc: Optional[int] = None,
             accumulate: bool = False, split_k: int = 1, lengths: Optional[torch.Tensor] = None, T: int = 0,
             a_off: int = 0, b_off: int = 0, c_off: int = 0, a_amax: Optional[torch.Tensor] = None):
    """sgemm's contract on the tensor cores (split-fp16 x3): the backward GEMMs of the GRU.  a_amax = absmax(A)."""
    import ctypes
    check(_lib.lib().ssb_gemm_tcx(ctypes.c_void_p(A.data_ptr() + 4 * a_off), i64(sa[0]), i64(sa[1]),
                                  ctypes.c_void_p(Bm.data_ptr() + 4 * b_off), i64(sb[0]), i64(sb[1]),
                                  ctypes.c_void_p(C.data_ptr() + 4 * c_off), i64(ldc), i64(M), i64(N), i64(K),
                                  i32(1 if accumulate else 0), i32(split_k), ptr(lengths), i64(T), ptr(a_amax),
                                  stream_ptr()))
    return C


def absmax(x: torch.Tensor) -> torch.Tensor:
    """Device scalar max|x| (no host sync); the scale source of gemm_tcx."""
    _need(x, torch.float32, "x")
    out = torch.empty(1, device=x.device, dtype=torch.float32)
    check(_lib.lib().ssb_absmax(ptr(x), i64(x.numel()), ptr(out), stream_ptr()))
    return out


def linear_nt(A2d: torch.Tensor, W: torch.Tensor, bias, C2d: torch.Tensor, accumulate=False,
              lengths=None, T=0, w_col0: int = 0, k: Optional[int] = None, c_col0: int = 0):
    """C[:, c_col0:c_col0+N] (+)= A @ W[:, w_col0:w_col0+k]^T (+ bias).  A [M,k] rows contiguous."""
    M, Ka = A2d.shape
    N = W.shape[0]
    k = Ka if k is None else k
    return sgemm(A2d, W, C2d, M, N, k, sa=(A2d.stride(0), 1), sb=(1, W.stride(0)), ldc=C2d.stride(0), bias=bias,
                 accumulate=accumulate, lengths=lengths, T=T, b_off=w_col0, c_off=c_col0)


# ----------------------------------------------------------------------------- K3b
def gru_layer_fwd(gi: torch.Tensor, lengths: torch.Tensor, whh_f, bhh_f, whh_r, bhh_r,
                  out: Optional[torch.Tensor] = None, gates: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need(gi, torch.float32, "gi")
    _need(lengths, torch.int64, "lengths")
    B, T, ld = gi.shape
    if out is None:
        out = torch.empty(B, T, 2 * HIDDEN, device=gi.device, dtype=torch.float32)
    check(_lib.lib().ssb_gru_layer_fwd(ptr(gi), i64(ld), ptr(lengths), i64(B), i64(T), i32(HIDDEN),
                                       ptr(whh_f), ptr(bhh_f), ptr(whh_r), ptr(bhh_r), ptr(out), ptr(gates),
                                       stream_ptr()))
    return out


# ----------------------------------------------------------------------------- K4
def head_fwd(seq: torch.Tensor, lengths: torch.Tensor, weights: Sequence[torch.Tensor], num_classes: int,
             saved: Optional[Sequence[torch.Tensor]] = None, want_probs: bool = False, k_top: int = 0,
             y: Optional[torch.Tensor] = None, label_smoothing: float = 0.0, dropout_p: float = 0.0,
             seed: int = 0):
    _need(seq, torch.float32, "seq")
    _need(lengths, torch.int64, "lengths")
    B, T, D = seq.shape
    dev = seq.device
    logits = torch.empty(B, num_classes, device=dev, dtype=torch.float32)
    probs = torch.empty(B, num_classes, device=dev, dtype=torch.float32) if want_probs else None
    ids = torch.empty(B, k_top, device=dev, dtype=torch.int32) if k_top else None
    tp = torch.empty(B, k_top, device=dev, dtype=torch.float32) if k_top else None
    loss = torch.empty(B, device=dev, dtype=torch.float32) if y is not None else None
    check(_lib.lib().ssb_head_fwd(ptr(seq), ptr(lengths), i64(B), i64(T), i32(D), ptr_array(weights),
                                  i32(num_classes), ptr_array(saved) if saved is not None else None,
                                  ptr(logits), ptr(probs), ptr(ids), ptr(tp), i32(k_top), ptr(y),
                                  f32(label_smoothing), ptr(loss), f32(dropout_p), u64(seed), stream_ptr()))
    return {"logits": logits, "probs": probs, "topk_ids": ids, "topk_probs": tp, "loss": loss}


def head_bwd_logits(logits, dlogits_in, y, label_smoothing, grad_scale, w2, hid, dropout_p):
    B, C = (logits if logits is not None else dlogits_in).shape
    dev = hid.device
    dlogits = torch.empty(B, C, device=dev, dtype=torch.float32)
    dhid = torch.empty(B, 128, device=dev, dtype=torch.float32)
    check(_lib.lib().ssb_head_bwd_logits(ptr(logits), ptr(dlogits_in), ptr(y), f32(label_smoothing), f32(grad_scale),
                                         ptr(w2), ptr(hid), f32(dropout_p), i64(B), i32(C), ptr(dlogits), ptr(dhid),
                                         stream_ptr()))
    return dlogits, dhid


def head_bwd_pool(dln, xhat, rstd, attn, seq, lengths, ln_g, ws, d_ln_g, d_ln_b, d_ws, d_bs, dseq=None):
    B, T, D = seq.shape
    if dseq is None:
        dseq = torch.empty_like(seq)
    check(_lib.lib().ssb_head_bwd_pool(ptr(dln), ptr(xhat), ptr(rstd), ptr(attn), ptr(seq), ptr(lengths), i64(B), i64(T),
                                       ptr(ln_g), ptr(ws), ptr(dseq), ptr(d_ln_g), ptr(d_ln_b), ptr(d_ws), ptr(d_bs),
                                       stream_ptr()))
    return dseq


def colsum(A2d: torch.Tensor, out: torch.Tensor):
    M, N = A2d.shape
    check(_lib.lib().ssb_colsum(ptr(A2d), i64(A2d.stride(0)), i64(M), i64(N), ptr(out), stream_ptr()))
    return out


# ----------------------------------------------------------------------------- backward / optimiser
def roi_cnn_saved(B: int, T: int, device) -> torch.Tensor:
    """Buffer for the activations the split-mode forward hands to roi_cnn_bwd (60,000 B per frame)."""
    L = _lib.lib()
    L.ssb_roi_cnn_saved_bytes.restype = i64
    return torch.empty(B * T, int(L.ssb_roi_cnn_saved_bytes()), device=device, dtype=torch.uint8)


def roi_cnn_bwd(R, lengths, weights, standardize, d_out: torch.Tensor, d_col: int, grads, saved: Optional[torch.Tensor] = None):
    """Accumulates the 8 ROI-CNN parameter gradients from d_out[..., d_col:d_col+32].  saved = the buffer the forward
    filled (roi_cnn_fwd(..., passes=3, saved=...)): skips the recompute."""
    import ctypes
    B, T, Hh, Ww = R.shape
    base = d_out.data_ptr() + 4 * d_col
    check(_lib.lib().ssb_roi_cnn_bwd(ptr(R), ptr(lengths), i64(B), i64(T), i32(Hh), i32(Ww), ptr_array(weights),
                                     i32(1 if standardize else 0), ctypes.c_void_p(base), i64(d_out.shape[-1]),
                                     ptr_array(grads), ptr(saved), stream_ptr()))


def transpose(x2d: torch.Tensor) -> torch.Tensor:
    r, c = x2d.shape
    out = torch.empty(c, r, device=x2d.device, dtype=torch.float32)
    check(_lib.lib().ssb_transpose(ptr(x2d), ptr(out), i64(r), i64(c), stream_ptr()))
    return out


def gru_layer_bwd(dout, out, gates, whhT_f, whhT_r, lengths):
    B, T, _ = out.shape
    dev = out.device
    dgi = torch.empty(B, T, 2 * 3 * HIDDEN, device=dev, dtype=torch.float32)
    dghn = torch.empty(B, T, 2, HIDDEN, device=dev, dtype=torch.float32)
    hprev = torch.empty(B, T, 2, HIDDEN, device=dev, dtype=torch.float32)
    check(_lib.lib().ssb_gru_layer_bwd(ptr(dout), ptr(out), ptr(gates), ptr(whhT_f), ptr(whhT_r), ptr(lengths),
                                       i64(B), i64(T), i32(HIDDEN), ptr(dgi), ptr(dghn), ptr(hprev), stream_ptr()))
    return dgi, dghn, hprev


def dropout(x: torch.Tensor, p: float, seed: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need(x, torch.float32, "x")
    y = torch.empty_like(x) if out is None else out
    check(_lib.lib().ssb_dropout(ptr(x), ptr(y), i64(x.numel()), f32(p), u64(seed), stream_ptr()))
    return y


def grad_sqnorm(g: torch.Tensor, sqnorm: torch.Tensor):
    check(_lib.lib().ssb_grad_sqnorm(ptr(g), i64(g.numel()), ptr(sqnorm), stream_ptr()))


def clip_adam_step(p, g, m, v, sqnorm, max_norm, lr, beta1, beta2, eps, step, write_back_grad=False, norm_out=None):
    check(_lib.lib().ssb_clip_adam_step(ptr(p), ptr(g), ptr(m), ptr(v), i64(p.numel()), ptr(sqnorm), f32(max_norm),
                                        f32(lr), f32(beta1), f32(beta2), f32(eps), i64(step),
                                        i32(1 if write_back_grad else 0), ptr(norm_out), stream_ptr()))


# ----------------------------------------------------------------------------- tensor-core path
def gemm_tc(A1: torch.Tensor, W: torch.Tensor, bias, C2d: torch.Tensor, A2: Optional[torch.Tensor] = None,
            passes: int = 3, accumulate: bool = False, lengths=None, T: int = 0, c_col0: int = 0):
    """C[:, c_col0:c_col0+N] = [A1 | A2] @ W^T + bias on tcgen05 (bf16 x passes, fp32 accumulate)."""
    import ctypes
    M, K1 = A1.shape
    N = W.shape[0]
    K2 = 0 if A2 is None else A2.shape[1]
    check(_lib.lib().ssb_gemm_tc(ptr(A1), i64(A1.stride(0)), i64(K1), ptr(A2), i64(0 if A2 is None else A2.stride(0)),
                                 i64(K2), ptr(W), i64(W.stride(0)), ptr(bias),
                                 ctypes.c_void_p(C2d.data_ptr() + 4 * c_col0), i64(C2d.stride(0)), i64(M), i64(N),
                                 i32(passes), i32(1 if accumulate else 0), ptr(lengths), i64(T), stream_ptr()))
    return C2d


def proj_prepack(W_a: torch.Tensor, W_b: Optional[torch.Tensor], K1: int, K2: int = 0, gru_blocked: bool = False,
                 split: bool = False) -> torch.Tensor:
    """[W_a ; W_b] (each [n_each, K1+K2] fp32) -> fp16 block layout streamed by proj_gemm (a derived cache).
    split: hi | lo blocks for proj_gemm(..., split=True) (fp32-parity grade)."""
    _need(W_a, torch.float32, "W_a")
    if W_b is not None:
        _need(W_b, torch.float32, "W_b")
    n_each = W_a.shape[0]
    N = n_each * (2 if W_b is not None else 1)
    L = _lib.lib()
    L.ssb_proj_packed_bytes.restype = i64
    L.ssb_proj_packed_bytes_split.restype = i64
    nbytes = (L.ssb_proj_packed_bytes_split if split else L.ssb_proj_packed_bytes)(i64(N), i64(K1), i64(K2))
    packed = torch.empty(nbytes, device=W_a.device, dtype=torch.uint8)
    fn = L.ssb_proj_prepack_split if split else L.ssb_proj_prepack
    check(fn(ptr(W_a), ptr(W_b), i64(W_a.stride(0)), i64(n_each), i64(K1), i64(K2),
             i32(1 if gru_blocked else 0), ptr(packed), stream_ptr()))
    return packed


def proj_gemm(A1: torch.Tensor, packed: torch.Tensor, bias, C2d: torch.Tensor, N: int, A2: Optional[torch.Tensor] = None,
              lengths=None, T: int = 0, split: bool = False, act_split: bool = True):
    # C2d may be float32 or float16 (fp16 gi for the single-pass recurrence)
    """C[:, :N] = [A1 | A2] @ packed^T + bias: both directions' input projections in one persistent tcgen05 GEMM.
    split: fp32-parity grade (hi/lo split of both operands; packed from proj_prepack(..., split=True); C fp32);
    act_split=False: the inference form (weights split, activations rounded to fp16 once, 128-row tiles)."""
    M, K1 = A1.shape
    K2 = 0 if A2 is None else A2.shape[1]
    if split:
        if not (C2d.is_cuda and C2d.dtype == torch.float32 and C2d.stride(1) == 1):
            raise _lib.SsbError("split projection: C must be a CUDA float32 matrix with unit column stride")
        fn = _lib.lib().ssb_proj_gemm_split if act_split else _lib.lib().ssb_proj_gemm_wsplit
        check(fn(ptr(A1), i64(A1.stride(0)), i64(K1), ptr(A2), i64(0 if A2 is None else A2.stride(0)),
                                             i64(K2), ptr(packed), ptr(bias), ptr(C2d), i64(C2d.stride(0)), i64(M), i64(N),
                                             ptr(lengths), i64(T), stream_ptr()))
        return C2d
    check(_lib.lib().ssb_proj_gemm(ptr(A1), i64(A1.stride(0)), i64(K1), ptr(A2), i64(0 if A2 is None else A2.stride(0)),
                                   i64(K2), ptr(packed), ptr(bias), ptr(C2d), i64(C2d.stride(0)), i64(M), i64(N),
                                   ptr(lengths), i64(T), i32(1 if C2d.dtype == torch.float16 else 0), stream_ptr()))
    return C2d


def gru_fold_bias(bih_f, bhh_f, bih_r, bhh_r, gru_blocked: bool = False) -> torch.Tensor:
    out = torch.empty(2 * 3 * HIDDEN, device=bih_f.device, dtype=torch.float32)
    check(_lib.lib().ssb_gru_fold_bias(ptr(bih_f), ptr(bhh_f), ptr(bih_r), ptr(bhh_r), ptr(out), i32(1 if gru_blocked else 0), stream_ptr()))
    return out


def gru_layer_fwd_tc(gi, lengths, whh_f, bhh_f, whh_r, bhh_r, passes=3, order=None, out=None, gates=None, gi_blocked=False):
    """Tensor-core recurrence; gi must include the folded bias (gru_fold_bias).  A float16 gi (GRU-blocked columns,
    written by proj_gemm into a half tensor) selects the double-buffered fp16 staging (gi_layout 2)."""
    half = gi.dtype == torch.float16
    if half and not gi_blocked:
        raise _lib.SsbError("fp16 gi exists in the GRU-blocked column order only")
    _need(gi, torch.float16 if half else torch.float32, "gi")
    _need(lengths, torch.int64, "lengths")
    B, T, ld = gi.shape
    if out is None:
        out = torch.empty(B, T, 2 * HIDDEN, device=gi.device, dtype=torch.float32)
    if order is not None:
        _need(order, torch.int32, "order")
    check(_lib.lib().ssb_gru_layer_fwd_tc(ptr(gi), i64(ld), ptr(lengths), ptr(order), i64(B), i64(T), i32(HIDDEN),
                                          ptr(whh_f), ptr(bhh_f), ptr(whh_r), ptr(bhh_r), ptr(out), ptr(gates),
                                          i32(passes), i32(2 if half else (1 if gi_blocked else 0)), stream_ptr()))
    return out


# ----------------------------------------------------------------------------- K6
def roi_crop(frames: torch.Tensor, centers: torch.Tensor, mouth_w: torch.Tensor, flavor: str = "live",
             frame_index: Optional[torch.Tensor] = None):
    """frames u8 [F,Hf,Wf,3] (BGR), centers f32 [N,2], mouth_w f64 [N] -> (roi u8 [N,48,96], ok u8 [N]).
    flavor 'live' = crop_roi_gray (INTER_AREA), 'record' = crop_roi (INTER_LINEAR); bit-exact with OpenCV."""
    _need(frames, torch.uint8, "frames")
    _need(centers, torch.float32, "centers")
    _need(mouth_w, torch.float64, "mouth_w")
    if frame_index is not None:
        _need(frame_index, torch.int32, "frame_index")
    F, Hf, Wf, C = frames.shape
    if C != 3:
        raise _lib.SsbError("frames must be [F,H,W,3] BGR")
    N = centers.shape[0]
    if frame_index is None and N != F:
        raise _lib.SsbError("without frame_index there must be one frame per crop")
    out = torch.empty(N, ROI_H, ROI_W, device=frames.device, dtype=torch.uint8)
    ok = torch.empty(N, device=frames.device, dtype=torch.uint8)
    check(_lib.lib().ssb_roi_crop(ptr(frames), ptr(frame_index), i64(Hf), i64(Wf), ptr(centers), ptr(mouth_w), i64(N),
                                  i32({"live": 0, "record": 1}[flavor]), ptr(out), ptr(ok), stream_ptr()))
    return out, ok


# ----------------------------------------------------------------------------- stream rings
def ring_push(ring: torch.Tensor, sid: torch.Tensor, pos: torch.Tensor, x: torch.Tensor, e: Optional[torch.Tensor] = None):
    _need(ring, torch.float32, "ring"); _need(sid, torch.int32, "sid"); _need(pos, torch.int32, "pos"); _need(x, torch.float32, "x")
    S, W, D = ring.shape
    check(_lib.lib().ssb_ring_push(ptr(ring), i64(W), i64(D), ptr(sid), ptr(pos), ptr(x), i64(x.shape[1]), ptr(e),
                                   i64(0 if e is None else e.shape[1]), i64(x.shape[0]), stream_ptr()))


def ring_gather(ring: torch.Tensor, sid: torch.Tensor, start: torch.Tensor, count: torch.Tensor, T: int, dx: int, de: int):
    _need(ring, torch.float32, "ring"); _need(sid, torch.int32, "sid"); _need(start, torch.int32, "start"); _need(count, torch.int32, "count")
    S, W, D = ring.shape
    m = sid.shape[0]
    Xw = torch.empty(m, T, dx, device=ring.device, dtype=torch.float32)
    Ew = torch.empty(m, T, de, device=ring.device, dtype=torch.float32) if de else None
    lengths = torch.empty(m, device=ring.device, dtype=torch.int64)
    check(_lib.lib().ssb_ring_gather(ptr(ring), i64(W), i64(D), ptr(sid), ptr(start), ptr(count), i64(m), ptr(Xw), i64(dx),
                                     ptr(Ew), i64(de), i64(T), ptr(lengths), stream_ptr()))
    return Xw, Ew, lengths


# ----------------------------------------------------------------------------- K7 (small-hidden GRU), mouth-open EMA
def gru_small_fwd(gi: torch.Tensor, whh_f, bhh_f, whh_r=None, bhh_r=None, lengths: Optional[torch.Tensor] = None,
                  hidden: int = 64, want_out: bool = False):
    """gi f32 [B,T,ndir*3H] -> (pooled f32 [B,ndir*H] = time mean, out f32 [B,T,ndir*H] or None)."""
    _need(gi, torch.float32, "gi")
    B, T, ld = gi.shape
    ndir = 1 if whh_r is None else 2
    if lengths is not None:
        _need(lengths, torch.int64, "lengths")
    pooled = torch.empty(B, ndir * hidden, device=gi.device, dtype=torch.float32)
    out = torch.empty(B, T, ndir * hidden, device=gi.device, dtype=torch.float32) if want_out else None
    check(_lib.lib().ssb_gru_small_fwd(ptr(gi), i64(ld), ptr(whh_f), ptr(bhh_f), ptr(whh_r), ptr(bhh_r), ptr(lengths),
                                       i64(B), i64(T), i64(hidden), i32(ndir), ptr(out), ptr(pooled), stream_ptr()))
    return pooled, out


def mouth_open_ema(lm: torch.Tensor, valid: Optional[torch.Tensor], ema_state: torch.Tensor, open_state: torch.Tensor,
                   alpha: float = 0.25, open_thr: float = 0.02, close_thr: float = 0.02, want_trace: bool = True):
    """lm f32 [S,F,n_lm,3]; updates ema_state f64 [S] / open_state u8 [S] in place -> (ema f64 [S,F], open u8 [S,F]) or None."""
    import ctypes
    _need(lm, torch.float32, "lm"); _need(ema_state, torch.float64, "ema_state"); _need(open_state, torch.uint8, "open_state")
    S, F, n_lm, _ = lm.shape
    if valid is not None:
        _need(valid, torch.uint8, "valid")
    ema = torch.empty(S, F, device=lm.device, dtype=torch.float64) if want_trace else None
    opn = torch.empty(S, F, device=lm.device, dtype=torch.uint8) if want_trace else None
    f64 = ctypes.c_double
    check(_lib.lib().ssb_mouth_open_ema(ptr(lm), ptr(valid), i64(S), i64(F), i64(n_lm), f64(alpha), f64(open_thr), f64(close_thr),
                                        ptr(ema_state), ptr(open_state), ptr(ema), ptr(opn), stream_ptr()))
    return (ema, opn) if want_trace else None


# ----------------------------------------------------------------------------- clip store
def gather_rows(src: torch.Tensor, index: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[i] = src[index[i]] along dim 0 (whole contiguous rows; index i64 on the device, values validated by the caller)."""
    if not src.is_cuda or not src.is_contiguous():
        raise _lib.SsbError("gather_rows: src must be a contiguous CUDA tensor")
    _need(index, torch.int64, "index")
    n = index.shape[0]
    if out is None:
        out = torch.empty((n,) + tuple(src.shape[1:]), device=src.device, dtype=src.dtype)
    row_bytes = (src[0].numel() if src.shape[0] else 1) * src.element_size()
    check(_lib.lib().ssb_gather_rows(ptr(src), i64(row_bytes), ptr(index), i64(n), ptr(out), stream_ptr()))
    return out


# ----------------------------------------------------------------------------- augmentation
def augment_clips(X: torch.Tensor, lengths: torch.Tensor, R: Optional[torch.Tensor], seed: int, p_noise: float = 0.7,
                  noise_std: float = 0.01, p_drop: float = 0.35, drop_max: int = 2):
    """NPZWordDataset's augmentation (train_model_official.py:143-152) for a whole batch on the device.
    -> (X_aug f32 [B,T,D], lengths_out i64 [B], decisions i32 [B,4]); R (u8 [B,T,48,96] or None) is cut in place."""
    _need(X, torch.float32, "X"); _need(lengths, torch.int64, "lengths")
    B, T, D = X.shape
    fb = 0
    if R is not None:
        _need(R, torch.uint8, "R")
        fb = R[0, 0].numel()
    Xo = torch.empty_like(X)
    lo = torch.empty_like(lengths)
    dec = torch.empty(B, 4, device=X.device, dtype=torch.int32)
    check(_lib.lib().ssb_augment_clips(ptr(X), ptr(lengths), i64(B), i64(T), i64(D), ptr(R), i64(fb), u64(seed & 0xFFFFFFFFFFFFFFFF),
                                       f32(p_noise), f32(noise_std), f32(p_drop), i32(drop_max), ptr(Xo), ptr(lo), ptr(dec),
                                       stream_ptr()))
    return Xo, lo, dec


# ----------------------------------------------------------------------------- no-allocation projection
def proj_gemm_static(A1: torch.Tensor, W_a: torch.Tensor, W_b: Optional[torch.Tensor], bias, C2d: torch.Tensor,
                     A2: Optional[torch.Tensor] = None, gru_blocked: bool = False, lengths=None, T: int = 0):
    """proj_prepack + proj_gemm in one C-ABI call that allocates nothing: the packed weights live in the library's
    static device array (884,736 bytes; N <= 1152, K <= 384).  Calls share that array: do not overlap them on the GPU.
    C2d (caller-owned, fp32 or fp16) receives [A1 | A2] @ [W_a ; W_b]^T + bias."""
    K1 = A1.shape[1]
    K2 = 0 if A2 is None else A2.shape[1]
    check(_lib.lib().ssb_proj_gemm_static(ptr(A1), i64(A1.stride(0)), i64(K1), ptr(A2), i64(0 if A2 is None else A2.stride(0)),
                                          i64(K2), ptr(W_a), ptr(W_b), i64(W_a.stride(0)), i64(W_a.shape[0]),
                                          i32(1 if gru_blocked else 0), ptr(bias), ptr(C2d), i64(C2d.stride(0)),
                                          i64(A1.shape[0]), ptr(lengths), i64(T),
                                          i32(1 if C2d.dtype == torch.float16 else 0), stream_ptr()))
    _lib._launches += 1          # two kernels: pack, GEMM
    return C2d
