"""The reference's earlier model generation and its mouth-open detector.

``GRUClassifier``       inactive/train_reduced.py:129-145 -- nn.GRU(in, H, 1 layer, bidirectional) over ALL max_t rows
                        (no packing), time mean, Linear(2H, C).  The only trained weights shipped with the reference,
                        inactive/word_model_5.pt (in 83, H 64, 5 words, max_t 60), load into it unchanged.
``pad_or_trim``         inactive/train_reduced.py pad/trim to max_t (zero rows at the end).
``MouthOpenDetector``   important_landmarks.py:130-144 -- EMA-smoothed lip gap / inter-eye distance with hysteresis, for
                        many streams at once.

Inference only (the reference's current trainer is train_model_official.py; this class exists to check the GRU kernels
against real trained weights).  Every FLOP runs behind the C ABI: ssb_sgemm (input projection, head) and
ssb_gru_small_fwd (recurrence + time mean); there is no CPU path.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib, ops


def pad_or_trim(X: np.ndarray, max_t: int) -> np.ndarray:
    """[T,D] -> [max_t,D]: first max_t rows, zero rows appended (inactive/train_reduced.py `pad_or_trim`)."""
    T = X.shape[0]
    if T >= max_t:
        return X[:max_t]
    out = np.zeros((max_t, X.shape[1]), dtype=X.dtype)
    out[:T] = X
    return out


class _Bag(nn.Module):
    def __init__(self, **shapes):
        super().__init__()
        for name, shp in shapes.items():
            self.register_parameter(name, nn.Parameter(torch.empty(*shp)))


class GRUClassifier(nn.Module):
    """Same constructor and state-dict names as the reference class: gru.weight_ih_l0[_reverse] (3H,in),
    gru.weight_hh_l0[_reverse] (3H,H), gru.bias_ih_l0[_reverse], gru.bias_hh_l0[_reverse], head.0.weight (C,2H), head.0.bias."""

    def __init__(self, input_dim: int, num_classes: int, hidden: int = 64):
        super().__init__()
        if hidden not in (32, 64, 96, 128):
            raise ValueError("ssb_gru_small_fwd keeps W_hh resident in one SM: hidden must be 32, 64, 96 or 128")
        self.input_dim, self.num_classes, self.hidden = int(input_dim), int(num_classes), int(hidden)
        H = hidden
        self.gru = _Bag(weight_ih_l0=(3 * H, input_dim), weight_hh_l0=(3 * H, H), bias_ih_l0=(3 * H,), bias_hh_l0=(3 * H,),
                        weight_ih_l0_reverse=(3 * H, input_dim), weight_hh_l0_reverse=(3 * H, H),
                        bias_ih_l0_reverse=(3 * H,), bias_hh_l0_reverse=(3 * H,))
        with torch.no_grad():
            for p in self.gru.parameters():                     # nn.GRU.reset_parameters
                p.uniform_(-1.0 / math.sqrt(H), 1.0 / math.sqrt(H))
        self.head = nn.Module()
        fc = _Bag(weight=(num_classes, 2 * H), bias=(num_classes,))
        with torch.no_grad():                                   # nn.Linear default init
            bound = 1.0 / math.sqrt(2 * H)
            fc.weight.uniform_(-bound, bound); fc.bias.uniform_(-bound, bound)
        self.head.add_module("0", fc)

    @torch.no_grad()
    def forward(self, x: torch.Tensor, lengths: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x f32 [B,T,in] -> logits f32 [B,C].  lengths=None is the reference's call (every row, padding included);
        passing lengths switches the recurrence to packed-sequence semantics (not used by the reference class)."""
        if not x.is_cuda:
            raise _lib.SsbError("GRUClassifier needs CUDA tensors (no CPU path)")
        B, T, D = x.shape
        H, g = self.hidden, self.gru
        x2 = x.contiguous().float().view(B * T, D)
        gi = torch.empty(B, T, 6 * H, device=x.device, dtype=torch.float32)
        gi2 = gi.view(B * T, 6 * H)
        ops.linear_nt(x2, g.weight_ih_l0, g.bias_ih_l0, gi2, c_col0=0)
        ops.linear_nt(x2, g.weight_ih_l0_reverse, g.bias_ih_l0_reverse, gi2, c_col0=3 * H)
        pooled, _ = ops.gru_small_fwd(gi, g.weight_hh_l0, g.bias_hh_l0, g.weight_hh_l0_reverse, g.bias_hh_l0_reverse,
                                      lengths=lengths, hidden=H)
        fc = getattr(self.head, "0")
        logits = torch.empty(B, self.num_classes, device=x.device, dtype=torch.float32)
        ops.linear_nt(pooled, fc.weight, fc.bias, logits)
        return logits


def load_word_model(path: str, device="cuda"):
    """Reads a checkpoint written by inactive/train_reduced.py (keys model, id_to_label, label_to_id, input_dim, max_t,
    words) the way inactive/live_test_5.py:162 does.  -> (model, id_to_label, max_t)."""
    ck = torch.load(path, map_location="cpu")
    sd = ck["model"]
    hidden = sd["gru.weight_hh_l0"].shape[1]
    model = GRUClassifier(int(ck["input_dim"]), sd["head.0.weight"].shape[0], hidden=hidden)
    model.load_state_dict(sd)
    return model.eval().to(device), ck["id_to_label"], int(ck["max_t"])


class MouthOpenDetector:
    """important_landmarks.py:118-144 for S independent streams: state (ema, open) lives on the device."""

    EMA_ALPHA, OPEN_THR, CLOSE_THR = 0.25, 0.02, 0.02          # important_landmarks.py:58-62

    def __init__(self, num_streams: int, device="cuda", alpha: float = EMA_ALPHA, open_thr: float = OPEN_THR,
                 close_thr: float = CLOSE_THR):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.SsbError("MouthOpenDetector needs a CUDA device (no CPU path)")
        self.alpha, self.open_thr, self.close_thr = float(alpha), float(open_thr), float(close_thr)
        self.ema = torch.zeros(num_streams, device=dev, dtype=torch.float64)           # mouth_ema = 0.0 (:109)
        self.open = torch.zeros(num_streams, device=dev, dtype=torch.uint8)            # mouth_state_open = False (:108)

    def update(self, landmarks: torch.Tensor, valid: Optional[torch.Tensor] = None):
        """landmarks f32 [S,F,n_lm,3] normalised coordinates, F consecutive frames per stream; valid u8 [S,F].
        -> (ema f64 [S,F], open u8 [S,F]): the values the reference draws on each frame."""
        return ops.mouth_open_ema(landmarks, valid, self.ema, self.open, self.alpha, self.open_thr, self.close_thr)
