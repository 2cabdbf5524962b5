"""CPU restatement of the training augmentation (TEST INFRASTRUCTURE ONLY).

Follows NPZWordDataset.__getitem__, train_model_official.py:143-172, with the random DECISIONS injected instead of
drawn (the reference draws them from Python's / numpy's global generators): noise on/off, the k frames to drop.
Same numpy operations as the reference for everything else (boolean keep mask, clip_pad_trim / roi_pad_trim padding,
T_use = min(T_eff, Tr, max_t) -- the ROI clip is cut, not thinned).  Pinned by tests/golden/augment.npz, produced by
running the reference's own __getitem__ with its random calls scripted (oracle/make_golden.py)."""
from __future__ import annotations

import numpy as np


def pad_trim(a: np.ndarray, t: int, max_t: int):
    """clip_pad_trim / roi_pad_trim (train_model_official.py:93-117)."""
    t = min(t, max_t)
    out = np.zeros((max_t,) + a.shape[1:], a.dtype)
    out[:t] = a[:t]
    return out, t


def augment_clip(X: np.ndarray, R, max_t: int, noise, drops):
    """X f32 [T,D] (unpadded), R u8 [Tr,H,W] or None, noise: None or f32 [T,D] added first, drops: iterable of frame
    indices in [1, T-2].  -> (X_pad [max_t,D], T_use, R_pad or None)."""
    X = X.astype(np.float32)
    T = X.shape[0]
    if noise is not None:
        X = X + noise.astype(np.float32)
    drops = list(drops)
    if drops:
        keep = np.ones(T, dtype=bool)
        keep[drops] = False
        X = X[keep]
        T = X.shape[0]
    Xp, T_eff = pad_trim(X, T, max_t)
    if R is None:
        return Xp, T_eff, None
    T_use = min(T_eff, R.shape[0], max_t)
    Xp, T_use = pad_trim(Xp[:T_use], T_use, max_t)
    Rp, _ = pad_trim(R[:T_use], T_use, max_t)
    return Xp, T_use, Rp
