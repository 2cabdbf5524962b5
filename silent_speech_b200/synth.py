"""Seeded synthetic inputs of the reference's shapes.

There is no dataset (``clips_npz/`` is not shipped) and no trained checkpoint,
so tests and ``bench.py`` drive the path with deterministic numpy (PCG64)
data: landmark tracks of a jittering face whose mouth width falls inside the
recorder's 60..150 px distance band, blurred-noise mouth ROIs, and clip
lengths either all = T or uniform in [6, T].

Shapes follow train_model_official.py:1-9,34,38 (T=90, ROI 48x96 u8) and
record_landmarks_official.py:30-46 (K=88 landmarks of 478).
"""
from __future__ import annotations

import math

import numpy as np

N_LANDMARKS = 478
FRAME_W, FRAME_H = 1280, 720
ROI_H, ROI_W = 48, 96
MAX_T = 90


def landmarks(B: int, T: int, seed: int = 42, n_lm: int = N_LANDMARKS, drop: float = 0.02):
    """-> (lm float32 [B,T,n_lm,3], present uint8 [B,T]).

    A fixed base face (uniform 0.30..0.70) plus per-frame jitter N(0, 0.004).
    The mouth corners (ids 61, 291) are placed so that the mouth width at
    1280x720 is 70..140 px; ``drop`` of the frames are marked absent to exercise
    the velocity reset (live_infer_official.py:294-296)."""
    rng = np.random.default_rng(seed)
    base = rng.uniform(0.30, 0.70, size=(B, 1, n_lm, 3)).astype(np.float32)
    width_px = rng.uniform(70.0, 140.0, size=(B, 1)).astype(np.float32)
    cx = rng.uniform(0.45, 0.55, size=(B, 1)).astype(np.float32)
    cy = rng.uniform(0.55, 0.65, size=(B, 1)).astype(np.float32)
    half = width_px / FRAME_W / 2
    base[:, :, 61, 0] = cx - half
    base[:, :, 291, 0] = cx + half
    base[:, :, 61, 1] = cy
    base[:, :, 291, 1] = cy
    base[:, :, 13, 0] = cx; base[:, :, 14, 0] = cx
    base[:, :, 13, 1] = cy - 0.01; base[:, :, 14, 1] = cy + 0.01
    lm = base + rng.normal(0.0, 0.004, size=(B, T, n_lm, 3)).astype(np.float32)
    present = (rng.uniform(size=(B, T)) >= drop).astype(np.uint8)
    return lm.astype(np.float32), present


def lengths(B: int, T: int, mode: str = "full", seed: int = 42) -> np.ndarray:
    """'full': all = T (L1, the headline); 'uniform': integers in [6, T] with at
    least one clip of full length (L2; the recorder keeps clips of >= 6 frames,
    record_landmarks_official.py:235)."""
    if mode == "full":
        return np.full(B, T, np.int64)
    rng = np.random.default_rng(seed + 1)
    out = rng.integers(min(6, T), T + 1, size=B).astype(np.int64)
    out[rng.integers(0, B)] = T
    return out


def point_features(B: int, T: int, lens: np.ndarray, x_dim: int = 180, seed: int = 42) -> np.ndarray:
    """Model-only runs: X ~ N(0, 0.25), zero rows past each clip's length
    (what clip_pad_trim produces, train_model_official.py:93-103)."""
    rng = np.random.default_rng(seed + 2)
    X = rng.normal(0.0, 0.25, size=(B, T, x_dim)).astype(np.float32)
    X[np.arange(T)[None, :] >= lens[:, None]] = 0.0
    return X


def rois(B: int, T: int, lens: np.ndarray, seed: int = 42) -> np.ndarray:
    """uint8 [B,T,48,96]: uniform noise with a 3x3 box blur (realistic per-frame
    std), a per-frame brightness offset, zero frames past each clip's length
    (roi_pad_trim, train_model_official.py:105-117)."""
    rng = np.random.default_rng(seed + 3)
    out = np.zeros((B, T, ROI_H, ROI_W), np.uint8)
    chunk = max(1, min(B, (1 << 26) // (T * ROI_H * ROI_W)))
    for b0 in range(0, B, chunk):
        b1 = min(B, b0 + chunk)
        n = rng.integers(0, 256, size=(b1 - b0, T, ROI_H + 2, ROI_W + 2)).astype(np.float32)
        acc = np.zeros((b1 - b0, T, ROI_H, ROI_W), np.float32)
        for dy in range(3):
            for dx in range(3):
                acc += n[:, :, dy:dy + ROI_H, dx:dx + ROI_W]
        acc = acc / 9.0 + rng.uniform(-40, 40, size=(b1 - b0, T, 1, 1)).astype(np.float32)
        out[b0:b1] = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
    out[np.arange(T)[None, :] >= lens[:, None]] = 0
    return out


def labels(B: int, num_classes: int = 10, seed: int = 42) -> np.ndarray:
    return np.random.default_rng(seed + 4).integers(0, num_classes, size=B).astype(np.int64)


def clip_batch(B: int, T: int = MAX_T, mode: str = "full", seed: int = 42, use_roi: bool = True,
               x_dim: int = 180, num_classes: int = 10):
    """-> dict(X, lengths, R or None, y) as numpy arrays in the collate_fn layout
    (train_model_official.py:174-204)."""
    L = lengths(B, T, mode, seed)
    return {"X": point_features(B, T, L, x_dim, seed), "lengths": L,
            "R": rois(B, T, L, seed) if use_roi else None, "y": labels(B, num_classes, seed)}


# --------------------------------------------------------------------------
# synthetic weights (no trained checkpoint is shipped: .MISSING_LARGE_BLOBS lists word_model_points_roi.pt)
# --------------------------------------------------------------------------
def param_shapes(x_dim: int = 180, num_classes: int = 10, use_roi: bool = True, roi_emb: int = 32, hidden: int = 192):
    """State-dict names and shapes in the order the reference's ``state_dict()`` lists them
    (TinyROICNN train_model_official.py:212-221, nn.GRU 261-267, AttnPool 236, head 271-277)."""
    s = {}
    if use_roi:
        for idx, (co, ci) in (("0", (8, 1)), ("3", (16, 8)), ("6", (24, 16))):
            s[f"roi_cnn.net.{idx}.weight"] = (co, ci, 3, 3)
            s[f"roi_cnn.net.{idx}.bias"] = (co,)
        s["roi_cnn.fc.weight"] = (roi_emb, 24)
        s["roi_cnn.fc.bias"] = (roi_emb,)
    for layer, width in ((0, x_dim + (roi_emb if use_roi else 0)), (1, 2 * hidden)):
        for sfx in ("", "_reverse"):
            s[f"gru.weight_ih_l{layer}{sfx}"] = (3 * hidden, width)
            s[f"gru.weight_hh_l{layer}{sfx}"] = (3 * hidden, hidden)
            s[f"gru.bias_ih_l{layer}{sfx}"] = (3 * hidden,)
            s[f"gru.bias_hh_l{layer}{sfx}"] = (3 * hidden,)
    s["pool.score.weight"] = (1, 2 * hidden)
    s["pool.score.bias"] = (1,)
    s["head.0.weight"] = (2 * hidden,)
    s["head.0.bias"] = (2 * hidden,)
    s["head.1.weight"] = (128, 2 * hidden)
    s["head.1.bias"] = (128,)
    s["head.4.weight"] = (num_classes, 128)
    s["head.4.bias"] = (num_classes,)
    return s


def synth_params(seed: int, x_dim: int = 180, num_classes: int = 10, use_roi: bool = True, scale: float = 1.0):
    """Machine-independent weights (numpy PCG64) with the reference constructor's default-init magnitudes:
    U(+-1/sqrt(fan_in)) for convolutions / linears, U(+-1/sqrt(H)) for the GRU, LayerNorm near (1, 0) but perturbed.
    -> {state-dict name: torch.float32 tensor}.  The same generator as the test oracle's (tests assert equality), kept
    here so that bench.py and the examples never import test infrastructure."""
    import torch
    shapes = param_shapes(x_dim, num_classes, use_roi)
    rng = np.random.default_rng(seed)
    out = {}
    for name, shp in shapes.items():
        if name == "head.0.weight":
            out[name] = torch.from_numpy(1.0 + 0.1 * rng.standard_normal(shp)).float()
            continue
        if name == "head.0.bias":
            out[name] = torch.from_numpy(0.1 * rng.standard_normal(shp)).float()
            continue
        if name.startswith("gru."):
            b = 1.0 / math.sqrt(192)
        else:
            wshape = shapes[name.rsplit(".", 1)[0] + ".weight"]
            b = 1.0 / math.sqrt(int(np.prod(wshape[1:])))
        out[name] = torch.from_numpy(rng.uniform(-b * scale, b * scale, size=shp)).float()
    return out
