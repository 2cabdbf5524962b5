"""The model's landmark subset and the batched feature entry point.

FIXED_IDXS is the union of four anatomical groups (lower lip, upper lip, chin arc,
cheeks) of MediaPipe FaceLandmarker ids -- live_infer_official.py:48-57 ==
record_landmarks_official.py:30-44 -- 88 ids, hence x_dim = 2*88 + 4 = 180.
(important_landmarks.py holds viewer masks only and feeds no model)
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import ops

LOWER_LIP = (61, 146, 91, 181, 84, 17, 314, 405, 321, 375, 291, 308, 324, 318, 402, 317, 14, 87, 178, 88, 95)
UPPER_LIP = (185, 40, 39, 37, 0, 267, 269, 270, 409, 415, 310, 311, 312, 13, 82, 81, 42, 183, 78)
CHIN_ARC = (152, 377, 400, 378, 379, 394, 148, 176, 149, 150, 169)
CHEEKS = (214, 212, 57, 186, 202, 210, 204, 211, 194, 32, 83, 201, 208, 18, 200, 199, 313, 421, 428, 396,
          406, 418, 262, 335, 424, 431, 273, 422, 430, 287, 432, 434, 364, 410, 322, 436, 416)
FIXED_IDXS = sorted(set(LOWER_LIP + UPPER_LIP + CHIN_ARC + CHEEKS))
LEFT_CORNER, RIGHT_CORNER, UPPER_INNER, LOWER_INNER = 61, 291, 13, 14      # live_infer_official.py:34-37
MOUTH_W_MIN_PX, MOUTH_W_MAX_PX = 60.0, 150.0                               # live_infer_official.py:40-41
ANCHORS = (LEFT_CORNER, RIGHT_CORNER, UPPER_INNER, LOWER_INNER)

_ids_cache = {}


def ids_tensor(device, idxs=None) -> torch.Tensor:
    key = (str(device), tuple(idxs) if idxs is not None else None)
    if key not in _ids_cache:
        _ids_cache[key] = torch.tensor(list(idxs) if idxs is not None else FIXED_IDXS, dtype=torch.int32, device=device)
    return _ids_cache[key]


def extract_features(landmarks: torch.Tensor, w: int, h: int, idxs=None, present: Optional[torch.Tensor] = None,
                     gate=(MOUTH_W_MIN_PX, MOUTH_W_MAX_PX), compact: bool = False, want_xy_px: bool = False):
    """Batched form of ``extract_feature(face, w, h, idxs, prev_xy)`` (live_infer_official.py:146-169)
    plus the capture loop's gate / velocity reset (live:275-296).

    landmarks  f32 CUDA [B,T,N_lm,>=2] normalised (x, y[, z]) as MediaPipe reports them
    present    u8 [B,T] face detected (None = all)
    gate       (min_px, max_px) band on the mouth width, None to disable
    compact    pack used frames to the front of each clip, as the reference's bufX does
    want_xy_px also return the gathered pixel positions xy_px [B,T,K,2] (`xy` of live:147, before any arithmetic)
    returns    dict X [B,T,2K+4], center [B,T,2], mouth_w [B,T], valid [B,T], lengths [B]

    An undetected frame (present = 0) is skipped without touching the velocity history; a detected face outside the
    band is skipped and resets it (live:272-296: the reset sits inside ``if res.face_landmarks``)."""
    with torch.cuda.device(landmarks.device):
        return ops.features_fwd(landmarks, ids_tensor(landmarks.device, idxs), ANCHORS, float(w), float(h),
                                present=present, gate=gate, compact=compact, want_xy_px=want_xy_px)


class _Pt:
    __slots__ = ("x", "y")


def extract_feature(face, w, h, idxs, prev_xy=None):
    """Per-frame shim with the reference's exact signature and return tuple
    (feat[2K+4], xy_n[K,2], center[2], mouth_w_px), computed by the CUDA kernel.
    ``face`` is any sequence of objects with .x/.y (MediaPipe landmarks).  When
    ``prev_xy`` is given the previous frame is re-derived from it only through the
    velocity term, exactly as live_infer_official.py:155-158 uses it."""
    n = len(face)
    cur = np.array([[p.x, p.y] for p in face], np.float32)
    K = len(idxs)
    dev = torch.device("cuda")
    lm = torch.from_numpy(cur).to(dev).view(1, 1, n, 2)
    out = ops.features_fwd(lm, ids_tensor(dev, idxs), ANCHORS, float(w), float(h), gate=None)
    feat = out["X"][0, 0].cpu().numpy()
    xy_n = feat[:2 * K].reshape(K, 2).copy()
    if prev_xy is not None:
        feat[2 * K] = float(np.mean(np.linalg.norm(xy_n - np.asarray(prev_xy, np.float32), axis=1)))
    return feat, xy_n, out["center"][0, 0].cpu().numpy(), float(out["mouth_w"][0, 0])


def crop_rois(frames, centers, mouth_w, flavor: str = "live", frame_index=None):
    """Batched crop_roi_gray (flavor 'live', live_infer_official.py:171-187) / crop_roi (flavor 'record',
    record_landmarks_official.py:102-118) on the GPU: frames u8 [F,H,W,3] BGR, centers f32 [N,2] and mouth_w [N]
    as returned by extract_features -> (R u8 [N,48,96], ok u8 [N]); rows where the reference returns None are zero
    frames (what live_infer_official.py:290-293 substitutes)."""
    with torch.cuda.device(frames.device):
        return ops.roi_crop(frames, centers.contiguous().float(), mouth_w.to(torch.float64).contiguous(), flavor, frame_index)
