"""Oracle (TEST INFRASTRUCTURE ONLY): per-frame landmark features on the CPU.

Restates, in numpy and with tensor inputs instead of MediaPipe landmark
objects, what the reference computes per camera frame:

* the landmark subset                live_infer_official.py:48-57  (== record_landmarks_official.py:30-44)
* ``mouth_width_px``                 live_infer_official.py:141-144 (float32 flavour)
                                     record_landmarks_official.py:52-59 (float64 flavour)
* ``extract_feature``                live_infer_official.py:146-169 (== record_landmarks_official.py:62-100)
* the capture loop's gating / velocity-reset policy
                                     live_infer_official.py:275-296, record_landmarks_official.py:183-201

Pinned against the reference's own functions by ``oracle/make_golden.py``
(fixtures ``tests/golden/features_*.npz``).
"""
from __future__ import annotations

import numpy as np

# The four anatomical groups whose union is the model's landmark set.  The
# reference spells these out as literals (live_infer_official.py:48-56); the
# union has no duplicates, so K = 21 + 19 + 11 + 37 = 88 and x_dim = 2K + 4.
_LOWER_LIP = (61, 146, 91, 181, 84, 17, 314, 405, 321, 375, 291, 308, 324, 318, 402, 317, 14, 87, 178, 88, 95)
_UPPER_LIP = (185, 40, 39, 37, 0, 267, 269, 270, 409, 415, 310, 311, 312, 13, 82, 81, 42, 183, 78)
_CHIN_ARC = (152, 377, 400, 378, 379, 394, 148, 176, 149, 150, 169)
_CHEEKS = (214, 212, 57, 186, 202, 210, 204, 211, 194, 32, 83, 201, 208, 18, 200, 199, 313, 421, 428,
           396, 406, 418, 262, 335, 424, 431, 273, 422, 430, 287, 432, 434, 364, 410, 322, 436, 416)

LANDMARK_IDS = np.array(sorted(set(_LOWER_LIP + _UPPER_LIP + _CHIN_ARC + _CHEEKS)), dtype=np.int32)
ID_LEFT_CORNER, ID_RIGHT_CORNER, ID_UPPER_INNER, ID_LOWER_INNER = 61, 291, 13, 14
GATE_MIN_PX, GATE_MAX_PX = 60.0, 150.0      # live_infer_official.py:40-41


def _px(lm_frame: np.ndarray, idx: int, w: int, h: int, dtype) -> np.ndarray:
    """Pixel position of one landmark.  The reference multiplies a Python float
    (the float32 landmark widened to double) by an int and then stores into an
    array of ``dtype`` -- for float32 that is one correctly-rounded multiply."""
    x = float(lm_frame[idx, 0]) * w
    y = float(lm_frame[idx, 1]) * h
    return np.array([x, y], dtype)


def mouth_width(lm_frame: np.ndarray, w: int, h: int, flavour: str = "live") -> float:
    """live: float32 arithmetic (live_infer_official.py:141-144);
    record: float64 arithmetic (record_landmarks_official.py:52-59)."""
    dt = np.float32 if flavour == "live" else np.float64
    a = _px(lm_frame, ID_LEFT_CORNER, w, h, dt)
    b = _px(lm_frame, ID_RIGHT_CORNER, w, h, dt)
    return float(np.linalg.norm(a - b))


def frame_feature(lm_frame: np.ndarray, w: int, h: int, ids: np.ndarray,
                  prev_xy: np.ndarray | None, flavour: str = "live"):
    """One frame -> (feat[2K+4], xy_n[K,2], center[2], mouth_w_px).

    lm_frame: float32 [N_lm, >=2] normalised landmark coordinates.
    Follows live_infer_official.py:146-169 step by step.
    """
    pts = np.array([[float(lm_frame[i, 0]) * w, float(lm_frame[i, 1]) * h] for i in ids], np.float32)
    center = pts.mean(0)
    mw = mouth_width(lm_frame, w, h, flavour)
    scale = float(mw + 1e-6)
    xy_n = (pts - center) / scale
    if prev_xy is None:
        vel = 0.0
    else:
        vel = float(np.mean(np.linalg.norm(xy_n - prev_xy, axis=1)))
    up = _px(lm_frame, ID_UPPER_INNER, w, h, np.float32)
    lo = _px(lm_frame, ID_LOWER_INNER, w, h, np.float32)
    open_px = float(np.linalg.norm(up - lo))
    aspect = float(open_px / (mw + 1e-6))
    feat = np.concatenate([xy_n.reshape(-1), np.array([vel, open_px, mw, aspect], np.float32)])
    return feat.astype(np.float32), xy_n, center, mw


def clip_features(lm: np.ndarray, w: int, h: int, ids: np.ndarray = LANDMARK_IDS,
                  present: np.ndarray | None = None, gate=(GATE_MIN_PX, GATE_MAX_PX),
                  flavour: str = "live", compact: bool = False):
    """A whole clip through the capture loop's policy.

    lm       float32 [T, N_lm, >=2]
    present  uint8  [T]   face detected in this frame (default: all ones)
    gate     (min_px, max_px) distance band on mouth width, or None to disable
             (live_infer_official.py:275-276: a frame is used only inside the band)

    A frame without a detected face is skipped and leaves the velocity history
    untouched (the reference's reset, live_infer_official.py:294-296 /
    record_landmarks_official.py:199-201, sits inside ``if res.face_landmarks:``,
    live:272 / record:178); a detected face outside the band is skipped and
    resets it.

    Returns dict with
      X       float32 [T, 2K+4]  feature rows (zeros where the frame was skipped;
                                 when ``compact`` the used frames are packed to the front,
                                 which is what the reference's ``bufX`` holds)
      valid   uint8   [T]        frame was used
      center  float32 [T, 2], mouth_w float32 [T]   (inputs of the ROI cropper)
      xy_px   float32 [T, K, 2]  the gathered pixel positions (``xy`` of live:147; rows as X)
      length  int                number of used frames
    """
    T = lm.shape[0]
    K = len(ids)
    X = np.zeros((T, 2 * K + 4), np.float32)
    valid = np.zeros(T, np.uint8)
    center = np.zeros((T, 2), np.float32)
    mouth_w = np.zeros(T, np.float32)
    xy_px = np.zeros((T, K, 2), np.float32)
    prev = None
    n = 0
    for t in range(T):
        if present is not None and not bool(present[t]):
            continue                      # `if res.face_landmarks:` is false: nothing happens, prev_xy survives
        if gate is not None:
            mw = mouth_width(lm[t], w, h, flavour)
            if not (gate[0] <= mw <= gate[1]):
                prev = None               # live:294-296
                continue
        feat, prev, c, mw = frame_feature(lm[t], w, h, ids, prev, flavour)
        row = n if compact else t
        X[row] = feat
        center[row] = c
        mouth_w[row] = mw
        xy_px[row] = np.array([[float(lm[t, i, 0]) * w, float(lm[t, i, 1]) * h] for i in ids], np.float32)   # live:147
        valid[t] = 1
        n += 1
    return {"X": X, "valid": valid, "center": center, "mouth_w": mouth_w, "xy_px": xy_px, "length": n}
