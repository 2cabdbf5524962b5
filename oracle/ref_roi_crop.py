"""CPU restatement of the mouth-ROI crop (TEST INFRASTRUCTURE ONLY -- never imported by the product path).

Follows crop_roi_gray (live_infer_official.py:171-187: crop box centre +- (1.2, 1.0) x mouth width in Python
double arithmetic, skip when x2 <= x1+2 or y2 <= y1+2, BGR->gray, cv2.resize INTER_AREA) and crop_roi
(record_landmarks_official.py:102-118: the same box in float32 arithmetic under numpy 2 promotion, skip when
x2 <= x1 or y2 <= y1, cv2.resize default INTER_LINEAR).  The image arithmetic lives in OpenCV 4.13
(requirements.txt:18-19, not vendored); it is restated here from OpenCV's published algorithm -- fixed-point
BGR2GRAY (coefficients 3735/19235/9798, shift 15), resizeArea_ with its float accumulation order, the integer
"area fast" path, and the 11-bit fixed-point bilinear -- and pinned against cv2 itself (present in the build
container) by tests/test_oracle_golden.py on seeded crops plus the reference-generated tests/golden/roi_crop.npz.
"""
from __future__ import annotations

import math

import numpy as np

ROI_W, ROI_H = 96, 48


def bgr2gray(img: np.ndarray) -> np.ndarray:
    b, g, r = (img[..., i].astype(np.int64) for i in range(3))
    return ((b * 3735 + g * 19235 + r * 9798 + (1 << 14)) >> 15).astype(np.uint8)


def crop_box(flavor: str, w: int, h: int, center, mouth_w: float):
    """-> (x1, y1, x2, y2) or None.  `live`: double arithmetic; `record`: float32 (numpy-2 weak-scalar promotion)."""
    if flavor == "live":
        cx, cy = float(center[0]), float(center[1])
        hw, hh = 1.2 * mouth_w, 1.0 * mouth_w
        x1, x2 = int(max(0, cx - hw)), int(min(w, cx + hw))
        y1, y2 = int(max(0, cy - hh)), int(min(h, cy + hh))
        if x2 <= x1 + 2 or y2 <= y1 + 2:
            return None
    else:
        cx, cy = np.float32(center[0]), np.float32(center[1])
        hw, hh = np.float32(1.2 * mouth_w), np.float32(1.0 * mouth_w)
        x1, x2 = int(max(0, cx - hw)), int(min(w, cx + hw))
        y1, y2 = int(max(0, cy - hh)), int(min(h, cy + hh))
        if x2 <= x1 or y2 <= y1:
            return None
    return x1, y1, x2, y2


def _area_tab(ssize: int, dsize: int, scale: float):
    """computeResizeAreaTab: list of (di, si, alpha float32)."""
    tab = []
    for dx in range(dsize):
        fsx1 = dx * scale
        fsx2 = fsx1 + scale
        cell = min(scale, ssize - fsx1)
        sx1, sx2 = math.ceil(fsx1), math.floor(fsx2)
        sx2 = min(sx2, ssize - 1)
        sx1 = min(sx1, sx2)
        if sx1 - fsx1 > 1e-3:
            tab.append((dx, sx1 - 1, np.float32((sx1 - fsx1) / cell)))
        for sx in range(sx1, sx2):
            tab.append((dx, sx, np.float32(1.0 / cell)))
        if fsx2 - sx2 > 1e-3:
            tab.append((dx, sx2, np.float32(min(min(fsx2 - sx2, 1.0), cell) / cell)))
    return tab


def resize_area(src: np.ndarray, dw: int = ROI_W, dh: int = ROI_H) -> np.ndarray:
    """cv2.resize(src u8 [h,w], (dw,dh), INTER_AREA)."""
    sh, sw = src.shape
    scale_x, scale_y = 1.0 / (dw / sw), 1.0 / (dh / sh)
    if scale_x < 1.0 or scale_y < 1.0:            # crops clipped by the frame border can be smaller than 96x48
        return resize_linear(src, dw, dh, area_mode=True)
    ix, iy = int(math.floor(scale_x + 0.5)), int(math.floor(scale_y + 0.5))   # saturate_cast<int>
    if abs(scale_x - ix) < 2.220446049250313e-16 and abs(scale_y - iy) < 2.220446049250313e-16:
        # integer "area fast" path: block sums; 2x2 uses (s+2)>>2, everything else saturate_cast(sum * (1.f/area))
        blk = src[:dh * iy, :dw * ix].astype(np.int64).reshape(dh, iy, dw, ix).sum(axis=(1, 3))
        if ix == 2 and iy == 2:
            return ((blk + 2) >> 2).astype(np.uint8)
        sc = np.float32(1.0) / np.float32(ix * iy)
        return np.clip(np.rint(blk.astype(np.float32) * sc), 0, 255).astype(np.uint8)
    xtab, ytab = _area_tab(sw, dw, scale_x), _area_tab(sh, dh, scale_y)
    xd = np.array([t[0] for t in xtab]); xs = np.array([t[1] for t in xtab]); xa = np.array([t[2] for t in xtab], np.float32)
    out = np.zeros((dh, dw), np.uint8)
    acc = np.zeros(dw, np.float32)
    prev = ytab[0][0]
    first = True
    S = src.astype(np.float32)
    for dy, sy, beta in ytab:
        buf = np.zeros(dw, np.float32)
        prod = S[sy, xs] * xa                       # float32 products, added in table order
        for k in range(len(xd)):
            buf[xd[k]] = np.float32(buf[xd[k]] + prod[k])
        if dy != prev:
            out[prev] = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
            acc = beta * buf
            prev = dy
        elif first:
            acc = beta * buf
        else:
            acc = acc + beta * buf
        first = False
    out[prev] = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
    return out


def _linear_tab(ssize: int, dsize: int, clamp_frac: bool = True, area_mode: bool = False):
    """INTER_LINEAR index / 11-bit coefficient table of cv::resize for 8-bit images.  Along x, taps that fall
    outside the image get the fraction forced to 0 (clamp_frac); along y OpenCV keeps the fraction and only clips
    the ROW INDICES, so a border row is blended with itself through two separately truncated products."""
    inv_scale = dsize / ssize          # cv::resize: inv_scale = (double)dsize/ssize; scale = 1./inv_scale
    scale = 1.0 / inv_scale
    idx = np.zeros(dsize, np.int64)
    c0 = np.zeros(dsize, np.int64)
    c1 = np.zeros(dsize, np.int64)
    for d in range(dsize):
        if area_mode:                 # INTER_AREA when an axis up-scales: bilinear with cell-overlap fractions
            s = int(math.floor(d * scale))
            f = np.float32((d + 1) - (s + 1) * inv_scale)
            f = np.float32(0) if f <= 0 else np.float32(f - math.floor(f))
        else:
            f = np.float32((d + 0.5) * scale - 0.5)
            s = int(math.floor(f))
            f = np.float32(f - s)
        if clamp_frac:
            if s < 0:
                s, f = 0, np.float32(0)
            if s >= ssize - 1:
                s, f = ssize - 1, np.float32(0)
        idx[d] = s
        c0[d] = int(np.rint(np.float32(np.float32(1.0) - f) * np.float32(2048)))
        c1[d] = int(np.rint(f * np.float32(2048)))
    return idx, c0, c1


def resize_linear(src: np.ndarray, dw: int = ROI_W, dh: int = ROI_H, area_mode: bool = False) -> np.ndarray:
    """cv2.resize(src u8, (dw,dh)) with the default INTER_LINEAR (exact 2x down-scaling is routed to INTER_AREA);
    area_mode: the bilinear variant INTER_AREA falls back to when either axis up-scales."""
    sh, sw = src.shape
    if sw == 2 * dw and sh == 2 * dh:
        return resize_area(src, dw, dh)
    xi, xa0, xa1 = _linear_tab(sw, dw, area_mode=area_mode)
    yi, yb0, yb1 = _linear_tab(sh, dh, clamp_frac=False, area_mode=area_mode)
    S = src.astype(np.int64)
    x1 = np.minimum(xi + 1, sw - 1)
    rows = S[:, xi] * xa0[None, :] + S[:, x1] * xa1[None, :]            # horizontal pass, 11-bit fixed point
    r0, r1 = rows[np.clip(yi, 0, sh - 1)], rows[np.clip(yi + 1, 0, sh - 1)]
    out = (((yb0[:, None] * (r0 >> 4)) >> 16) + ((yb1[:, None] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def crop_roi(flavor: str, frame_bgr: np.ndarray, center, mouth_w: float):
    """-> uint8 [48,96] or None: crop_roi_gray (flavor 'live') / crop_roi (flavor 'record')."""
    h, w = frame_bgr.shape[:2]
    box = crop_box(flavor, w, h, center, mouth_w)
    if box is None:
        return None
    x1, y1, x2, y2 = box
    gray = bgr2gray(frame_bgr[y1:y2, x1:x2])
    return resize_area(gray) if flavor == "live" else resize_linear(gray)
