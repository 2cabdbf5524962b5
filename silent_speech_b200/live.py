"""Camera-side entry points: from landmark tensors + camera frames to logits in one call, the press-to-record clip of
live_infer_official.py, and the trainer's no-grad validation pass.

    clips_from_camera      features (K1) -> ROI crops (K6) -> the collate_fn layout, all on the device
                           (live_infer_official.py:272-293 per frame: extract_feature, crop_roi_gray, zero frame when the crop
                           is rejected; record_landmarks_official.py:183-197 for flavor="record")
    classify_camera_clips  ... -> model(X, lengths, R) -> logits                       (live:344-358)
    ClipRecorder           press-to-record: start / push one camera frame / stop -> "too short" (< 5 used frames) or one
                           classification of the clip trimmed to max_t                    (live:330-360)
    evaluate               model.eval(), no_grad, loss + argmax accuracy + confusions   (train_model_official.py:449-484)

Nothing here computes on the host: every stage is a libssb200 kernel; torch is used for index bookkeeping only."""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch

from . import ops
from .landmarks import crop_rois, extract_features

MIN_CLIP_FRAMES = 5          # live_infer_official.py:340-342 ("Too short.")


def clips_from_camera(landmarks: torch.Tensor, frames: Optional[torch.Tensor], present: Optional[torch.Tensor] = None,
                      flavor: str = "live", max_t: Optional[int] = None, idxs=None, gate=(60.0, 150.0)):
    """landmarks f32 [B,T,n_lm,>=2] (MediaPipe's normalised x, y[, z] per camera frame), frames u8 [B,T,H,W,3] BGR or None
    (landmarks-only model), present u8 [B,T] (face detected; None = all).
    -> dict X f32 [B,T',2K+4], lengths i64 [B], R u8 [B,T',48,96] or None, valid u8 [B,T]; the used frames of each clip are
    packed to the front exactly as the capture loop's bufX / bufR hold them, T' = min(T, max_t) (live:345)."""
    if not landmarks.is_cuda:
        raise ops._lib.SsbError("clips_from_camera needs CUDA tensors (this package has no CPU path)")
    B, T = landmarks.shape[:2]
    if frames is None:
        raise ops._lib.SsbError("frames give the frame size (w, h) the landmarks are scaled by; for a landmarks-only "
                                "pipeline call extract_features(landmarks, w, h, compact=True) directly")
    with torch.cuda.device(landmarks.device):
        Hf, Wf = int(frames.shape[2]), int(frames.shape[3])
        f = extract_features(landmarks.contiguous(), Wf, Hf, idxs=idxs, present=present, gate=gate, compact=True)
        X, lengths, valid = f["X"], f["lengths"], f["valid"]
        # compact row r of clip b is the r-th used camera frame: stable sort of "not used" gives that order on the device
        order = torch.argsort(valid.logical_not().to(torch.uint8), dim=1, stable=True).to(torch.int32)
        frame_index = (order + (torch.arange(B, device=order.device, dtype=torch.int32) * T)[:, None]).reshape(-1).contiguous()
        # rows past a clip's length carry center = mouth_w = 0: the crop box is empty, the kernel returns a zero frame
        R, _ = crop_rois(frames.reshape(B * T, Hf, Wf, 3).contiguous(), f["center"].reshape(B * T, 2), f["mouth_w"].reshape(B * T),
                         flavor=flavor, frame_index=frame_index)
        R = R.view(B, T, ops.ROI_H, ops.ROI_W)
        if max_t is not None and T > max_t:
            X, R = X[:, :max_t].contiguous(), R[:, :max_t].contiguous()
            lengths = lengths.clamp(max=max_t)
        return {"X": X, "lengths": lengths, "R": R, "valid": valid}


@torch.no_grad()
def classify_camera_clips(model, landmarks, frames, present=None, flavor: str = "live", max_t: Optional[int] = None, k_top: int = 0):
    """Camera-side tensors -> logits f32 [B,C] (and top-k when k_top > 0).  Clips with no used frame get lengths 0 and must be
    filtered by the caller (the reference never classifies them: live:340)."""
    c = clips_from_camera(landmarks, frames, present, flavor, max_t)
    R = c["R"] if model.use_roi else None
    L = c["lengths"].clamp(min=1)                 # the kernels honour 1 <= lengths; empty clips are reported through `lengths`
    if k_top:
        ids, probs, logits = model.predict_topk(c["X"], L, R, k=k_top)
        return {"logits": logits, "topk_ids": ids, "topk_probs": probs, "lengths": c["lengths"]}
    return {"logits": model(c["X"], L, R), "lengths": c["lengths"]}


class ClipRecorder:
    """Press-to-record, as the live demo's key handler does it (live_infer_official.py:330-360): `start()` empties the
    buffers and the velocity history, `push()` is one camera frame while recording, `stop()` classifies the clip once.

    The per-frame work of the reference (extract_feature, crop_roi_gray) is deferred to `stop()`, where the whole clip
    goes through the batched kernels in one go; the result is the same because both are functions of the frame alone plus
    the velocity term, which the feature kernel carries across frames by the capture loop's own rule."""

    def __init__(self, model, id_to_label: Dict[int, str], max_t: int = 90, frame_size: Tuple[int, int] = (1280, 720),
                 n_landmarks: int = 478, device=None):
        self.model, self.id_to_label, self.max_t = model, id_to_label, int(max_t)
        self.dev = torch.device(device) if device is not None else next(model.parameters()).device
        if self.dev.type != "cuda":
            raise ops._lib.SsbError("ClipRecorder needs the model on a CUDA device (no CPU path)")
        self.w, self.h = frame_size
        self.n_lm = n_landmarks
        self.recording = False
        self._lm: List[torch.Tensor] = []
        self._fr: List[torch.Tensor] = []
        self._present: List[int] = []

    def start(self):
        self.recording = True
        self._lm, self._fr, self._present = [], [], []

    def push(self, face_landmarks: Optional[torch.Tensor], frame_bgr: torch.Tensor):
        """face_landmarks f32 [n_lm, >=2] or None when no face was detected; frame_bgr u8 [H,W,3] (host or device)."""
        if not self.recording:
            return
        if face_landmarks is None:
            self._lm.append(torch.zeros(self.n_lm, 3))
            self._present.append(0)
        else:
            lm = torch.as_tensor(face_landmarks, dtype=torch.float32)
            self._lm.append(lm[:, :3] if lm.shape[1] >= 3 else torch.cat([lm, lm.new_zeros(lm.shape[0], 3 - lm.shape[1])], 1))
            self._present.append(1)
        self._fr.append(torch.as_tensor(frame_bgr, dtype=torch.uint8))

    def stop(self, k: int = 3):
        """-> None when fewer than 5 frames were used ("Too short.", live:340-342), else
        [(label, prob), ...] top-k like topk_from_logits (live:223-226) plus the logits."""
        self.recording = False
        if not self._lm:
            return None
        lm = torch.stack([t.to(self.dev) for t in self._lm])[None].contiguous()
        fr = torch.stack([t.to(self.dev) for t in self._fr])[None].contiguous()
        present = torch.tensor(self._present, dtype=torch.uint8, device=self.dev)[None]
        c = clips_from_camera(lm, fr, present, flavor="live", max_t=None)
        n = int(c["lengths"][0])
        if n < MIN_CLIP_FRAMES:
            return None
        T = min(n, self.max_t)                                     # feats[:T], bufR[:T]  (live:344-353)
        X, R = c["X"][:, :T].contiguous(), c["R"][:, :T].contiguous()
        L = torch.tensor([T], dtype=torch.int64, device=self.dev)
        ids, probs, logits = self.model.predict_topk(X, L, R if self.model.use_roi else None, k=k)
        top = [(self.id_to_label[int(i)], float(p)) for i, p in zip(ids[0].tolist(), probs[0].tolist())]
        return {"top": top, "logits": logits[0], "frames_used": n}


@torch.no_grad()
def evaluate(model, batches: Iterable[Sequence[torch.Tensor]], label_smoothing: float = 0.05, num_classes: Optional[int] = None):
    """The trainer's validation pass (train_model_official.py:449-475): eval mode, no gradients, per batch
    `logits = model(X, lengths, R)`, loss summed with the batch size as weight, argmax accuracy, (true, predicted) pairs.
    batches yield (X, lengths, R or None, y) on the model's device.  -> dict loss, acc, n, confusion i64 [C,C] (rows = true).
    One host synchronisation at the end (the reference syncs every batch through .item())."""
    was_training = model.training
    model.eval()
    C = num_classes or model.num_classes
    dev = next(model.parameters()).device
    loss_sum = torch.zeros((), device=dev, dtype=torch.float64)
    correct = torch.zeros((), device=dev, dtype=torch.int64)
    conf = torch.zeros(C * C, device=dev, dtype=torch.int64)
    n = 0
    from .functional import forward_impl
    for X, lengths, R, y in batches:
        y = y.to(device=dev, dtype=torch.int64)
        with torch.cuda.device(dev):
            out, _ = forward_impl(model, X, lengths, R if model.use_roi else None, training=False, save=False,
                                  y=y.contiguous(), label_smoothing=label_smoothing)      # CE fused into the head kernel
        pred = out["logits"].argmax(dim=1)
        loss_sum += out["loss"].double().sum()
        correct += (pred == y).sum()
        conf += torch.bincount(y * C + pred, minlength=C * C)
        n += int(X.shape[0])
    if was_training:
        model.train()
    if n == 0:
        return {"loss": float("nan"), "acc": float("nan"), "n": 0, "confusion": conf.view(C, C).cpu()}
    return {"loss": float(loss_sum) / n, "acc": float(correct) / n, "n": n, "confusion": conf.view(C, C).cpu()}


def top_confusions(confusion: torch.Tensor, id_to_label: Dict[int, str], k: int = 6):
    """The k most frequent mistakes as "actual→predicted(n)" strings, the format train_model_official.py:79-91 prints after
    each epoch (k = 6 at its call site, train:480)."""
    c = confusion.clone()
    c.fill_diagonal_(0)
    flat = c.flatten()
    kk = min(k, int((flat > 0).sum()))
    if kk == 0:
        return []
    vals, idx = torch.topk(flat, kk)
    C = confusion.shape[0]
    return [f"{id_to_label[int(i) // C]}→{id_to_label[int(i) % C]}({int(v)})" for v, i in zip(vals.tolist(), idx.tolist())]
