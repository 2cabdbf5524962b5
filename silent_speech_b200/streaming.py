"""Sliding-window inference for many concurrent camera streams (BASELINE.json configs[4]).

Windowing policy = the reference's only streaming loop, inactive/live_feed.py:155-213, driving the official
points(+ROI) model:

    buf = deque(maxlen=max_t)                 # per stream
    every camera frame:  frame_idx += 1
        no face found -> continue                      (nothing appended, NO prediction; live_feed.py:178-184)
        buf.append(feature row)
        if len(buf) >= WARMUP_MIN and frame_idx % PRED_EVERY == 0:
            window = buf zero-padded to max_t ; logits = model(window, lengths=len(buf))

Design: the per-stream ring of rows lives in HBM ([S, max_t, 180 + 32] fp32 = 76 KB per stream, 95 MB for 1,250
streams); a frame's ROI-CNN embedding is frame-local, so it is computed ONCE when the frame arrives (one K2 launch
over the tick's new frames) and cached in the ring next to the point features -- a window re-uses 89 of its 90
embeddings from earlier ticks.  Per tick: K2 on the new frames -> ring push -> gather the due windows ->
projection GEMMs + recurrences + pooled head on the gathered batch.  The integer state (head, count, frame counter)
is host-side numpy: it is deterministic from the calls, so no device synchronisation is needed to schedule.
Streams are independent: shard them over GPUs by stream id (no collective).
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import functional as F_ssb
from . import ops


class WindowState:
    """Host-side integer state of the per-stream deques (numpy only; covered by the CPU tests)."""

    def __init__(self, num_streams: int, max_t: int, pred_every: int = 2, warmup_min: Optional[int] = None):
        self.S, self.W = int(num_streams), int(max_t)
        self.pred_every = int(pred_every)
        self.warmup_min = min(10, max_t) if warmup_min is None else int(warmup_min)     # live_feed.py:162
        self.head = np.zeros(self.S, np.int64)          # next write position
        self.count = np.zeros(self.S, np.int64)         # rows buffered (<= W)
        self.frame_idx = np.zeros(self.S, np.int64)     # camera frames seen (live_feed.py:167)

    def reset(self, stream_ids):
        ids = np.asarray(stream_ids, np.int64)
        self.head[ids] = 0; self.count[ids] = 0; self.frame_idx[ids] = 0

    def advance(self, ids: np.ndarray, found: np.ndarray):
        """One camera frame per listed stream.  -> (appended ids, their ring positions, due ids, window start, window length)."""
        if np.unique(ids).shape[0] != ids.shape[0]:
            raise ValueError("stream ids must be unique within a tick")
        self.frame_idx[ids] += 1
        a = ids[found]
        pos = self.head[a].copy()
        self.head[a] = (pos + 1) % self.W
        self.count[a] = np.minimum(self.count[a] + 1, self.W)
        # a frame without a face hits `continue` before the predict block (live_feed.py:178-184): it advances the frame
        # counter but never produces a prediction, however full the window is
        due = found & (self.count[ids] >= self.warmup_min) & (self.frame_idx[ids] % self.pred_every == 0)
        d = ids[due]
        cnt = self.count[d]
        return a, pos, d, (self.head[d] - cnt) % self.W, cnt


class StreamScheduler:
    def __init__(self, model, num_streams: int, max_t: int = 90, pred_every: int = 2, warmup_min: Optional[int] = None):
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise ops._lib.SsbError("StreamScheduler needs the model on a CUDA device (no CPU path)")
        self.model, self.dev = model, dev
        self.state = WindowState(num_streams, max_t, pred_every, warmup_min)
        self.S, self.W = self.state.S, self.state.W
        self.dx = model.x_dim
        self.de = 32 if model.use_roi else 0
        self.ring = torch.zeros(self.S, self.W, self.dx + self.de, device=dev, dtype=torch.float32)

    def reset(self, stream_ids):
        self.state.reset(stream_ids)

    def _i32(self, a):
        return torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(self.dev, non_blocking=True)

    @torch.no_grad()
    def tick(self, stream_ids, X: torch.Tensor, R: Optional[torch.Tensor] = None, face_found=None, k_top: int = 0):
        """One camera frame for each listed stream (ids unique within a call).
        X f32 [n, x_dim] feature rows, R u8 [n,48,96] mouth ROIs (ROI model), face_found bool [n] (default all).
        Returns (due_ids int64 ndarray [m], logits f32 [m,C] on the device) for the streams whose window was
        classified this tick (and top-k ids/probs instead of logits when k_top > 0)."""
        with torch.cuda.device(self.dev):
            return self._tick(stream_ids, X, R, face_found, k_top)

    def _tick(self, stream_ids, X, R, face_found, k_top):
        ids = np.asarray(stream_ids, np.int64)
        n = ids.shape[0]
        found = np.ones(n, bool) if face_found is None else np.asarray(face_found, bool)
        a, pos, d, start, cnt = self.state.advance(ids, found)
        if a.size:
            all_found = a.size == n
            sel = np.nonzero(found)[0]
            Xs = X if all_found else X[torch.from_numpy(sel).to(X.device)]
            emb = None
            if self.de:
                Rs = R if all_found else R[torch.from_numpy(sel).to(R.device)]
                passes = F_ssb.cnn_passes(self.model)
                emb = ops.roi_cnn_fwd(Rs.view(a.size, 1, *Rs.shape[1:]).contiguous(), None, self.model.roi_cnn.tensors(),
                                      standardize=self.model.standardize_roi, passes=passes).view(a.size, self.de)
            ops.ring_push(self.ring, self._i32(a), self._i32(pos), Xs.contiguous().float(), emb)
        if d.size == 0:
            return d, None
        Xw, Ew, lengths = ops.ring_gather(self.ring, self._i32(d), self._i32(start), self._i32(cnt), self.W, self.dx, self.de)
        was_training = self.model.training
        self.model.eval()
        out, _ = F_ssb.forward_impl(self.model, Xw, lengths, None, training=False, save=False, k_top=k_top, emb=Ew)
        if was_training:
            self.model.train()
        if k_top:
            return d, (out["topk_ids"], out["topk_probs"])
        return d, out["logits"]
