"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (build container only).

    python oracle/make_golden.py            # needs /root/reference; writes tests/golden/

The reference (davdwan21/Silent-Speech) ships no tests or golden vectors, so
the pin for this path is its own code executed here on seeded inputs:

* train_model_official.BiGRUClassifier                (train_model_official.py:253-310)
  forward, CE(label_smoothing=.05) backward, clip_grad_norm_(1.0), Adam(3e-4) step
* live_infer_official.BiGRUClassifier                 (live_infer_official.py:101-138; no ROI standardise)
* live_infer_official.extract_feature / mouth_width_px / crop_roi_gray / topk_from_logits
* record_landmarks_official.extract_feature / crop_roi

``mediapipe`` is absent; four empty stub modules let the two capture scripts
import (their feature functions never touch it).  The reference cannot travel
to the GPU box, hence fixtures.  Weights are NOT stored: they are regenerated
from ``oracle.ref_model.synth_params(seed)`` (numpy PCG64, machine-independent)
and loaded into the reference classes with ``load_state_dict``.  Large
gradients are stored as (norm, sum, every-97th-element) summaries.
"""
from __future__ import annotations

import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from oracle import ref_model  # noqa: E402
from silent_speech_b200 import synth  # noqa: E402

STRIDE = 97


def import_reference():
    for name in ("mediapipe", "mediapipe.tasks", "mediapipe.tasks.python", "mediapipe.tasks.python.vision"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["mediapipe"].tasks = sys.modules["mediapipe.tasks"]
    sys.modules["mediapipe.tasks"].python = sys.modules["mediapipe.tasks.python"]
    sys.modules["mediapipe.tasks.python"].vision = sys.modules["mediapipe.tasks.python.vision"]
    sys.path.insert(0, REF)
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())           # record_landmarks_official makedirs("clips_npz") at import
    try:
        import train_model_official as train
        import live_infer_official as live
        import record_landmarks_official as record
    finally:
        os.chdir(cwd)
    return train, live, record


class _Pt:
    __slots__ = ("x", "y")

    def __init__(self, x, y):
        self.x, self.y = float(x), float(y)


def as_face(lm_frame):
    return [_Pt(p[0], p[1]) for p in lm_frame]


def summarise(t: torch.Tensor):
    f = t.detach().reshape(-1).double()
    return np.array([f.norm().item(), f.sum().item()]), t.detach().reshape(-1)[::STRIDE].numpy().copy()


def golden_features(live, record):
    B, T = 3, 14
    lm, present = synth.landmarks(B, T, seed=7, drop=0.15)
    lm[1, 5, 291, 0] += 0.2          # one frame outside the 60..150 px band
    out = {"lm": lm, "present": present, "w": np.int64(synth.FRAME_W), "h": np.int64(synth.FRAME_H),
           "ids": np.array(live.FIXED_IDXS, np.int32)}
    assert live.FIXED_IDXS == record.FIXED_IDXS and len(live.FIXED_IDXS) == 88
    # make sure the fixture holds the case the capture loop treats specially: a dropped detection BETWEEN two used
    # frames (velocity must bridge it) and a dropped detection right after an out-of-band frame (history stays reset)
    present[0, 3] = 0; present[0, 2] = 1; present[0, 4] = 1
    present[1, 6] = 0; present[1, 4] = 1; present[1, 7] = 1
    w, h = synth.FRAME_W, synth.FRAME_H
    for tag, mod in (("live", live), ("record", record)):
        X = np.zeros((B, T, 180), np.float32); C = np.zeros((B, T, 2), np.float32)
        MW = np.zeros((B, T), np.float64); V = np.zeros((B, T), np.uint8); XY = np.zeros((B, T, 88, 2), np.float32)
        for b in range(B):
            prev_xy = None
            recording = True
            for t in range(T):
                # the body of the capture loop, live_infer_official.py:272-296 / record_landmarks_official.py:178-201,
                # statement for statement: `res.face_landmarks` is empty when no face was detected
                face_landmarks = [as_face(lm[b, t])] if present[b, t] else []
                if face_landmarks:
                    face = face_landmarks[0]
                    mw = mod.mouth_width_px(face, w, h)
                    in_range = mod.MOUTH_W_MIN_PX <= mw <= mod.MOUTH_W_MAX_PX
                    if recording and in_range:
                        feat, prev_xy, center, mw_ret = mod.extract_feature(face, w, h, mod.FIXED_IDXS, prev_xy)
                        X[b, t] = feat; C[b, t] = center; MW[b, t] = mw_ret; V[b, t] = 1
                        # `xy` of live:147 / record:64, the gather itself (not returned by extract_feature)
                        XY[b, t] = np.array([[face[i].x * w, face[i].y * h] for i in mod.FIXED_IDXS], dtype=np.float32)
                    else:
                        if recording:
                            prev_xy = None
        out[f"X_{tag}"] = X; out[f"center_{tag}"] = C; out[f"mouth_w_{tag}"] = MW; out[f"valid_{tag}"] = V
        out[f"xy_px_{tag}"] = XY
    # a bridged frame really carries a velocity (otherwise the fixture would not distinguish the two readings)
    assert out["valid_live"][0, 2] and out["valid_live"][0, 4] and out["X_live"][0, 4, 176] > 0
    np.savez_compressed(os.path.join(OUT, "features.npz"), **out)
    print("features.npz", {k: getattr(v, "shape", v) for k, v in out.items()})


def golden_model(train, live, tag, use_roi, B, T, lens, seed, scale=1.0):
    p = ref_model.synth_params(seed, use_roi=use_roi, scale=scale)
    lens = np.asarray(lens, np.int64)
    X = synth.point_features(B, T, lens, seed=seed)
    R = synth.rois(B, T, lens, seed=seed) if use_roi else None
    y = synth.labels(B, seed=seed)
    tX, tL, ty = torch.from_numpy(X), torch.from_numpy(lens), torch.from_numpy(y)
    tR = torch.from_numpy(R) if use_roi else None
    out = {"param_seed": np.int64(seed), "param_scale": np.float64(scale), "use_roi": np.bool_(use_roi),
           "X": X, "lengths": lens, "y": y}
    if use_roi:
        out["R"] = R

    m = train.BiGRUClassifier(180, 10, use_roi=use_roi, roi_emb=32, hidden=192)
    m.load_state_dict(p)
    m.eval()
    with torch.no_grad():
        logits = m(tX, tL, tR)
        out["logits_train_semantics"] = logits.numpy()
        if use_roi:
            r = (tR.float() / 255.0).unsqueeze(2)
            mu = r.mean(dim=(2, 3, 4), keepdim=True); sd = r.std(dim=(2, 3, 4), keepdim=True).clamp_min(1e-6)
            out["roi_emb"] = m.roi_cnn((r - mu) / sd).numpy()
    ml = live.BiGRUClassifier(180, 10, use_roi=use_roi, roi_emb=32, hidden=192, gru_layers=2)
    ml.load_state_dict(p)
    ml.eval()
    with torch.no_grad():
        ll = ml(tX, tL, tR)
        out["logits_live_semantics"] = ll.numpy()
    top = live.topk_from_logits(ll[:1], {i: i for i in range(10)}, k=3)
    out["top3_ids_clip0"] = np.array([t[0] for t in top], np.int64)
    out["top3_probs_clip0"] = np.array([t[1] for t in top], np.float64)

    # one train step exactly as train_model_official.py:402-405,433-439, dropout inactive (eval)
    opt = torch.optim.Adam(m.parameters(), lr=train.LR)
    loss_fn = torch.nn.CrossEntropyLoss(label_smoothing=0.05)
    loss = loss_fn(m(tX, tL, tR), ty)
    opt.zero_grad()
    loss.backward()
    out["loss"] = np.float64(loss.item())
    for name, prm in m.named_parameters():
        out[f"grad_stat/{name}"], out[f"grad_samp/{name}"] = summarise(prm.grad)
    gn = torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
    out["grad_norm"] = np.float64(gn.item())
    opt.step()
    for name, prm in m.named_parameters():
        out[f"step_stat/{name}"], out[f"step_samp/{name}"] = summarise(prm)
    np.savez_compressed(os.path.join(OUT, f"model_{tag}.npz"), **out)
    print(f"model_{tag}.npz loss={loss.item():.6f} gnorm={gn.item():.4f}")


def golden_roi_crop(live, record):
    rng = np.random.default_rng(11)
    frame = rng.integers(0, 256, size=(360, 640, 3)).astype(np.uint8)
    # smooth it a little so INTER_AREA / INTER_LINEAR differences are meaningful but bounded
    import cv2
    frame = cv2.GaussianBlur(frame, (7, 7), 2.0)
    centers = np.array([[320.3, 200.7], [30.2, 20.9], [600.0, 340.0], [100.5, 180.25]], np.float32)
    widths = np.array([80.0, 61.5, 120.0, 1.0], np.float32)
    outs_live, outs_rec, ok_live, ok_rec = [], [], [], []
    for c, mw in zip(centers, widths):
        a = live.crop_roi_gray(frame, c, float(mw)); b = record.crop_roi(frame, c, float(mw))
        ok_live.append(a is not None); ok_rec.append(b is not None)
        outs_live.append(a if a is not None else np.zeros((48, 96), np.uint8))
        outs_rec.append(b if b is not None else np.zeros((48, 96), np.uint8))
    np.savez_compressed(os.path.join(OUT, "roi_crop.npz"), frame=frame, centers=centers, widths=widths,
                        roi_live=np.stack(outs_live), roi_record=np.stack(outs_rec),
                        ok_live=np.array(ok_live), ok_record=np.array(ok_rec))
    print("roi_crop.npz", ok_live, ok_rec)


def golden_roi_crop_many(live, record):
    """40 seeded crops on a 240x320 frame: interior, clipped by each border, smaller than 96x48 (up-scaling
    fallbacks), exact integer ratios (area-fast, and the 2x case INTER_LINEAR routes to INTER_AREA), rejected boxes."""
    import cv2
    rng = np.random.default_rng(12)
    frame = cv2.GaussianBlur(rng.integers(0, 256, size=(240, 320, 3)).astype(np.uint8), (5, 5), 1.2)
    centers = np.stack([rng.uniform(-5, 325, 40), rng.uniform(-5, 245, 40)], 1).astype(np.float32)
    widths = rng.uniform(2.0, 75.0, 40)
    centers[:6] = [[160, 120], [160.5, 120.5], [160, 120], [3.2, 100.0], [318.0, 236.0], [160, 120]]
    widths[:6] = [80.0, 40.0, 48.0, 30.0, 25.0, 0.5]              # 192x160, 96x80, 2x-ish, clipped, clipped, rejected
    widths = np.array([float(np.float32(w)) for w in widths])     # mouth widths are float32 norms turned into Python floats
    outs_live, outs_rec, ok_live, ok_rec = [], [], [], []
    for c, mw in zip(centers, widths):
        a = live.crop_roi_gray(frame, c, float(mw)); b = record.crop_roi(frame, c, float(mw) + 1e-6)
        ok_live.append(a is not None); ok_rec.append(b is not None)
        outs_live.append(a if a is not None else np.zeros((48, 96), np.uint8))
        outs_rec.append(b if b is not None else np.zeros((48, 96), np.uint8))
    np.savez_compressed(os.path.join(OUT, "roi_crop_many.npz"), frame=frame, centers=centers, widths=widths,
                        roi_live=np.stack(outs_live), roi_record=np.stack(outs_rec),
                        ok_live=np.array(ok_live), ok_record=np.array(ok_rec))
    print("roi_crop_many.npz", int(np.sum(ok_live)), int(np.sum(ok_rec)))


def golden_augment(train):
    """NPZWordDataset.__getitem__ (train:133-172) on temporary .npz clips with its random calls scripted, so that the
    decisions are known: (noise?, drops) per clip.  Stores inputs, decisions, the injected noise and the outputs."""
    import random as pyrandom
    rng = np.random.default_rng(13)
    tmp = tempfile.mkdtemp()
    cases = [(30, False, []), (30, True, [5]), (40, True, [1, 38]), (12, True, []), (90, False, [44, 45]), (20, False, [18])]
    rec = {"X": [], "R": [], "T": [], "noise_on": [], "noise": [], "drops": [], "X_out": [], "T_out": [], "R_out": []}
    real = (pyrandom.random, pyrandom.randint, np.random.normal, np.random.choice)
    try:
        for i, (T, noise_on, drops) in enumerate(cases):
            X = rng.normal(0, 0.3, (T, 180)).astype(np.float32)
            R = ((np.arange(T)[:, None, None] * 7 + np.arange(48)[None, :, None] + 3 * np.arange(96)[None, None, :]) % 256).astype(np.uint8)
            path = os.path.join(tmp, f"c{i}.npz")
            np.savez(path, X=X, roi=R, label="w")
            noise = rng.normal(0, 0.01, (T, 180))
            seq = iter([0.0 if noise_on else 0.99, 0.0 if drops else 0.99])
            pyrandom.random = lambda: next(seq)
            pyrandom.randint = lambda a, b, k=len(drops): k
            np.random.normal = lambda m, s_, size, nz=noise: nz
            np.random.choice = lambda a, size, replace, d=drops: np.array(d)
            ds = train.NPZWordDataset([path], {"w": 0}, max_t=90, augment=True, use_roi=True)
            Xo, To, Ro, _ = ds[0]
            rec["X"].append(np.pad(X, ((0, 90 - T), (0, 0)))); rec["R"].append(np.pad(R, ((0, 90 - T), (0, 0), (0, 0))))
            rec["T"].append(T); rec["noise_on"].append(noise_on); rec["noise"].append(np.pad(noise.astype(np.float32), ((0, 90 - T), (0, 0))))
            rec["drops"].append(drops + [-1] * (2 - len(drops)))
            rec["X_out"].append(Xo.numpy()); rec["T_out"].append(int(To)); rec["R_out"].append(Ro.numpy())
    finally:
        pyrandom.random, pyrandom.randint, np.random.normal, np.random.choice = real
    np.savez_compressed(os.path.join(OUT, "augment.npz"), **{k: np.array(v) for k, v in rec.items()})
    print("augment.npz", rec["T_out"])


def golden_dataset(train):
    """The ingestion path on a small synthetic clip directory (schema train:1-9): ragged lengths around max_t, a ROI
    shorter than X, a clip without ROI, a clip of another feature dim (dropped by the most-common-dim rule, train:338-351).
    Expected values come from the reference's own split_by_label, NPZWordDataset(augment=False) + collate_fn and
    torch's WeightedRandomSampler."""
    from collections import Counter
    from torch.utils.data import WeightedRandomSampler
    rng = np.random.default_rng(29)
    tmp = tempfile.mkdtemp()
    words = ["water"] * 6 + ["hello"] * 4 + ["apple"] * 3 + ["water"]
    Ts = [5, 30, 90, 95, 100, 13, 44, 89, 91, 7, 60, 61, 12, 20]
    rec, names = {}, []
    for i, (w, T) in enumerate(zip(words, Ts)):
        D = 120 if i == 13 else 180
        X = (rng.integers(-64, 64, (T, D)) / 64.0).astype(np.float32)          # low-entropy: keeps the fixture small
        Tr = T - 3 if i in (1, 4) else (T + 2 if i == 6 else T)
        R = ((np.arange(Tr)[:, None, None] * 5 + i * 11 + np.arange(48)[None, :, None] + 2 * np.arange(96)[None, None, :]) % 256).astype(np.uint8)
        name = f"clip_{i:02d}.npz"
        kw = dict(X=X, ts=np.arange(T, dtype=np.int32), label=w, speaker="s0", idxs=np.arange(88, dtype=np.int32))
        if i != 9:
            kw["roi"] = R
        np.savez(os.path.join(tmp, name), **kw)
        names.append(name)
        rec[f"X_{i}"] = X; rec[f"label_{i}"] = np.array(w)
        rec[f"roi_tr_{i}"] = np.array(Tr if i != 9 else -1)
    files = [os.path.join(tmp, n) for n in names if n != "clip_13.npz"]
    labels = [w for n, w in zip(names, words) if n != "clip_13.npz"]
    uniq = sorted(set(labels))
    l2i = {lab: k for k, lab in enumerate(uniq)}
    import io, contextlib
    with contextlib.redirect_stdout(io.StringIO()):
        tr, va = train.split_by_label(list(files), list(labels), 0.15, seed=42)
    ds = train.NPZWordDataset(files, l2i, max_t=90, augment=False, use_roi=True)
    X, T, R, y = train.collate_fn([ds[k] for k in range(len(files))])
    tl = [str(np.load(f, allow_pickle=True)["label"]) for f in tr]
    cnt = Counter(tl)
    w8 = torch.tensor([1.0 / cnt[lab] for lab in tl], dtype=torch.double)
    g = torch.Generator(); g.manual_seed(123)
    order = list(iter(WeightedRandomSampler(w8, num_samples=len(w8), replacement=True, generator=g)))
    rec.update(names=np.array(names), kept=np.array([os.path.basename(f) for f in files]), classes=np.array(uniq),
               train=np.array([os.path.basename(f) for f in tr]), val=np.array([os.path.basename(f) for f in va]),
               X=X.numpy(), T=T.numpy(), R=R.numpy(), y=y.numpy(), weights=w8.numpy(), order=np.array(order, np.int64))
    np.savez_compressed(os.path.join(OUT, "dataset.npz"), **rec)
    print("dataset.npz", T.tolist(), [os.path.basename(f) for f in va], order)


def golden_legacy():
    """(1) the reference's GRUClassifier (inactive/train_reduced.py:129-145) with the shipped trained
    weights inactive/word_model_5.pt on synthetic zero-padded clips -> out / pooled / logits; the weights travel in the
    fixture because /root/reference does not.  (2) the mouth-open EMA + hysteresis loop of important_landmarks.py:118-144,
    re-run here with the reference's own dist2d and constants (the loop body is inline in its main(), not callable)."""
    sys.path.insert(0, os.path.join(REF, "inactive"))
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())
    try:
        import train_reduced as tr5
        import important_landmarks as il
    finally:
        os.chdir(cwd)
    ck = torch.load(os.path.join(REF, "inactive", "word_model_5.pt"), map_location="cpu")
    model = tr5.GRUClassifier(int(ck["input_dim"]), len(ck["words"]), hidden=64)
    model.load_state_dict(ck["model"])
    model.eval()
    rng = np.random.default_rng(31)
    B, T, D = 7, int(ck["max_t"]), int(ck["input_dim"])
    lens = [60, 41, 7, 60, 23, 1, 55]
    x = (rng.integers(-96, 96, (B, T, D)) / 64.0).astype(np.float32)
    for b, n in enumerate(lens):
        x[b, n:] = 0.0                                           # pad_or_trim's zero rows (the GRU still runs over them)
    with torch.no_grad():
        xt = torch.from_numpy(x)
        out, _ = model.gru(xt)
        pooled = out.mean(dim=1)
        logits = model(xt)
    rec = {"sd_" + k: v.numpy() for k, v in ck["model"].items()}
    rec.update(x=x, out=out.numpy(), pooled=pooled.numpy(), logits=logits.numpy(), max_t=np.array(T),
               words=np.array(ck["words"]))
    # ---- mouth-open detector: S streams x F frames of 478 landmarks; only ids 13, 14, 33, 263 matter
    S, F = 5, 80
    lm = np.zeros((S, F, 478, 3), np.float32)
    lm[..., :2] = rng.uniform(0.3, 0.7, (S, F, 478, 2)).astype(np.float32)
    tt = np.arange(F)[None, :] / 9.0 + np.arange(S)[:, None]
    gap = (0.012 + 0.012 * np.sin(tt) + rng.normal(0, 0.002, (S, F))).astype(np.float32)      # straddles 0.02 * eye span
    lm[:, :, il.LEFT_EYE_CORNER, 0] = 0.40; lm[:, :, il.RIGHT_EYE_CORNER, 0] = 0.60
    lm[:, :, il.LEFT_EYE_CORNER, 1] = 0.45; lm[:, :, il.RIGHT_EYE_CORNER, 1] = 0.46
    lm[:, :, il.MOUTH_TOP, 1] = 0.60
    lm[:, :, il.MOUTH_BOTTOM, 1] = 0.60 + gap * 0.4
    valid = (rng.uniform(size=(S, F)) > 0.1).astype(np.uint8)
    ema_tr, open_tr = np.zeros((S, F)), np.zeros((S, F), np.uint8)
    for s_ in range(S):
        mouth_state_open, mouth_ema = False, 0.0                 # important_landmarks.py:108-109
        for f in range(F):
            if valid[s_, f]:
                face = as_face(lm[s_, f])
                lip_gap = abs(face[il.MOUTH_BOTTOM].y - face[il.MOUTH_TOP].y)
                eye_span = il.dist2d(face[il.LEFT_EYE_CORNER], face[il.RIGHT_EYE_CORNER]) + 1e-6
                openness = lip_gap / eye_span
                mouth_ema = (1 - il.EMA_ALPHA) * mouth_ema + il.EMA_ALPHA * openness
                if mouth_state_open:
                    if mouth_ema < il.CLOSE_THR:
                        mouth_state_open = False
                else:
                    if mouth_ema > il.OPEN_THR:
                        mouth_state_open = True
            ema_tr[s_, f], open_tr[s_, f] = mouth_ema, mouth_state_open
    rec.update(m_lm4=lm[:, :, [il.MOUTH_TOP, il.MOUTH_BOTTOM, il.LEFT_EYE_CORNER, il.RIGHT_EYE_CORNER], :2].copy(),
               m_ids=np.array([il.MOUTH_TOP, il.MOUTH_BOTTOM, il.LEFT_EYE_CORNER, il.RIGHT_EYE_CORNER]),
               m_valid=valid, m_ema=ema_tr, m_open=open_tr,
               m_const=np.array([il.EMA_ALPHA, il.OPEN_THR, il.CLOSE_THR]))
    np.savez_compressed(os.path.join(OUT, "legacy.npz"), **rec)
    print("legacy.npz logits", logits.numpy().round(3).tolist()[:2], "open frames", int(open_tr.sum()), "of", S * F,
          "toggles", int(np.abs(np.diff(open_tr.astype(int), axis=1)).sum()))


def golden_checkpoint(train, live):
    """The checkpoint contract (train_model_official.py:489-500 written, live_infer_official.py:198-221 read) and the
    constructor's seed-42 initialisation (train:45-47, 402), both directions, executed against the reference here:
      * the reference constructor under torch.manual_seed(42): per-tensor (norm, sum) + strided samples -> `init/...`
      * a file written by silent_speech_b200.checkpoint.save is loaded by the reference's own load_classifier and run
        on a fixed input -> `ref_loaded_ours_logits` (+ the returned id_to_label / max_t / use_roi)
      * a file written the way train:489-500 writes it (the reference's model, the same dict literal) is loaded by
        silent_speech_b200.load_classifier(device="cpu") and must hold identical tensors
    The .pt files themselves are 4.8 MB each and are not committed; the CPU tests rebuild both files from seeds."""
    import silent_speech_b200 as ssb
    from silent_speech_b200 import checkpoint as ckpt_io
    out = {}
    torch.manual_seed(42)
    ref = train.BiGRUClassifier(180, 10, use_roi=True, roi_emb=32, hidden=192)
    sd = ref.state_dict()
    out["names"] = np.array(list(sd.keys()))
    for k, v in sd.items():
        out[f"init_stat/{k}"], out[f"init_samp/{k}"] = summarise(v)
        out[f"shape/{k}"] = np.array(v.shape, np.int64)
    torch.manual_seed(42)
    ours = ssb.BiGRUClassifier(180, 10, use_roi=True)
    osd = ours.state_dict()
    assert list(osd.keys()) == list(sd.keys())
    assert all(torch.equal(osd[k], sd[k]) for k in sd), "seed-42 init differs from the reference constructor"
    labels = ["hello", "yes", "no", "please", "thanks", "stop", "go", "left", "right", "help"]
    tmp = tempfile.mkdtemp()
    # ours -> reference loader
    path_ours = os.path.join(tmp, "ours.pt")
    ckpt_io.save(path_ours, ours, labels, max_t=90, seed=42)
    model_r, id_to_label, max_t, use_roi = live.load_classifier(path_ours)
    rng = np.random.default_rng(5)
    X = rng.normal(0, 0.25, (2, 9, 180)).astype(np.float32)
    R = rng.integers(0, 256, (2, 9, 48, 96)).astype(np.uint8)
    L = np.array([9, 6], np.int64)
    with torch.no_grad():
        lg = model_r(torch.from_numpy(X), torch.from_numpy(L), torch.from_numpy(R))
    out.update(X=X, R=R, lengths=L, ref_loaded_ours_logits=lg.numpy(), ref_loaded_max_t=np.int64(max_t),
               ref_loaded_use_roi=np.bool_(use_roi), ref_loaded_labels=np.array([id_to_label[i] for i in range(10)]),
               labels=np.array(labels))
    # reference -> our loader (dict literal of train:489-500)
    path_ref = os.path.join(tmp, "ref.pt")
    label_to_id = {lab: i for i, lab in enumerate(labels)}
    id_to_label_w = {i: lab for lab, i in label_to_id.items()}
    torch.save({"model": ref.state_dict(), "x_dim": 180, "max_t": train.MAX_T, "use_roi": True, "roi_w": train.ROI_W,
                "roi_h": train.ROI_H, "labels": labels, "label_to_id": label_to_id, "id_to_label": id_to_label_w,
                "seed": train.SEED}, path_ref)
    out["ref_ckpt_keys"] = np.array(sorted(torch.load(path_ref, map_location="cpu").keys()))
    m2, id2, mt2, ur2 = ssb.load_classifier(path_ref, device="cpu")
    assert all(torch.equal(m2.state_dict()[k], sd[k]) for k in sd) and mt2 == train.MAX_T and ur2 and id2 == id_to_label_w
    out["ours_ckpt_keys"] = np.array(sorted(torch.load(path_ours, map_location="cpu").keys()))
    np.savez_compressed(os.path.join(OUT, "checkpoint.npz"), **out)
    print("checkpoint.npz: init identical, ours->reference loader ok, reference->our loader ok; keys", out["ref_ckpt_keys"].tolist())


def main():
    os.makedirs(OUT, exist_ok=True)
    train, live, record = import_reference()
    torch.set_num_threads(1)                 # deterministic reduction order for the fixtures
    golden_features(live, record)
    golden_checkpoint(train, live)
    golden_model(train, live, "roi", True, B=4, T=12, lens=[12, 5, 1, 9], seed=101)
    golden_model(train, live, "points", False, B=5, T=20, lens=[20, 6, 13, 1, 17], seed=102)
    golden_model(train, live, "roi_hot", True, B=3, T=10, lens=[7, 10, 3], seed=103, scale=2.0)
    golden_roi_crop(live, record)
    golden_roi_crop_many(live, record)
    golden_augment(train)
    golden_dataset(train)
    golden_legacy()


if __name__ == "__main__":
    main()
