"""Camera-side entry points on the GPU (silent_speech_b200.live): landmarks + frames -> collate layout -> logits in one call,
the press-to-record clip (live_infer_official.py:330-360) and the validation pass (train_model_official.py:449-484),
each against the CPU oracle composed the way the reference's loop composes its own functions."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_features, ref_model, ref_roi_crop  # noqa: E402


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


def _camera(B, T, seed, Hf=360, Wf=640):
    from silent_speech_b200 import synth
    lm, present = synth.landmarks(B, T, seed=seed, drop=0.12)
    # the synthetic faces are sized for 1280x720; at 640x360 the mouth would be 35..70 px wide: stretch x about the centre
    lm[..., 0] = 0.5 + (lm[..., 0] - 0.5) * 2.0
    lm[0, 3, 291, 0] += 0.4                                  # one detected frame outside the 60..150 px band
    rng = np.random.default_rng(seed)
    frames = rng.integers(0, 256, (B, T, Hf, Wf, 3)).astype(np.uint8)
    frames[:, :, ::3] //= 2
    return lm, present, frames


def _oracle_clip(lm, present, frames, max_t=None):
    """What the capture loop leaves in bufX / bufR for one clip (live:272-293), with the oracle's functions."""
    Hf, Wf = frames.shape[1:3]
    o = ref_features.clip_features(lm, Wf, Hf, present=present, compact=True)
    n = o["length"]
    used = np.nonzero(o["valid"])[0]
    R = np.zeros((lm.shape[0], 48, 96), np.uint8)
    for r, t in enumerate(used):
        roi = ref_roi_crop.crop_roi("live", frames[t], o["center"][r], float(o["mouth_w"][r]))
        if roi is not None:
            R[r] = roi
    return o["X"], R, n


def test_clips_from_camera_match_capture_loop():
    dev = _cuda()
    from silent_speech_b200 import clips_from_camera
    B, T = 3, 14
    lm, present, frames = _camera(B, T, seed=31)
    c = clips_from_camera(torch.from_numpy(lm).to(dev), torch.from_numpy(frames).to(dev), torch.from_numpy(present).to(dev))
    for b in range(B):
        X, R, n = _oracle_clip(lm[b], present[b], frames[b])
        assert int(c["lengths"][b]) == n and 0 < n < T
        got = c["X"][b].cpu().numpy()
        assert np.abs(got - X).max() <= 1e-5 * np.abs(X).max()
        assert np.all(got[n:] == 0)
        # crops are byte work: bit-exact (boxes come from the same centre / width to within a float ulp; a box edge that
        # falls within an ulp of an integer would move one pixel -- none does on this seed)
        assert np.array_equal(c["R"][b].cpu().numpy(), R), b
        assert np.all(c["R"][b, n:].cpu().numpy() == 0)


def test_classify_camera_clips_and_recorder():
    dev = _cuda()
    from silent_speech_b200 import BiGRUClassifier, ClipRecorder, classify_camera_clips
    B, T = 2, 16
    lm, present, frames = _camera(B, T, seed=32)
    p = ref_model.synth_params(41)
    m = BiGRUClassifier(180, 10, use_roi=True, standardize_roi=False)        # live_infer_official semantics (live:126-127)
    m.load_state_dict(p)
    m = m.to(dev).eval()
    out = classify_camera_clips(m, torch.from_numpy(lm).to(dev), torch.from_numpy(frames).to(dev), torch.from_numpy(present).to(dev))
    labels = {i: f"w{i}" for i in range(10)}
    for b in range(B):
        X, R, n = _oracle_clip(lm[b], present[b], frames[b])
        with torch.no_grad():
            ref = ref_model.forward(p, torch.from_numpy(X[None, :n]), torch.tensor([n]), torch.from_numpy(R[None, :n]), standardize_roi=False)
        got = out["logits"][b].cpu()
        assert float((got - ref[0]).abs().max() / ref.abs().max()) < 1e-4
        # the same clip through the press-to-record helper, trimmed to max_t = 9 (feats[:T], bufR[:T], live:344-353)
        rec = ClipRecorder(m, labels, max_t=9)
        rec.push(torch.from_numpy(lm[b, 0]), torch.from_numpy(frames[b, 0]))          # not recording yet: ignored
        rec.start()
        for t in range(T):
            rec.push(torch.from_numpy(lm[b, t]) if present[b, t] else None, torch.from_numpy(frames[b, t]))
        res = rec.stop()
        tt = min(n, 9)
        with torch.no_grad():
            ref9 = ref_model.forward(p, torch.from_numpy(X[None, :tt]), torch.tensor([tt]), torch.from_numpy(R[None, :tt]), standardize_roi=False)
        ids, probs = ref_model.topk(ref9)
        assert res["frames_used"] == n and [lab for lab, _ in res["top"]] == [labels[int(i)] for i in ids[0]]
        np.testing.assert_allclose([pr for _, pr in res["top"]], probs[0].numpy(), atol=1e-5)
    # "Too short." (live:340-342): fewer than five used frames -> no classification
    rec = ClipRecorder(m, labels)
    rec.start()
    for t in range(4):
        rec.push(torch.from_numpy(lm[0, 4 + t]), torch.from_numpy(frames[0, 4 + t]))
    assert rec.stop() is None


def test_evaluate_matches_reference_validation_loop():
    dev = _cuda()
    from silent_speech_b200 import BiGRUClassifier, evaluate, synth, top_confusions
    p = ref_model.synth_params(43)
    m = BiGRUClassifier(180, 10, use_roi=True)
    m.load_state_dict(p)
    m = m.to(dev).train()
    batches, loss, ok, n, pairs = [], 0.0, 0, 0, []
    for i, B in enumerate((5, 3)):
        d = synth.clip_batch(B, 10, mode="uniform", seed=50 + i)
        X, L, R, y = (torch.from_numpy(d[k]) for k in ("X", "lengths", "R", "y"))
        batches.append((X.to(dev), L.to(dev), R.to(dev), y.to(dev)))
        with torch.no_grad():                                             # train:457-471
            logits = ref_model.forward(p, X, L, R)
            loss += ref_model.loss_fn(logits, y).item() * B
            pred = logits.argmax(1)
            ok += int((pred == y).sum()); n += B
            pairs += list(zip(y.tolist(), pred.tolist()))
    res = evaluate(m, batches)
    assert m.training                                                     # mode restored
    assert res["n"] == n and abs(res["loss"] - loss / n) < 1e-5 and res["acc"] == ok / n
    conf = np.zeros((10, 10), np.int64)
    for t, q in pairs:
        conf[t, q] += 1
    assert np.array_equal(res["confusion"].numpy(), conf)
    strs = top_confusions(res["confusion"], {i: f"w{i}" for i in range(10)})
    assert len(strs) == min(6, int((conf - np.diag(np.diag(conf)) > 0).sum())) and all("→" in s for s in strs)
