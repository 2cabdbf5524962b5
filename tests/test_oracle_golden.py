"""The oracle against the committed outputs of the reference itself
(tests/golden/*.npz, produced by oracle/make_golden.py from /root/reference)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_explicit, ref_features, ref_model


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_landmark_set_known_answers():
    # record_landmarks_official.py:44-46 prints "should be 88 points"; x_dim = 2K+4 = 180
    ids = ref_features.LANDMARK_IDS
    assert len(ids) == 88 and ids.max() == 436 and list(ids) == sorted(set(ids.tolist()))
    pos = {int(v): i for i, v in enumerate(ids)}
    assert (pos[61], pos[291], pos[13], pos[14]) == (11, 50, 1, 2)


@pytest.mark.parametrize("flavour", ["live", "record"])
def test_features_match_reference(golden_dir, flavour):
    g = _load(golden_dir, "features.npz")
    assert np.array_equal(g["ids"], ref_features.LANDMARK_IDS)
    lm, present = g["lm"], g["present"]
    for b in range(lm.shape[0]):
        o = ref_features.clip_features(lm[b], int(g["w"]), int(g["h"]), present=present[b], flavour=flavour)
        assert np.array_equal(o["valid"], g[f"valid_{flavour}"][b])
        assert np.array_equal(o["X"], g[f"X_{flavour}"][b])          # same numpy ops -> bit-exact
        assert np.array_equal(o["center"], g[f"center_{flavour}"][b])
        assert np.array_equal(o["xy_px"], g[f"xy_px_{flavour}"][b])   # the gather (live:147): pure data movement
    assert g["valid_live"].sum() < g["valid_live"].size              # gate / absent frames exercised
    # the capture loop's history rule (live:272-296): a missed detection between two used frames does not reset the
    # velocity (clip 0: frames 2 and 4 used, 3 undetected); an out-of-band face does, and a missed detection after it
    # keeps the history empty (clip 1: frame 5 out of band, 6 undetected, 7 used)
    K2 = 2 * 88
    assert present[0, 3] == 0 and g["valid_live"][0, 2] and g["valid_live"][0, 4] and g["X_live"][0, 4, K2] > 0
    assert present[1, 6] == 0 and not g["valid_live"][1, 5] and g["valid_live"][1, 7] and g["X_live"][1, 7, K2] == 0


def test_product_synth_params_equal_the_oracles():
    """bench.py and the examples take their weights from silent_speech_b200.synth (never from test infrastructure);
    the fixtures were generated with the oracle's generator: they must be the same function."""
    from silent_speech_b200 import synth
    for kw in (dict(), dict(use_roi=False), dict(scale=2.0), dict(num_classes=7)):
        a, b = ref_model.synth_params(11, **kw), synth.synth_params(11, **kw)
        assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)


def _inputs(g):
    p = ref_model.synth_params(int(g["param_seed"]), use_roi=bool(g["use_roi"]), scale=float(g["param_scale"]))
    X, L, y = torch.from_numpy(g["X"]), torch.from_numpy(g["lengths"]), torch.from_numpy(g["y"])
    R = torch.from_numpy(g["R"]) if bool(g["use_roi"]) else None
    return p, X, L, R, y


@pytest.mark.parametrize("name", ["model_roi.npz", "model_points.npz", "model_roi_hot.npz"])
def test_forward_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    p, X, L, R, y = _inputs(g)
    torch.set_num_threads(1)
    out = ref_model.forward(p, X, L, R, standardize_roi=True, want_intermediates=True)
    np.testing.assert_allclose(out["logits"].numpy(), g["logits_train_semantics"], rtol=0, atol=2e-6)
    live = ref_model.forward(p, X, L, R, standardize_roi=False)
    np.testing.assert_allclose(live.numpy(), g["logits_live_semantics"], rtol=0, atol=2e-6)
    if R is not None:
        np.testing.assert_allclose(out["roi_emb"].numpy(), g["roi_emb"], rtol=0, atol=2e-6)
        assert np.abs(g["logits_train_semantics"] - g["logits_live_semantics"]).max() > 1e-4   # the skew is real
    ids, probs = ref_model.topk(live[:1])
    assert ids[0].tolist() == g["top3_ids_clip0"].tolist()
    np.testing.assert_allclose(probs[0].numpy(), g["top3_probs_clip0"], atol=1e-6)
    # float64 explicit math agrees with ATen fp32 to fp32 noise
    ex = ref_explicit.forward(p, X, L, R)
    np.testing.assert_allclose(ex["logits"].numpy(), g["logits_train_semantics"], rtol=0, atol=2e-5)


@pytest.mark.parametrize("name", ["model_roi.npz", "model_points.npz", "model_roi_hot.npz"])
def test_train_step_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    p, X, L, R, y = _inputs(g)
    torch.set_num_threads(1)
    loss, _, grads = ref_model.loss_and_grads(p, X, L, R, y)
    assert abs(loss.item() - float(g["loss"])) < 1e-6
    for k, gr in grads.items():
        samp = gr.reshape(-1)[::97].numpy()
        ref = g[f"grad_samp/{k}"]
        assert np.abs(samp - ref).max() <= 1e-6 + 1e-5 * np.abs(ref).max(), k
        assert abs(gr.double().norm().item() - g[f"grad_stat/{k}"][0]) <= 1e-5 * max(1.0, g[f"grad_stat/{k}"][0]), k
    st = ref_model.AdamState(p)
    l2, _, gn = st.step(X, L, R, y)
    assert abs(gn.item() - float(g["grad_norm"])) < 1e-4 * float(g["grad_norm"])
    for k, v in st.params().items():
        ref = g[f"step_samp/{k}"]
        assert np.abs(v.reshape(-1)[::97].numpy() - ref).max() <= 1e-6, k


# ------------------------------------------------------------------ K6 ROI crop
@pytest.mark.parametrize("name", ["roi_crop.npz", "roi_crop_many.npz"])
def test_roi_crop_oracle_matches_reference_outputs(golden_dir, name):
    """oracle/ref_roi_crop.py against crops produced by the reference's own crop_roi_gray / crop_roi (cv2 4.13)."""
    from oracle import ref_roi_crop as rc
    g = _load(golden_dir, name)
    n_ok = 0
    for i, (c, mw) in enumerate(zip(g["centers"], g["widths"])):
        for flavor, key, okk, eps in (("live", "roi_live", "ok_live", 0.0), ("record", "roi_record", "ok_record", 1e-6)):
            scale = float(mw) + (eps if name == "roi_crop_many.npz" else 0.0)
            o = rc.crop_roi(flavor, g["frame"], c, scale)
            assert (o is not None) == bool(g[okk][i]), (i, flavor)
            if o is not None:
                assert np.array_equal(o, g[key][i]), (i, flavor, int(np.abs(o.astype(int) - g[key][i]).max()))
                n_ok += 1
    assert n_ok >= 6


def test_roi_crop_oracle_matches_cv2_on_seeded_sizes():
    """The restated OpenCV arithmetic (gray, INTER_AREA incl. its fast and up-scaling paths, fixed-point bilinear)
    against cv2 itself, where cv2 is importable (it is in the build container)."""
    cv2 = pytest.importorskip("cv2")
    from oracle import ref_roi_crop as rc
    rng = np.random.default_rng(5)
    img = cv2.GaussianBlur(rng.integers(0, 256, (300, 460, 3)).astype(np.uint8), (5, 5), 1.5)
    g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    assert np.array_equal(g, rc.bgr2gray(img))
    sizes = [(297, 37), (68, 171), (17, 280), (192, 96), (288, 240), (96, 48), (3, 3)]
    for it in range(60):
        w, h = sizes[it] if it < len(sizes) else (int(rng.integers(3, 420)), int(rng.integers(3, 290)))
        x, y = int(rng.integers(0, 460 - w)), int(rng.integers(0, 300 - h))
        src = np.ascontiguousarray(g[y:y + h, x:x + w])
        assert np.array_equal(cv2.resize(src, (96, 48), interpolation=cv2.INTER_AREA), rc.resize_area(src)), (w, h)
        assert np.array_equal(cv2.resize(src, (96, 48)), rc.resize_linear(src)), (w, h)


# ------------------------------------------------------------------ augmentation
def test_augment_oracle_matches_reference_getitem(golden_dir):
    """oracle/ref_augment.py against NPZWordDataset.__getitem__ run with scripted random decisions."""
    from oracle import ref_augment as ra
    g = _load(golden_dir, "augment.npz")
    for i in range(len(g["T"])):
        T = int(g["T"][i])
        drops = [int(d) for d in g["drops"][i] if d >= 0]
        noise = g["noise"][i][:T] if g["noise_on"][i] else None
        Xo, To, Ro = ra.augment_clip(g["X"][i][:T], g["R"][i][:T], 90, noise, drops)
        assert To == int(g["T_out"][i])
        assert np.array_equal(Xo, g["X_out"][i])
        assert np.array_equal(Ro, g["R_out"][i])
