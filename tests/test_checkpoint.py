"""The checkpoint contract and the constructor's initialisation (CPU; no compute kernels involved).

Reference: the trainer writes a dict (train_model_official.py:489-500), the live demo reads it back
(live_infer_official.py:198-221, strict load_state_dict), and the trainer seeds torch with 42 before building the model
(train:45-47, 402).  tests/golden/checkpoint.npz was produced by oracle/make_golden.py::golden_checkpoint, which ran the
reference's own constructor and its own load_classifier on a file written by silent_speech_b200.checkpoint.save."""
import os

import numpy as np
import pytest
import torch

from silent_speech_b200 import BiGRUClassifier, checkpoint, load_classifier

LABELS = ["hello", "yes", "no", "please", "thanks", "stop", "go", "left", "right", "help"]


@pytest.fixture(scope="module")
def g(golden_dir):
    return np.load(os.path.join(golden_dir, "checkpoint.npz"))


def _seed42_model(use_roi=True):
    torch.manual_seed(42)
    return BiGRUClassifier(180, 10, use_roi=use_roi)


def test_seed42_init_equals_reference_constructor(g):
    sd = _seed42_model().state_dict()
    assert list(sd.keys()) == g["names"].tolist()                    # same 32 names, same order
    for k, v in sd.items():
        assert tuple(v.shape) == tuple(g[f"shape/{k}"]) and v.dtype == torch.float32, k
        samp = v.reshape(-1)[::97].numpy()
        assert np.array_equal(samp, g[f"init_samp/{k}"]), k          # bit-identical draws from the shared torch stream
        f = v.reshape(-1).double()
        assert abs(f.norm().item() - g[f"init_stat/{k}"][0]) <= 1e-12 * max(1.0, g[f"init_stat/{k}"][0]), k
        assert abs(f.sum().item() - g[f"init_stat/{k}"][1]) <= 1e-9, k
    assert sum(v.numel() for v in sd.values()) == 1_190_819


def test_saved_file_has_the_reference_dict_layout(tmp_path, g):
    m = _seed42_model()
    path = str(tmp_path / "ours.pt")
    checkpoint.save(path, m, LABELS, max_t=90, seed=42)
    ck = torch.load(path, map_location="cpu")
    assert sorted(ck.keys()) == g["ref_ckpt_keys"].tolist() == g["ours_ckpt_keys"].tolist()
    assert ck["x_dim"] == 180 and ck["max_t"] == 90 and ck["use_roi"] is True and (ck["roi_w"], ck["roi_h"]) == (96, 48)
    assert ck["labels"] == LABELS and ck["label_to_id"] == {lab: i for i, lab in enumerate(LABELS)}
    assert ck["id_to_label"] == {i: lab for i, lab in enumerate(LABELS)} and ck["seed"] == 42
    assert list(ck["model"].keys()) == g["names"].tolist()
    assert all(isinstance(v, torch.Tensor) and v.dtype == torch.float32 and v.device.type == "cpu" for v in ck["model"].values())
    # what the reference's loader returned for this very file (recorded at fixture time)
    assert g["ref_loaded_labels"].tolist() == LABELS and int(g["ref_loaded_max_t"]) == 90 and bool(g["ref_loaded_use_roi"])


def test_reference_written_file_loads_unchanged(tmp_path, g):
    """A file with the dict literal of train:489-500 (weights = the seed-42 constructor's, which
    test_seed42_init_equals_reference_constructor ties to the reference's) -> load_classifier -> identical tensors."""
    src = _seed42_model()
    label_to_id = {lab: i for i, lab in enumerate(LABELS)}
    path = str(tmp_path / "ref.pt")
    torch.save({"model": src.state_dict(), "x_dim": 180, "max_t": 90, "use_roi": True, "roi_w": 96, "roi_h": 48,
                "labels": LABELS, "label_to_id": label_to_id, "id_to_label": {i: l for l, i in label_to_id.items()},
                "seed": 42}, path)
    m, id_to_label, max_t, use_roi = load_classifier(path, device="cpu")
    assert max_t == 90 and use_roi and id_to_label == {i: l for i, l in enumerate(LABELS)} and not m.training
    assert m.standardize_roi is False                                  # live_infer_official semantics by default (live:126-127)
    a, b = m.state_dict(), src.state_dict()
    assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)
    m2, ck = checkpoint.load(path, device="cpu")
    assert m2.standardize_roi is True and all(torch.equal(m2.state_dict()[k], b[k]) for k in b)
    # strictness, as live:220: a missing tensor is an error, not a silent default
    bad = torch.load(path, map_location="cpu")
    del bad["model"]["head.4.bias"]
    torch.save(bad, path)
    with pytest.raises(RuntimeError):
        load_classifier(path, device="cpu")


def test_landmarks_only_checkpoint_roundtrip(tmp_path):
    m = _seed42_model(use_roi=False)
    path = str(tmp_path / "p.pt")
    checkpoint.save(path, m, LABELS)
    m2, _, _, use_roi = load_classifier(path, device="cpu")
    assert not use_roi and m2.roi_cnn is None and m2.gru.weight_ih_l0.shape == (576, 180)
    assert all(torch.equal(v, m.state_dict()[k]) for k, v in m2.state_dict().items())
