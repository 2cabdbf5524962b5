"""GPU side of the checkpoint contract: a file the reference's load_classifier accepted gives the reference's logits
when run through the CUDA path, and Trainer state survives save / load bit for bit."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LABELS = ["hello", "yes", "no", "please", "thanks", "stop", "go", "left", "right", "help"]


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


def test_our_checkpoint_through_reference_loader_gives_our_logits(golden_dir, tmp_path):
    """golden: our seed-42 file, loaded by live_infer_official.load_classifier and run by the reference's own forward
    (no ROI standardisation).  Here: the same file through silent_speech_b200.load_classifier on the GPU."""
    dev = _cuda()
    from silent_speech_b200 import BiGRUClassifier, checkpoint, load_classifier
    g = np.load(os.path.join(golden_dir, "checkpoint.npz"))
    torch.manual_seed(42)
    m = BiGRUClassifier(180, 10, use_roi=True)
    path = str(tmp_path / "ours.pt")
    checkpoint.save(path, m, LABELS, max_t=90, seed=42)
    model, id_to_label, max_t, use_roi = load_classifier(path, device=dev)
    with torch.no_grad():
        out = model(torch.from_numpy(g["X"]).to(dev), torch.from_numpy(g["lengths"]).to(dev), torch.from_numpy(g["R"]).to(dev)).cpu().numpy()
    ref = g["ref_loaded_ours_logits"]
    assert np.abs(out - ref).max() <= 1e-4 * np.abs(ref).max()
    assert np.array_equal(out.argmax(1), ref.argmax(1))


@pytest.mark.parametrize("dropout", [False, True])
def test_trainer_resume(tmp_path, dropout):
    """two steps == one step + checkpoint.save / load + Trainer.load_state_dict + one step.  The restored state
    (parameters, Adam moments, step count, dropout-mask counter) must be bit-identical to the saved one; the step that
    follows is compared at 1e-6 because the backward pass sums per-CTA partial weight gradients with floating-point
    atomics, whose order is not fixed from run to run.  With dropout on, a lost mask counter would show as O(1e-3)."""
    dev = _cuda()
    from silent_speech_b200 import BiGRUClassifier, checkpoint, synth
    from silent_speech_b200.train import Trainer
    d = synth.clip_batch(6, 14, mode="uniform", seed=3)
    X, L, R, y = (torch.from_numpy(d[k]).to(dev) for k in ("X", "lengths", "R", "y"))

    def fresh():
        m = BiGRUClassifier(180, 10, use_roi=True)
        m.load_state_dict(synth.synth_params(9))
        return m.to(dev).train()
    a = fresh(); ta = Trainer(a, dropout=dropout)
    ta.step(X, L, R, y); ta.step(X, L, R, y)
    b = fresh(); tb = Trainer(b, dropout=dropout)
    tb.step(X, L, R, y)
    path = str(tmp_path / "resume.pt")
    ck = checkpoint.save(path, b, LABELS, trainer=tb)
    assert "ssb200_optimizer" in ck
    c, ck2 = checkpoint.load(path, device=dev)
    c.train()
    tc = Trainer(c, dropout=dropout)
    tc.load_state_dict({k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in ck2["ssb200_optimizer"].items()})
    assert torch.equal(tb.buckets.param, tc.buckets.param) and torch.equal(tb.buckets.m, tc.buckets.m)
    assert torch.equal(tb.buckets.v, tc.buckets.v) and tb.t == tc.t == 1 and b._seed == c._seed
    tc.step(X, L, R, y)
    assert ta.t == tc.t == 2
    for (n, pa), (_, pc) in zip(a.named_parameters(), c.named_parameters()):
        if n == "pool.score.bias":
            continue          # its true gradient is exactly 0 (softmax shift invariance): Adam turns rounding noise into +-lr
        assert float((pa.detach() - pc.detach()).abs().max()) <= 1e-6 * max(1.0, float(pa.detach().abs().max())), n
    assert float((ta.buckets.m - tc.buckets.m).abs().max()) <= 1e-6 * float(ta.buckets.m.abs().max())
