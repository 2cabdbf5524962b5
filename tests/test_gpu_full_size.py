"""GPU parity at BASELINE.json's FULL sizes (configs[1]: 4096 clips x 90 frames per GPU; configs[2]: 128 clips per GPU),
where the CPU oracle cannot run the whole batch in seconds: size-independent properties tie the full-size launch to
small launches that ARE checked against the oracle.

  * clips are independent (train_model_official.py:279-310 has no cross-clip op; nn.GRU on a packed batch is per sequence):
    a clip's logits must not depend on which clips share its launch, tile, cluster or frame slot -- checked bit-exactly by
    replicating 64 distinct clips 64 times across the batch, and against the oracle on the 64 originals;
  * the loss is a mean over clips, so gradients are additive over a partition of the batch.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_model  # noqa: E402


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


def _model(p, dev, precision):
    from silent_speech_b200 import BiGRUClassifier
    m = BiGRUClassifier(180, 10, use_roi=True, precision=precision)
    m.load_state_dict(p)
    return m.to(dev).eval()


@pytest.mark.parametrize("precision,tol", [("fp16", 1e-3), ("tc3", 1e-4)])
def test_full_batch_4096_replicas_identical_and_base_matches_oracle(precision, tol):
    from silent_speech_b200 import synth
    dev = _cuda()
    torch.set_num_threads(os.cpu_count() or 1)
    base, reps, T = 64, 64, 90
    p = ref_model.synth_params(42)
    d = synth.clip_batch(base, T, mode="uniform", seed=42)
    d["lengths"][:8] = T                                          # full-length clips are in the mix
    X, L, R = (torch.from_numpy(d[k]) for k in ("X", "lengths", "R"))
    for t in range(base):                                         # collate_fn zero-pads beyond each clip's length
        X[t, int(L[t]):] = 0; R[t, int(L[t]):] = 0
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    m = _model(p, dev, precision)
    # replica r of clip i sits at r * 64 + i, rotated by r so that a clip meets every tile row / lane quarter / frame slot
    idx = torch.cat([(torch.arange(base) + r) % base for r in range(reps)]).to(dev)
    Xd, Ld, Rd = X.to(dev)[idx].contiguous(), L.to(dev)[idx].contiguous(), R.to(dev)[idx].contiguous()
    assert Xd.shape[0] == 4096
    with torch.no_grad():
        full = m(Xd, Ld, Rd)
        small = m(X.to(dev), L.to(dev), R.to(dev))
    back = torch.empty_like(full)
    per_clip = full.view(reps, base, -1)
    for r in range(reps):                                          # undo the rotation: row i of every replica = clip i
        back.view(reps, base, -1)[r] = per_clip[r][(torch.arange(base, device=dev) - r) % base]
    assert torch.equal(back.view(reps, base, -1), small.unsqueeze(0).expand(reps, -1, -1)), "a clip's logits depend on its batch"
    err = float((small.cpu() - ref).abs().max() / ref.abs().max())
    assert err < tol and torch.equal(small.argmax(1).cpu(), ref.argmax(1)), err


def test_train_gradients_are_additive_over_a_partition_of_128_clips():
    from silent_speech_b200 import synth
    from silent_speech_b200.train import Trainer
    dev = _cuda()
    B, T = 128, 90
    p = ref_model.synth_params(42)
    d = synth.clip_batch(B, T, mode="uniform", seed=7)
    X, L, R, y = (torch.from_numpy(d[k]).to(dev) for k in ("X", "lengths", "R", "y"))
    tr = Trainer(_model(p, dev, "tc3"), dropout=False)
    loss_full, _ = tr.forward_backward(X, L, R, y, global_batch=B)
    g_full = tr.buckets.grad.clone()
    g_sum = torch.zeros_like(g_full)
    losses = []
    for lo, hi in ((0, 40), (40, 41), (41, 128)):                  # ragged partition, one single-clip shard
        l, _ = tr.forward_backward(X[lo:hi].contiguous(), L[lo:hi].contiguous(), R[lo:hi].contiguous(), y[lo:hi].contiguous(),
                                   global_batch=B)
        g_sum += tr.buckets.grad
        losses.append(l)
    assert torch.allclose(torch.cat(losses), loss_full, rtol=0, atol=1e-6)
    scale = float(g_full.abs().max())
    assert float((g_sum - g_full).abs().max()) < 1e-5 * scale


# --------------------------------------------------------------------------------------------- edges of the domain
@pytest.mark.parametrize("precision,tol", [("tc3", 1e-4), ("fp16", 1e-3), ("fp32", 1e-4)])
@pytest.mark.parametrize("B,T,lens", [(1, 1, [1]), (3, 90, [1, 90, 2]), (5, 7, [1, 1, 1, 1, 1]), (2, 256, [256, 6])])
def test_shortest_and_longest_clips(precision, tol, B, T, lens):
    """One-frame clips (pack_padded_sequence's minimum, train_model_official.py:301), a one-clip batch, T = 1, and the
    T = 256 stress length of configs[3], against the oracle."""
    from silent_speech_b200 import synth
    dev = _cuda()
    p = ref_model.synth_params(13)
    d = synth.clip_batch(B, T, mode="full", seed=3)
    X, R = torch.from_numpy(d["X"]), torch.from_numpy(d["R"])
    L = torch.tensor(lens, dtype=torch.int64)
    for b in range(B):
        X[b, lens[b]:] = 0; R[b, lens[b]:] = 0
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    m = _model(p, dev, precision)
    with torch.no_grad():
        out = m(X.to(dev), L.to(dev), R.to(dev)).cpu()
    assert torch.isfinite(out).all()
    assert float((out - ref).abs().max() / ref.abs().max()) < tol
    assert torch.equal(out.argmax(1), ref.argmax(1))


def test_empty_batch_returns_empty_logits():
    dev = _cuda()
    m = _model(ref_model.synth_params(13), dev, "tc3")
    X = torch.zeros(0, 90, 180, device=dev)
    L = torch.zeros(0, dtype=torch.int64, device=dev)
    R = torch.zeros(0, 90, 48, 96, dtype=torch.uint8, device=dev)
    with torch.no_grad():
        out = m(X, L, R)
    assert out.shape == (0, 10)
    assert m.infer_host(X.cpu(), L.cpu(), R.cpu()).shape == (0, 10)
