"""GPU checks of the on-device augmentation (train_model_official.py:143-152).  The reference draws from CPU
generators, so parity is (a) exact given the decisions the kernel reports -- compaction, padding, ROI cut -- against
the oracle, and (b) statistical for the decisions and the noise."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_augment_exact_given_decisions_and_statistics():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    dev = torch.device("cuda")
    from oracle import ref_augment as ra
    from silent_speech_b200 import ops, synth
    B, T = 4000, 90
    d = synth.clip_batch(64, T, mode="uniform", seed=2)
    reps = (B + 63) // 64
    X = np.concatenate([d["X"]] * reps)[:B]; L = np.concatenate([d["lengths"]] * reps)[:B]
    R = np.concatenate([d["R"]] * reps)[:B].copy()
    Xd, Ld, Rd = torch.from_numpy(X).to(dev), torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
    # (a) no noise: exact against the oracle for the reported decisions
    Xo, Lo, dec = ops.augment_clips(Xd, Ld, Rd, seed=123, p_noise=0.0)
    Xo, Lo, dec, Ro = Xo.cpu().numpy(), Lo.cpu().numpy(), dec.cpu().numpy(), Rd.cpu().numpy()
    for b in range(0, B, 13):
        n = int(L[b])
        drops = [int(v) for v in dec[b, 2:] if v >= 0][:int(dec[b, 1])]
        Xw, Tw, Rw = ra.augment_clip(X[b, :n], R[b, :n], T, None, drops)
        assert Tw == int(Lo[b]) and np.array_equal(Xw, Xo[b]) and np.array_equal(Rw, Ro[b]), b
    k = dec[:, 1]
    elig = L > 12
    assert np.all(k[~elig] == 0)
    rate = (k[elig] > 0).mean()
    assert abs(rate - 0.35) < 0.03, rate                                  # DROP_FRAMES_PROB
    assert abs((k[elig & (k > 0)] == 2).mean() - 0.5) < 0.06              # randint(1, 2)
    dr = dec[k > 0]
    assert np.all(dr[:, 2] >= 1) and np.all(dr[:, 2] <= L[k > 0] - 2)
    two = dec[k == 2]
    assert np.all(two[:, 3] > two[:, 2]) and np.all(two[:, 3] <= L[k == 2] - 2)
    # (b) noise: rate 0.7, zero mean, std 0.01, only on valid rows, same decisions for the same seed
    Xd2 = torch.from_numpy(X).to(dev)
    Xn, Ln, dec2 = ops.augment_clips(Xd2, Ld, None, seed=123, p_drop=0.0)
    Xn, dec2 = Xn.cpu().numpy(), dec2.cpu().numpy()
    assert abs(dec2[:, 0].mean() - 0.7) < 0.03
    diff = Xn - X
    mask = (np.arange(T)[None, :] < L[:, None])
    on = dec2[:, 0].astype(bool)
    assert np.all(diff[~on] == 0) and np.all(diff[~mask] == 0)
    v = diff[on][mask[on]]
    assert abs(v.mean()) < 2e-5 and abs(v.std() - 0.01) < 1e-4, (v.mean(), v.std())
    assert abs(np.mean(np.abs(v) > 0.0196) - 0.05) < 0.003                # Gaussian tails
    Xn2, _, _ = ops.augment_clips(Xd2, Ld, None, seed=123, p_drop=0.0)
    assert np.array_equal(Xn2.cpu().numpy(), Xn)                          # counter-based: reproducible
