""".npz ingestion / split / sampler against the reference's own split_by_label, NPZWordDataset +
collate_fn and WeightedRandomSampler, through the fixture written by oracle/make_golden.py::golden_dataset."""
import os

import numpy as np
import pytest
import torch

from silent_speech_b200 import dataset as D

GOLD = os.path.join(os.path.dirname(__file__), "golden", "dataset.npz")


@pytest.fixture()
def clip_dir(tmp_path):
    g = np.load(GOLD)
    for i, name in enumerate(g["names"]):
        X = g[f"X_{i}"]
        T, Tr = X.shape[0], int(g[f"roi_tr_{i}"])
        kw = dict(X=X, ts=np.arange(T, dtype=np.int32), label=str(g[f"label_{i}"]), speaker="s0", idxs=np.arange(88, dtype=np.int32))
        if Tr >= 0:          # the generator's pattern (oracle/make_golden.py::golden_dataset)
            kw["roi"] = ((np.arange(Tr)[:, None, None] * 5 + i * 11 + np.arange(48)[None, :, None] + 2 * np.arange(96)[None, None, :]) % 256).astype(np.uint8)
        np.savez(tmp_path / str(name), **kw)
    return str(tmp_path), g


def test_scan_split_collate_match_reference(clip_dir):
    path, g = clip_dir
    idx = D.scan_clips(path)
    assert [os.path.basename(f) for f in idx.files] == list(g["kept"])            # the 120-dim clip is dropped
    assert idx.x_dim == 180 and idx.use_roi and idx.idx_signatures == 1
    assert sorted(idx.label_to_id, key=idx.label_to_id.get) == list(g["classes"])
    tr, va = D.split_by_label(list(idx.files), list(idx.labels), 0.15, seed=42)
    assert [os.path.basename(f) for f in tr] == list(g["train"])
    assert [os.path.basename(f) for f in va] == list(g["val"])
    X, T, R, y = D.collate([D.load_clip(f, idx.label_to_id, 90, True) for f in idx.files])
    assert X.dtype == np.float32 and R.dtype == np.uint8 and T.dtype == np.int64
    np.testing.assert_array_equal(T, g["T"])
    np.testing.assert_array_equal(y, g["y"])
    np.testing.assert_array_equal(X, g["X"])
    np.testing.assert_array_equal(R, g["R"])          # incl. the ROI-less clip's zero frames and the Tr < T cut


def test_weighted_sampler_matches_torch(clip_dir):
    path, g = clip_dir
    idx = D.scan_clips(path)
    tr, _ = D.split_by_label(list(idx.files), list(idx.labels), 0.15, seed=42)
    lab = {f: l for f, l in zip(idx.files, idx.labels)}
    w = D.sampler_weights([lab[f] for f in tr])
    np.testing.assert_array_equal(w.numpy(), g["weights"])
    gen = torch.Generator(); gen.manual_seed(123)
    np.testing.assert_array_equal(D.weighted_indices(w, generator=gen).numpy(), g["order"])


def test_clip_pad_trim_edges():
    X = np.arange(12, dtype=np.float32).reshape(4, 3)
    out, T = D.clip_pad_trim(X, 4, 6)
    assert T == 4 and out.shape == (6, 3) and (out[4:] == 0).all() and (out[:4] == X).all()
    out, T = D.clip_pad_trim(X, 4, 4)
    assert T == 4 and out.shape == (4, 3)
    out, T = D.clip_pad_trim(X, 4, 2)
    assert T == 2 and (out == X[:2]).all()
    out, T = D.clip_pad_trim(X[:0], 0, 3)
    assert T == 0 and out.shape == (3, 3) and (out == 0).all()


def test_store_refuses_cpu():
    with pytest.raises(Exception):
        D.ClipStore(np.zeros((1, 4, 3), np.float32), np.ones(1, np.int64), None, np.zeros(1, np.int64), device="cpu")


@pytest.mark.gpu
def test_clip_store_batches_match_collate(clip_dir):
    path, g = clip_dir
    idx = D.scan_clips(path)
    store = D.ClipStore.from_files(idx.files, idx.label_to_id, 90, True, device="cuda")
    assert store.N == len(idx.files) and store.nbytes() == store.N * (90 * 180 * 4 + 90 * 4608 + 16)
    pick = [3, 0, 12, 12, 7, 1]
    X, L, R, y = store.batch(pick)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(X.cpu().numpy(), g["X"][pick])
    np.testing.assert_array_equal(R.cpu().numpy(), g["R"][pick])
    np.testing.assert_array_equal(L.cpu().numpy(), g["T"][pick])
    np.testing.assert_array_equal(y.cpu().numpy(), g["y"][pick])
    with pytest.raises(IndexError):
        store.batch([0, store.N])
    # stored order, ragged last batch
    got = list(store.epoch(5, weighted=False))
    assert [b[0].shape[0] for b in got] == [5, 5, 3]
    np.testing.assert_array_equal(torch.cat([b[1] for b in got]).cpu().numpy(), g["T"])
    # weighted epoch == torch's sampler order; augmentation keeps shapes and only ever shortens clips
    gen = torch.Generator(); gen.manual_seed(5)
    order = D.weighted_indices(D.sampler_weights(store.labels), store.N, torch.Generator().manual_seed(5))
    got = list(store.epoch(4, weighted=True, generator=gen, augment_seed=77))
    ys = torch.cat([b[3] for b in got]).cpu().numpy()
    np.testing.assert_array_equal(ys, g["y"][order.numpy()])
    Ls = torch.cat([b[1] for b in got]).cpu().numpy()
    base = g["T"][order.numpy()]
    assert ((Ls <= base) & (Ls >= base - 2)).all()
    assert all(b[0].shape[1:] == (90, 180) and b[2].shape[1:] == (90, 48, 96) for b in got)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,dtype", [((37, 90, 180), torch.float32), ((9, 3), torch.uint8), ((50,), torch.int64),
                                         ((11, 7), torch.float32), ((64, 90, 48, 96), torch.uint8), ((5, 1031), torch.uint8)])
def test_gather_rows_shapes(shape, dtype):
    from silent_speech_b200 import ops
    gen = torch.Generator().manual_seed(3)
    src = torch.randint(0, 250, shape, generator=gen).to(dtype).cuda()
    idx = torch.randint(0, shape[0], (2 * shape[0] + 1,), generator=gen)
    out = ops.gather_rows(src, idx.cuda())
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), src.cpu()[idx])
    assert ops.gather_rows(src, idx[:0].cuda()).shape[0] == 0


@pytest.mark.gpu
def test_train_from_clip_store_reduces_loss(clip_dir):
    """The whole training data path on the device: ingest -> class-balanced epoch -> augmentation -> fused train step."""
    from silent_speech_b200 import BiGRUClassifier
    from silent_speech_b200.train import Trainer
    path, _ = clip_dir
    idx = D.scan_clips(path)
    store = D.ClipStore.from_files(idx.files, idx.label_to_id, 90, True, device="cuda")
    torch.manual_seed(0)
    model = BiGRUClassifier(idx.x_dim, len(idx.label_to_id), use_roi=True).cuda()
    tr = Trainer(model, lr=3e-3, dropout=False)
    gen = torch.Generator().manual_seed(1)
    losses = []
    for ep in range(12):
        for X, L, R, y in store.epoch(13, weighted=True, generator=gen, augment_seed=100 * ep):
            losses.append(float(tr.step(X, L, R, y).mean()))
    assert all(np.isfinite(losses))
    assert np.mean(losses[-3:]) < 0.7 * np.mean(losses[:3]), losses
