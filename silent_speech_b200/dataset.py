""".npz ingestion, label split, class-balanced sampling and a device-resident clip store.

Host-side mirror of the data path of train_model_official.py:

    scan_clips        main(), train:316-368      file scan, most-common-dim filter, label maps, ROI decision
    split_by_label    train:52-78                per-label shuffled 15 % validation split (same RNG stream)
    load_clip         NPZWordDataset.__getitem__ train:134-172 with augment=False (clip_pad_trim / roi_pad_trim 93-117)
    collate           collate_fn train:174-204
    sampler_weights / weighted_indices           train:381-386 (WeightedRandomSampler == one torch.multinomial per epoch)

The reference re-opens every .npz and re-pads it for every item of every epoch.  ``ClipStore`` ingests the padded clips
ONCE into HBM (479,560 B per clip with ROI at max_t = 90; 10,000 clips = 4.8 GB of 180 GB) and then a batch is one
row gather per array on the device (ssb_gather_rows) followed by the on-device augmentation (ssb_augment_clips), so a
B = 1024 train step never waits for the host.

One documented deviation: the reference augments BEFORE trimming to max_t (train:143-155), so for a clip longer than
max_t a dropped frame pulls frame max_t (or max_t + 1) into the window; the store trims at ingestion, i.e. augmentation
sees at most max_t frames.  Clips of the reference's recorder are <= max_t by construction (train:112-113), where the
two orders coincide.
"""
from __future__ import annotations

import glob
import os
import random
from collections import Counter, defaultdict
from dataclasses import dataclass
from typing import Dict, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, ops

MAX_T = 90                      # train_model_official.py:34
ROI_W, ROI_H = 96, 48           # train_model_official.py:38
VAL_FRAC, SEED = 0.15, 42       # train_model_official.py:26-27


# ----------------------------------------------------------------------------- host side (numpy / python)
def clip_pad_trim(X: np.ndarray, T: int, max_t: int):
    """train_model_official.py:93-103 (also serves roi_pad_trim, 105-117: same rule on a [T,H,W] array)."""
    if T >= max_t:
        return X[:max_t], max_t
    out = np.zeros((max_t,) + X.shape[1:], dtype=X.dtype)
    out[:T] = X
    return out, T


def load_clip(path: str, label_to_id: Dict[str, int], max_t: int = MAX_T, use_roi: bool = True):
    """One un-augmented item: (X f32 [max_t,D], T_use, R u8 [max_t,48,96] or None, y)."""
    d = np.load(path, allow_pickle=True)
    X = d["X"].astype(np.float32)
    T = int(X.shape[0])
    y = int(label_to_id[str(d["label"])])
    X, T_eff = clip_pad_trim(X, T, max_t)
    if ("roi" in d.files) and use_roi:
        R = d["roi"]
        T_use = min(T_eff, int(R.shape[0]), max_t)              # align lengths conservatively (train:160-162)
        X_pad, T_use = clip_pad_trim(np.ascontiguousarray(X[:T_use]), T_use, max_t)
        R_pad, _ = clip_pad_trim(np.ascontiguousarray(R[:T_use]).astype(np.uint8, copy=False), T_use, max_t)
        return X_pad, T_use, R_pad, y
    return X, T_eff, None, y


def collate(items, roi_hw=(ROI_H, ROI_W)):
    """collate_fn: stack; an item without ROI in a batch that has some gets zero frames (train:192-199)."""
    X = np.stack([it[0] for it in items]).astype(np.float32, copy=False)
    T = np.asarray([it[1] for it in items], np.int64)
    y = np.asarray([it[3] for it in items], np.int64)
    R = None
    if any(it[2] is not None for it in items):
        R = np.stack([it[2] if it[2] is not None else np.zeros((X.shape[1],) + tuple(roi_hw), np.uint8) for it in items])
    return X, T, R, y


@dataclass
class ClipIndex:
    files: List[str]
    labels: List[str]
    x_dim: int
    use_roi: bool
    label_to_id: Dict[str, int]
    id_to_label: Dict[int, str]
    idx_signatures: int          # number of distinct landmark-index lists seen (>1: the reference warns, train:354-360)


def scan_clips(clip_dir: str, use_roi_if_present: bool = True) -> ClipIndex:
    """The scan at the top of main() (train:316-368): sorted glob, keep only the most common feature dim,
    labels -> ids in sorted order, ROI is used when at least one file carries it."""
    files = sorted(glob.glob(os.path.join(clip_dir, "*.npz")))
    if not files:
        raise RuntimeError(f"No .npz files found in {clip_dir}")
    labels, dims, sigs, has_roi = [], [], [], 0
    for f in files:
        d = np.load(f, allow_pickle=True)
        labels.append(str(d["label"]))
        dims.append(int(d["X"].shape[1]))
        has_roi += int("roi" in d.files)
        sigs.append(tuple(d["idxs"].tolist()) if "idxs" in d.files else None)
    x_dim = Counter(dims).most_common(1)[0][0]
    keep = [i for i, dm in enumerate(dims) if dm == x_dim]
    files, labels, sigs = [files[i] for i in keep], [labels[i] for i in keep], [sigs[i] for i in keep]
    uniq = sorted(set(labels))
    label_to_id = {lab: i for i, lab in enumerate(uniq)}
    return ClipIndex(files, labels, x_dim, bool(use_roi_if_present and has_roi > 0), label_to_id,
                     {i: lab for lab, i in label_to_id.items()}, len({s for s in sigs if s is not None}))


def split_by_label(files: Sequence[str], labels: Sequence[str], val_frac: float = VAL_FRAC, seed: int = SEED):
    """train:52-78 -- same random.Random(seed) call sequence, so the same files land in train / val."""
    rng = random.Random(seed)
    by_lab = defaultdict(list)
    for f, lab in zip(files, labels):
        by_lab[lab].append(f)
    train, val = [], []
    for lab, fs in by_lab.items():
        rng.shuffle(fs)
        n = len(fs)
        n_val = min(max(1, int(round(n * val_frac))), n - 1)
        val.extend(fs[:n_val])
        train.extend(fs[n_val:])
    rng.shuffle(train)
    rng.shuffle(val)
    return train, val


def sampler_weights(labels: Sequence) -> torch.Tensor:
    """1 / (count of the item's label), float64 (train:381-385)."""
    counts = Counter(labels)
    return torch.tensor([1.0 / counts[lab] for lab in labels], dtype=torch.double)


def weighted_indices(weights: torch.Tensor, num_samples: Optional[int] = None,
                     generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """One epoch of WeightedRandomSampler(weights, num_samples, replacement=True) (train:386): torch's sampler is a
    single torch.multinomial call, reproduced here so an equal generator state yields equal indices."""
    n = int(weights.shape[0]) if num_samples is None else int(num_samples)
    return torch.multinomial(weights, n, True, generator=generator)


# ----------------------------------------------------------------------------- device side
class ClipStore:
    """Padded clips resident in HBM: X f32 [N,max_t,D], lengths i64 [N], R u8 [N,max_t,48,96] (or None), y i64 [N]."""

    def __init__(self, X, lengths, R, y, device="cuda", labels: Optional[Sequence] = None):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.SsbError("ClipStore lives in device memory (this package has no CPU path)")
        to = lambda a, dt: None if a is None else torch.as_tensor(a).to(dt).contiguous().to(dev, non_blocking=True)
        self.X, self.lengths = to(X, torch.float32), to(lengths, torch.int64)
        self.R, self.y = to(R, torch.uint8), to(y, torch.int64)
        self.device, self.N = dev, int(self.X.shape[0])
        self.labels = list(labels) if labels is not None else [int(v) for v in torch.as_tensor(y).tolist()]

    @classmethod
    def from_files(cls, files: Sequence[str], label_to_id: Dict[str, int], max_t: int = MAX_T, use_roi: bool = True,
                   device="cuda") -> "ClipStore":
        """Ingest: one load_clip per file, stacked like collate_fn, copied to the device once."""
        X, T, R, y = collate([load_clip(f, label_to_id, max_t, use_roi) for f in files])
        return cls(X, T, R, y, device, labels=y.tolist())

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.X, self.lengths, self.R, self.y) if t is not None)

    def _index(self, index) -> torch.Tensor:
        idx = torch.as_tensor(index, dtype=torch.int64)
        if idx.device.type == "cpu":
            if idx.numel() and (int(idx.min()) < 0 or int(idx.max()) >= self.N):
                raise IndexError(f"clip index out of range [0, {self.N})")
            idx = idx.to(self.device, non_blocking=True)
        return idx.contiguous()

    def batch(self, index):
        """(X, lengths, R, y) of the listed clips -- what collate_fn yields, already on the device.
        index: host sequence / tensor (range-checked here) or an int64 device tensor (trusted)."""
        idx = self._index(index)
        return (ops.gather_rows(self.X, idx), ops.gather_rows(self.lengths, idx),
                ops.gather_rows(self.R, idx) if self.R is not None else None, ops.gather_rows(self.y, idx))

    def epoch(self, batch_size: int, weighted: bool = True, generator: Optional[torch.Generator] = None,
              augment_seed: Optional[int] = None, drop_last: bool = False) -> Iterator[Tuple]:
        """One pass in the reference's order: class-balanced sampling with replacement (train:381-395) or the
        stored order (the validation loader, train:397).  With augment_seed the batch goes through the on-device
        augmentation (train:143-152); batch k uses seed augment_seed + k."""
        if weighted:
            order = weighted_indices(sampler_weights(self.labels), self.N, generator)
        else:
            order = torch.arange(self.N)
        order_dev = order.to(self.device, non_blocking=True)         # drawn from [0, N) by construction
        for k, lo in enumerate(range(0, self.N, batch_size)):
            hi = min(self.N, lo + batch_size)
            if drop_last and hi - lo < batch_size:
                break
            X, L, R, y = self.batch(order_dev[lo:hi])
            if augment_seed is not None:
                X, L, _ = ops.augment_clips(X, L, R, seed=augment_seed + k)
            yield X, L, R, y
