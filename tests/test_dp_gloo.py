"""Host-side data-parallel logic on CPU (gloo, world_size 2): contiguous sharding of independent
clips and the 1/B_global pre-scaling under which a SUM all-reduce of per-shard gradients equals
the full-batch gradient (the invariant silent_speech_b200.train.Trainer relies on).
The per-shard gradients here come from the CPU oracle -- the CUDA kernels are covered by -m gpu."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_model
from silent_speech_b200 import synth
from silent_speech_b200.train import shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 128, 1000):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    p = ref_model.synth_params(3, use_roi=False)
    B = 6
    d = synth.clip_batch(B, 10, mode="uniform", seed=4, use_roi=False)
    lo, hi = shard_range(B, rank, world)
    X, L, y = (torch.from_numpy(d[k][lo:hi]) for k in ("X", "lengths", "y"))
    leaves = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    logits = ref_model.forward(leaves, X, L, None)
    per_clip = torch.nn.functional.cross_entropy(logits, y, label_smoothing=0.05, reduction="none")
    (per_clip.sum() / B).backward()                       # local shard, pre-scaled by 1/B_global
    flat = torch.cat([v.grad.reshape(-1) for v in leaves.values()])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.save(out, flat.numpy())
    dist.destroy_process_group()


def test_allreduced_shard_grads_equal_full_batch(tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "flat.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    p = ref_model.synth_params(3, use_roi=False)
    d = synth.clip_batch(6, 10, mode="uniform", seed=4, use_roi=False)
    _, _, grads = ref_model.loss_and_grads(p, torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"]), None,
                                           torch.from_numpy(d["y"]))
    ref = torch.cat([g.reshape(-1) for g in grads.values()]).numpy()
    assert np.abs(got - ref).max() <= 1e-6 + 1e-5 * np.abs(ref).max()
