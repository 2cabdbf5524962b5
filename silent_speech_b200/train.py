"""The train step of train_model_official.py:433-439 as one fused sequence on the GPU.

    logits = model(X, lengths, R); loss = CE(label_smoothing=0.05)(logits, y)
    opt.zero_grad(); loss.backward(); clip_grad_norm_(params, 1.0); opt.step()   # Adam(lr=3e-4)

``Trainer`` re-homes the model's parameters as views of ONE flat fp32 bucket (1,190,819
elements, 4.76 MB) with matching flat gradient / Adam-moment buckets, so that
  * the backward kernels accumulate straight into the gradient bucket,
  * data parallelism is a single NCCL all-reduce of that bucket over NVLink (clips are
    independent; the loss is pre-scaled by 1/B_global so a SUM reproduces full-batch gradients),
  * the joint-norm clip + Adam update is one kernel over the bucket (every rank computes the
    same norm from the same reduced gradients -- no second collective).
The state-dict layout is untouched: ``model.state_dict()`` still yields the reference's 32 tensors.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import _lib, ops
from .autograd import backward_impl
from .functional import forward_impl

LABEL_SMOOTHING = 0.05          # train_model_official.py:405
CLIP_NORM = 1.0                 # train_model_official.py:438


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced shard [lo, hi) of n independent clips for `rank` of `world`."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class FlatBuckets:
    """Flat parameter / gradient / moment storage with per-name views."""

    def __init__(self, model: torch.nn.Module):
        named = list(model.named_parameters())
        self.names = [n for n, _ in named]
        dev = named[0][1].device
        # 4-element (16-byte) alignment of every view: the recurrence kernels cp.async 16-byte rows
        self.offsets, off = {}, 0
        for n, p in named:
            self.offsets[n] = off
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        self.param = torch.zeros(off, device=dev, dtype=torch.float32)
        self.grad = torch.zeros_like(self.param)
        self.m = torch.zeros_like(self.param)
        self.v = torch.zeros_like(self.param)
        self.grads = {}
        for n, p in named:
            o = self.offsets[n]
            view = self.param[o:o + p.numel()].view_as(p)
            view.copy_(p.data)
            p.data = view
            self.grads[n] = self.grad[o:o + p.numel()].view_as(p)

    def attach_grads(self, model):
        for n, p in model.named_parameters():
            p.grad = self.grads[n]


class Trainer:
    def __init__(self, model, lr: float = 3e-4, betas=(0.9, 0.999), eps: float = 1e-8,
                 world_size: int = 1, clip_norm: float = CLIP_NORM, label_smoothing: float = LABEL_SMOOTHING,
                 dropout: bool = True):
        if not next(model.parameters()).is_cuda:
            raise _lib.SsbError("Trainer needs the model on a CUDA device")
        self.model, self.lr, self.betas, self.eps = model, lr, betas, eps
        self.world, self.clip_norm, self.ls = world_size, clip_norm, label_smoothing
        self.dropout = dropout
        self.rank = dist.get_rank() if (world_size > 1 and dist.is_available() and dist.is_initialized()) else 0
        if self.rank:              # per-rank dropout masks: identical replicas would otherwise mask (clip i, t, j) alike
            model._seed = (model._seed + 0xD1B54A32D192ED03 * self.rank) & 0xFFFFFFFFFFFFFFFF
        self.buckets = FlatBuckets(model)
        dev = self.buckets.param.device
        self.sqnorm = torch.zeros(1, device=dev, dtype=torch.float64)
        self.grad_norm = torch.zeros(1, device=dev, dtype=torch.float32)
        self.t = 0
        self.launches_per_step = 0
        self.allreduce_events = None        # set to [] to collect one (start, end) CUDA-event pair per all-reduce

    def state_dict(self):
        """Optimiser state for resume: Adam step count and moments plus the model's dropout-mask counter, so that
        two steps == one step + save/load + one step, bit for bit (tests/test_gpu_checkpoint.py)."""
        return {"step": self.t, "exp_avg": self.buckets.m.clone(), "exp_avg_sq": self.buckets.v.clone(),
                "dropout_seed": int(self.model._seed)}

    def load_state_dict(self, sd):
        self.t = int(sd["step"]); self.buckets.m.copy_(sd["exp_avg"]); self.buckets.v.copy_(sd["exp_avg_sq"])
        if "dropout_seed" in sd:
            self.model._seed = int(sd["dropout_seed"])

    def forward_backward(self, X, lengths, R, y, global_batch: Optional[int] = None):
        """Fills the flat gradient bucket with d(mean CE over the GLOBAL batch)/d(params) of the
        local shard; returns (per-clip loss [B_local], logits)."""
        with torch.cuda.device(X.device):
            return self._forward_backward(X, lengths, R, y, global_batch)

    def _forward_backward(self, X, lengths, R, y, global_batch):
        model = self.model
        training = self.dropout and model.training
        y = y.to(device=X.device, dtype=torch.int64).contiguous()
        out, ctx = forward_impl(model, X, lengths, R if model.use_roi else None, training=training, save=True,
                                y=y, label_smoothing=self.ls)
        self.buckets.grad.zero_()
        gb = global_batch if global_batch is not None else X.shape[0] * self.world
        backward_impl(model, ctx, self.buckets.grads, y=y, label_smoothing=self.ls, grad_scale=1.0 / gb)
        return out["loss"], out["logits"]

    def step(self, X, lengths, R, y, global_batch: Optional[int] = None, augment_seed: Optional[int] = None):
        """One full train step on this rank's shard; returns the local per-clip losses.
        augment_seed: apply NPZWordDataset's augmentation (train_model_official.py:143-152: noise p=0.7, 1-2 dropped
        frames p=0.35) to the batch on the device first; R is then cut in place to the new lengths."""
        with torch.cuda.device(X.device):
            return self._step(X, lengths, R, y, global_batch, augment_seed)

    def _step(self, X, lengths, R, y, global_batch, augment_seed):
        c0 = _lib.launch_count()
        if augment_seed is not None:
            lengths = lengths.to(device=X.device, dtype=torch.int64).contiguous()
            X, lengths, _ = ops.augment_clips(X.contiguous().float(), lengths, R if self.model.use_roi else None, augment_seed)
        loss, _ = self._forward_backward(X, lengths, R, y, global_batch)
        if self.world > 1:
            ev = None
            if self.allreduce_events is not None:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
            dist.all_reduce(self.buckets.grad, op=dist.ReduceOp.SUM)       # 4.76 MB over NVLink / NVSwitch
            if ev is not None:
                ev[1].record()
                self.allreduce_events.append(ev)
        self.t += 1
        self.sqnorm.zero_()
        ops.grad_sqnorm(self.buckets.grad, self.sqnorm)
        ops.clip_adam_step(self.buckets.param, self.buckets.grad, self.buckets.m, self.buckets.v, self.sqnorm,
                           self.clip_norm, self.lr, self.betas[0], self.betas[1], self.eps, self.t,
                           write_back_grad=True, norm_out=self.grad_norm)
        self.launches_per_step = _lib.launch_count() - c0
        return loss


class HostBatchPrefetcher:
    """Double-buffered host -> device staging for the training loop (train_model_official.py:428-439 iterates a
    DataLoader of CPU batches and moves each one to the device right before its step).

    ``submit(*host_tensors)`` starts the copy of the NEXT batch from pinned host memory on a side stream;
    ``get()`` makes the compute stream wait for it and returns the device tensors.  Calling ``submit`` for batch
    i+1 before ``Trainer.step`` on batch i hides the PCIe time (61 MB per 128 clips) behind the step's kernels.
    """

    def __init__(self, device):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self._pending = None

    def submit(self, *host_tensors):
        if self._pending is not None:
            raise RuntimeError("previous batch was not consumed: call get() first")
        for t in host_tensors:
            if t is not None and not t.is_pinned():
                raise ValueError("host batches must live in pinned memory for an asynchronous copy")
        with torch.cuda.stream(self.stream):
            dev = tuple(None if t is None else t.to(self.device, non_blocking=True) for t in host_tensors)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._pending = (dev, ev)

    def get(self):
        if self._pending is None:
            raise RuntimeError("no batch in flight: call submit() first")
        dev, ev = self._pending
        self._pending = None
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in dev:
            if t is not None:
                t.record_stream(cur)         # allocated on the side stream, used on the compute stream
        return dev
