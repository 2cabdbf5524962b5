"""Drop-in ``BiGRUClassifier`` whose forward/backward run in libssb200's CUDA kernels.

Same constructor, call signature and ``state_dict()`` layout as the reference class
(train_model_official.py:253-310; live_infer_official.py:101-138 adds ``gru_layers``),
so ``word_model_points_roi.pt`` checkpoints load into it unchanged and
``model(X, lengths, R)`` / ``loss.backward()`` / ``model.parameters()`` keep working in
``train_model_official.py``'s loop.  The sub-modules below are parameter containers
only: they never compute anything, every FLOP happens behind the C ABI.

Two forward semantics share one set of weights in the reference:
``standardize_roi=True``  per-frame ROI standardisation, train_model_official.py:289-291 (default)
``standardize_roi=False`` no standardisation,            live_infer_official.py:126-127
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import functional as F_ssb

HIDDEN = 192
ROI_EMB = 32


class _Bag(nn.Module):
    """A named bundle of parameters (no forward)."""

    def __init__(self, **shapes):
        super().__init__()
        for name, shp in shapes.items():
            self.register_parameter(name, nn.Parameter(torch.empty(*shp)))


def _uniform_(t: torch.Tensor, bound: float):
    with torch.no_grad():
        t.uniform_(-bound, bound)


def _init_affine(bag: _Bag, fan_in: int):
    """nn.Conv2d / nn.Linear default init: kaiming_uniform_(a=sqrt(5)) on the weight, then
    U(-1/sqrt(fan_in), 1/sqrt(fan_in)) on the bias -- reproduced call for call so that the
    same torch seed yields bit-identical tensors to the reference's constructor."""
    gain = math.sqrt(2.0 / (1 + math.sqrt(5) ** 2))
    std = gain / math.sqrt(fan_in)
    _uniform_(bag.weight, math.sqrt(3.0) * std)
    _uniform_(bag.bias, 1 / math.sqrt(fan_in))


class _RoiCnnParams(nn.Module):
    """Parameters named like TinyROICNN's (train_model_official.py:209-221):
    net.0 / net.3 / net.6 convolutions and fc."""

    def __init__(self, out_dim: int):
        super().__init__()
        self.net = nn.Module()
        for idx, (co, ci) in (("0", (8, 1)), ("3", (16, 8)), ("6", (24, 16))):
            bag = _Bag(weight=(co, ci, 3, 3), bias=(co,))
            _init_affine(bag, ci * 9)
            self.net.add_module(idx, bag)
        self.fc = _Bag(weight=(out_dim, 24), bias=(out_dim,))
        _init_affine(self.fc, 24)

    def tensors(self):
        n = self.net
        return [getattr(n, "0").weight, getattr(n, "0").bias, getattr(n, "3").weight, getattr(n, "3").bias,
                getattr(n, "6").weight, getattr(n, "6").bias, self.fc.weight, self.fc.bias]


class _GruParams(nn.Module):
    """Parameters named like nn.GRU's (weight_ih_l0, ..., bias_hh_l1_reverse), gate rows (r,z,n)."""

    def __init__(self, in_dim: int, hidden: int, layers: int):
        super().__init__()
        self.hidden, self.layers = hidden, layers
        stdv = 1.0 / math.sqrt(hidden)
        for layer in range(layers):
            width = in_dim if layer == 0 else 2 * hidden
            for sfx in ("", "_reverse"):
                for name, shp in ((f"weight_ih_l{layer}{sfx}", (3 * hidden, width)),
                                  (f"weight_hh_l{layer}{sfx}", (3 * hidden, hidden)),
                                  (f"bias_ih_l{layer}{sfx}", (3 * hidden,)),
                                  (f"bias_hh_l{layer}{sfx}", (3 * hidden,))):
                    self.register_parameter(name, nn.Parameter(torch.empty(*shp)))
        for p in self.parameters():          # nn.GRU.reset_parameters order == registration order
            _uniform_(p, stdv)

    def layer(self, layer: int):
        g = lambda n: getattr(self, n)
        return {k: (g(f"{k}_l{layer}"), g(f"{k}_l{layer}_reverse"))
                for k in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")}


class _PoolParams(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.score = _Bag(weight=(1, dim), bias=(1,))
        _init_affine(self.score, dim)


class BiGRUClassifier(nn.Module):
    def __init__(self, x_dim: int, num_classes: int, use_roi: bool = False, roi_emb: int = ROI_EMB,
                 hidden: int = HIDDEN, gru_layers: int = 2, standardize_roi: bool = True,
                 precision: str = "tc3"):
        super().__init__()
        if precision not in ("fp32", "tc3", "fp16", "cnn16", "gru16"):
            raise ValueError("precision must be 'fp32' (CUDA-core FMA), 'tc3' (tensor cores, split-fp16, "
                             "fp32-parity grade), 'fp16' (tensor cores, single fp16 pass) or one of the diagnostic "
                             "mixes 'cnn16' / 'gru16' (only that stage single-pass; tests/probes/precision_modes.py)")
        self.precision = precision
        if hidden != HIDDEN or roi_emb != ROI_EMB or gru_layers != 2:
            raise ValueError("libssb200 kernels are specialised for hidden=192, roi_emb=32, 2 GRU layers "
                             "(train_model_official.py:402, live_infer_official.py:209-217)")
        self.x_dim, self.num_classes = int(x_dim), int(num_classes)
        self.use_roi = bool(use_roi)
        self.standardize_roi = bool(standardize_roi)
        self.gru_dropout, self.head_dropout = 0.1, 0.2       # train:266, train:275
        # construction order == the reference's, so a shared torch seed gives identical weights
        self.roi_cnn = _RoiCnnParams(roi_emb) if use_roi else None
        self.gru = _GruParams(x_dim + (roi_emb if use_roi else 0), hidden, gru_layers)
        self.pool = _PoolParams(2 * hidden)
        self.head = nn.Module()
        ln = _Bag(weight=(2 * hidden,), bias=(2 * hidden,))
        with torch.no_grad():
            ln.weight.fill_(1.0); ln.bias.zero_()
        self.head.add_module("0", ln)
        fc1 = _Bag(weight=(128, 2 * hidden), bias=(128,)); _init_affine(fc1, 2 * hidden)
        self.head.add_module("1", fc1)
        fc2 = _Bag(weight=(num_classes, 128), bias=(num_classes,)); _init_affine(fc2, 128)
        self.head.add_module("4", fc2)
        # dropout-mask counter: follows torch.manual_seed like the reference's nn.Dropout / nn.GRU masks do; Trainer mixes
        # in the data-parallel rank (every rank must draw its own masks) and checkpoints it (Trainer.state_dict)
        self._seed = (0x5EED ^ (torch.initial_seed() * 0x9E3779B97F4A7C15)) & 0xFFFFFFFFFFFFFFFF

    # -- parameter groups in the order the C ABI takes them
    def head_tensors(self):
        h = self.head
        return [self.pool.score.weight, self.pool.score.bias, getattr(h, "0").weight, getattr(h, "0").bias,
                getattr(h, "1").weight, getattr(h, "1").bias, getattr(h, "4").weight, getattr(h, "4").bias]

    def _next_seed(self) -> int:
        self._seed = (self._seed * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
        return self._seed

    def forward(self, X: torch.Tensor, lengths: torch.Tensor, R: Optional[torch.Tensor] = None) -> torch.Tensor:
        """X f32 [B,T,x_dim], lengths i64 [B], R u8 [B,T,48,96] or None -> logits f32 [B,C]
        (train_model_official.py:279-310)."""
        return F_ssb.classifier_forward(self, X, lengths, R if self.use_roi else None)

    @torch.no_grad()
    def infer_host(self, X: torch.Tensor, lengths: torch.Tensor, R: Optional[torch.Tensor] = None,
                   chunk: int = 512, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``model(X, lengths, R)`` for HOST tensors (what collate_fn, train_model_official.py:174-204, yields):
        the batch is cut into chunks of at most ``chunk`` clips (clips are independent; the first and last two chunks are
        smaller, see the schedule below), each chunk's host->device copy
        runs on a copy stream while the previous chunk computes, and the logits come back into a host tensor.
        Pinned inputs give true overlap; pageable ones still work (the copies then serialise).  Returns f32 [B,C]
        on the host.  The PCIe copy of R (414,720 B/clip) is the floor of this call."""
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise F_ssb.ops._lib.SsbError("infer_host needs the model on a CUDA device (no CPU path)")
        with torch.cuda.device(dev):
            return self._infer_host(dev, X, lengths, R, chunk, out)

    def _infer_host(self, dev, X, lengths, R, chunk, out):
        B = X.shape[0]
        use_roi = self.use_roi and R is not None
        if out is None:
            out = torch.empty(B, self.num_classes, dtype=torch.float32).pin_memory()
        main = torch.cuda.current_stream(dev)
        st = self.__dict__.setdefault("_h2d_state", {})
        if st.get("dev") != dev:
            st.clear()
            st.update(dev=dev, copy=torch.cuda.Stream(dev), bufs={})
        copy = st["copy"]
        key = (chunk, tuple(X.shape[1:]), tuple(R.shape[1:]) if use_roi else None)
        if key not in st["bufs"]:
            st["bufs"].clear()
            st["bufs"][key] = [dict(X=torch.empty((chunk,) + tuple(X.shape[1:]), device=dev, dtype=torch.float32),
                                    L=torch.empty(chunk, device=dev, dtype=torch.int64),
                                    R=torch.empty((chunk,) + tuple(R.shape[1:]), device=dev, dtype=torch.uint8) if use_roi else None,
                                    ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(2)]
        bufs = st["bufs"][key]
        was_training = self.training
        self.eval()
        copy.wait_stream(main)
        # chunk schedule: the copy engine is the bottleneck, so the first chunk (whose copy nothing overlaps) and the last one
        # (whose compute nothing overlaps) are small -- chunk/4, chunk/2, chunk, ..., chunk, chunk/2, chunk/4
        sizes, left = [], B
        head = [max(1, chunk // 4), max(1, chunk // 2)] if B >= 4 * chunk else []
        for h in head:
            sizes.append(h); left -= h
        tail = list(reversed(head))
        left -= sum(tail)
        while left > 0:
            sizes.append(min(chunk, left)); left -= sizes[-1]
        sizes += tail
        lo = 0
        for i, n in enumerate(sizes):
            hi = lo + n
            b = bufs[i & 1]
            with torch.cuda.stream(copy):
                if i >= 2:
                    copy.wait_event(b["free"])             # the compute that read this buffer two chunks ago is done
                b["X"][:n].copy_(X[lo:hi], non_blocking=True)
                b["L"][:n].copy_(lengths[lo:hi], non_blocking=True)
                if use_roi:
                    b["R"][:n].copy_(R[lo:hi], non_blocking=True)
                b["ready"].record(copy)
            main.wait_event(b["ready"])
            logits = F_ssb.classifier_forward(self, b["X"][:n], b["L"][:n], b["R"][:n] if use_roi else None)
            out[lo:hi].copy_(logits, non_blocking=True)
            b["free"].record(main)
            lo = hi
        if was_training:
            self.train()
        main.synchronize()
        return out

    @torch.no_grad()
    def predict_topk(self, X, lengths, R=None, k: int = 3):
        """softmax + top-k fused into the head kernel (live_infer_official.py:223-226)."""
        return F_ssb.classifier_forward(self, X, lengths, R if self.use_roi else None, k_top=k)


def load_classifier(path: str, device="cuda", standardize_roi: bool = False):
    """Reads a checkpoint written by train_model_official.py:489-500 (or by
    silent_speech_b200.checkpoint.save) the way live_infer_official.py:198-221 does."""
    ckpt = torch.load(path, map_location="cpu")
    model = BiGRUClassifier(int(ckpt["x_dim"]), len(ckpt["labels"]), use_roi=bool(ckpt.get("use_roi", False)),
                            gru_layers=int(ckpt.get("gru_layers", 2)), standardize_roi=standardize_roi)
    model.load_state_dict(ckpt["model"])
    model.eval()
    return model.to(device), ckpt["id_to_label"], int(ckpt["max_t"]), bool(ckpt.get("use_roi", False))
