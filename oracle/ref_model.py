"""Oracle (TEST INFRASTRUCTURE ONLY): the clip classifier on the CPU via ATen.

A functional restatement of the reference's forward pass and train step that
goes through the *same ATen call sites* as the reference does (conv2d,
max_pool2d, packed-sequence GRU, layer_norm, linear, cross_entropy with label
smoothing, clip_grad_norm_, Adam), operating on a plain ``{name: tensor}``
dict laid out exactly like the reference's ``state_dict()``:

* ROI normalise + per-frame standardise  train_model_official.py:286-291
  (the live demo skips the standardise:  live_infer_official.py:126-127)
* TinyROICNN                             train_model_official.py:209-229
* concat, packed 2-layer BiGRU           train_model_official.py:297-305 (module 261-267)
* AttnPool                               train_model_official.py:231-248
* head                                   train_model_official.py:271-277
* loss / clip / Adam                     train_model_official.py:403-405, 433-439

Because it is the reference's own arithmetic on the reference's own backend
it doubles as the timed CPU arm of ``bench.py`` (``cpu_baseline.kind="port"``).
Pinned against the imported reference by ``oracle/make_golden.py``.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

HIDDEN = 192
ROI_EMB = 32
LABEL_SMOOTHING = 0.05      # train_model_official.py:405
CLIP_NORM = 1.0             # train_model_official.py:438
LR = 3e-4                   # train_model_official.py:31


# --------------------------------------------------------------------------
# parameter layout (the state-dict contract)
# --------------------------------------------------------------------------
def param_shapes(x_dim: int = 180, num_classes: int = 10, use_roi: bool = True,
                 roi_emb: int = ROI_EMB, hidden: int = HIDDEN) -> Dict[str, tuple]:
    """Names and shapes in the order ``state_dict()`` of the reference lists them."""
    s: Dict[str, tuple] = {}
    if use_roi:
        s["roi_cnn.net.0.weight"] = (8, 1, 3, 3);   s["roi_cnn.net.0.bias"] = (8,)
        s["roi_cnn.net.3.weight"] = (16, 8, 3, 3);  s["roi_cnn.net.3.bias"] = (16,)
        s["roi_cnn.net.6.weight"] = (24, 16, 3, 3); s["roi_cnn.net.6.bias"] = (24,)
        s["roi_cnn.fc.weight"] = (roi_emb, 24);     s["roi_cnn.fc.bias"] = (roi_emb,)
    in0 = x_dim + (roi_emb if use_roi else 0)
    for layer, width in ((0, in0), (1, 2 * hidden)):
        for sfx in ("", "_reverse"):
            s[f"gru.weight_ih_l{layer}{sfx}"] = (3 * hidden, width)
            s[f"gru.weight_hh_l{layer}{sfx}"] = (3 * hidden, hidden)
            s[f"gru.bias_ih_l{layer}{sfx}"] = (3 * hidden,)
            s[f"gru.bias_hh_l{layer}{sfx}"] = (3 * hidden,)
    s["pool.score.weight"] = (1, 2 * hidden); s["pool.score.bias"] = (1,)
    s["head.0.weight"] = (2 * hidden,);       s["head.0.bias"] = (2 * hidden,)
    s["head.1.weight"] = (128, 2 * hidden);   s["head.1.bias"] = (128,)
    s["head.4.weight"] = (num_classes, 128);  s["head.4.bias"] = (num_classes,)
    return s


def synth_params(seed: int, x_dim: int = 180, num_classes: int = 10, use_roi: bool = True,
                 scale: float = 1.0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Deterministic, machine-independent weights (numpy PCG64, not torch's RNG) with
    the reference's default-init magnitudes (U(-1/sqrt(fan_in), +1/sqrt(fan_in)),
    GRU U(-1/sqrt(H), 1/sqrt(H)), LayerNorm near (1, 0) but perturbed so its
    gradients are exercised).  ``scale`` > 1 makes the recurrence less contractive."""
    import numpy as np
    rng = np.random.default_rng(seed)
    out = {}
    for name, shp in param_shapes(x_dim, num_classes, use_roi).items():
        if name.startswith("gru."):
            b = 1.0 / math.sqrt(HIDDEN)
        elif name == "head.0.weight":
            out[name] = torch.from_numpy(1.0 + 0.1 * rng.standard_normal(shp)).to(dtype); continue
        elif name == "head.0.bias":
            out[name] = torch.from_numpy(0.1 * rng.standard_normal(shp)).to(dtype); continue
        else:
            wshape = param_shapes(x_dim, num_classes, use_roi)[name.rsplit(".", 1)[0] + ".weight"]
            fan_in = int(np.prod(wshape[1:]))
            b = 1.0 / math.sqrt(fan_in)
        out[name] = torch.from_numpy(rng.uniform(-b * scale, b * scale, size=shp)).to(dtype)
    return out


# --------------------------------------------------------------------------
# forward
# --------------------------------------------------------------------------
def roi_prepare(R: torch.Tensor, standardize: bool = True) -> torch.Tensor:
    """uint8 [B,T,H,W] -> float32 [B,T,1,H,W]; train_model_official.py:286-291."""
    r = (R.float() / 255.0).unsqueeze(2)
    if standardize:
        mu = r.mean(dim=(2, 3, 4), keepdim=True)
        sd = r.std(dim=(2, 3, 4), keepdim=True).clamp_min(1e-6)     # unbiased (N-1)
        r = (r - mu) / sd
    return r


def roi_cnn(p: Dict[str, torch.Tensor], r: torch.Tensor) -> torch.Tensor:
    """float32 [B,T,1,H,W] -> [B,T,E]; train_model_official.py:212-229."""
    B, T = r.shape[:2]
    x = r.reshape(B * T, *r.shape[2:])
    x = F.max_pool2d(F.relu(F.conv2d(x, p["roi_cnn.net.0.weight"], p["roi_cnn.net.0.bias"], padding=1)), 2)
    x = F.max_pool2d(F.relu(F.conv2d(x, p["roi_cnn.net.3.weight"], p["roi_cnn.net.3.bias"], padding=1)), 2)
    x = F.relu(F.conv2d(x, p["roi_cnn.net.6.weight"], p["roi_cnn.net.6.bias"], padding=1))
    x = F.adaptive_avg_pool2d(x, (1, 1)).reshape(B * T, -1)
    x = F.linear(x, p["roi_cnn.fc.weight"], p["roi_cnn.fc.bias"])
    return x.reshape(B, T, -1)


_GRU_ORDER = tuple(f"gru.{kind}_l{layer}{sfx}" for layer in (0, 1) for sfx in ("", "_reverse")
                   for kind in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"))


def bigru(p: Dict[str, torch.Tensor], Z: torch.Tensor, lengths: torch.Tensor,
          training: bool = False) -> torch.Tensor:
    """Packed variable-length 2-layer BiGRU; train_model_official.py:301-305.
    Calls the ATen packed-sequence GRU (the op ``nn.GRU.forward`` dispatches to)
    with the weights in ``nn.GRU``'s flat order, h0 = 0, inter-layer dropout 0.1.
    Returns [B, max(lengths), 2H] with zeros past each clip's length."""
    pk = torch.nn.utils.rnn.pack_padded_sequence(Z, lengths.cpu(), batch_first=True, enforce_sorted=False)
    H = p["gru.weight_hh_l0"].shape[1]
    h0 = torch.zeros(4, Z.shape[0], H, dtype=Z.dtype)
    flat = [p[n] for n in _GRU_ORDER]
    data, _ = torch._VF.gru(pk.data, pk.batch_sizes, h0, flat, True, 2, 0.1, training, True)
    out_pk = torch.nn.utils.rnn.PackedSequence(data, pk.batch_sizes, pk.sorted_indices, pk.unsorted_indices)
    out, _ = torch.nn.utils.rnn.pad_packed_sequence(out_pk, batch_first=True)
    return out


def attn_pool(p: Dict[str, torch.Tensor], h: torch.Tensor, lengths: torch.Tensor) -> torch.Tensor:
    """train_model_official.py:240-248."""
    T = h.shape[1]
    keep = torch.arange(T).unsqueeze(0) < lengths.unsqueeze(1)
    s = F.linear(h, p["pool.score.weight"], p["pool.score.bias"]).squeeze(-1)
    s = s.masked_fill(~keep, -1e9)
    a = torch.softmax(s, dim=1).unsqueeze(-1)
    return (h * a).sum(dim=1)


def head(p: Dict[str, torch.Tensor], pooled: torch.Tensor, training: bool = False) -> torch.Tensor:
    """train_model_official.py:271-277."""
    x = F.layer_norm(pooled, (pooled.shape[-1],), p["head.0.weight"], p["head.0.bias"], 1e-5)
    x = F.relu(F.linear(x, p["head.1.weight"], p["head.1.bias"]))
    x = F.dropout(x, 0.2, training)
    return F.linear(x, p["head.4.weight"], p["head.4.bias"])


def forward(p: Dict[str, torch.Tensor], X: torch.Tensor, lengths: torch.Tensor,
            R: Optional[torch.Tensor] = None, standardize_roi: bool = True,
            training: bool = False, want_intermediates: bool = False):
    """``BiGRUClassifier.forward`` (train_model_official.py:279-310).

    standardize_roi=True follows train_model_official.py, False follows
    live_infer_official.py:124-138 (same weights, no per-frame standardise)."""
    use_roi = "roi_cnn.fc.weight" in p and R is not None
    if use_roi:
        emb = roi_cnn(p, roi_prepare(R, standardize_roi))
        Z = torch.cat([X, emb], dim=2)
    else:
        emb, Z = None, X
    seq = bigru(p, Z, lengths, training)
    pooled = attn_pool(p, seq, lengths)
    logits = head(p, pooled, training)
    if want_intermediates:
        return {"roi_emb": emb, "gru_out": seq, "pooled": pooled, "logits": logits}
    return logits


def topk(logits: torch.Tensor, k: int = 3):
    """live_infer_official.py:223-226 -- softmax then descending argsort."""
    probs = torch.softmax(logits, dim=-1)
    order = torch.argsort(probs, dim=-1, descending=True)[..., :k]
    return order, torch.gather(probs, -1, order)


# --------------------------------------------------------------------------
# loss and train step
# --------------------------------------------------------------------------
def loss_fn(logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    return F.cross_entropy(logits, y, label_smoothing=LABEL_SMOOTHING)


def loss_and_grads(p: Dict[str, torch.Tensor], X, lengths, R, y, standardize_roi=True):
    """Forward + CE(label_smoothing=0.05) + backward with dropout inactive
    (gradient parity is defined with dropout off)."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    logits = forward(leaves, X, lengths, R, standardize_roi, training=False)
    loss = loss_fn(logits, y)
    loss.backward()
    return loss.detach(), logits.detach(), {k: v.grad for k, v in leaves.items()}


class AdamState:
    """The reference's optimiser (train_model_official.py:403) with its clip
    (train_model_official.py:438) on a dict of leaf tensors."""

    def __init__(self, p: Dict[str, torch.Tensor], lr: float = LR):
        self.leaves = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
        self.opt = torch.optim.Adam(list(self.leaves.values()), lr=lr)

    def step(self, X, lengths, R, y, standardize_roi=True, training=False):
        logits = forward(self.leaves, X, lengths, R, standardize_roi, training=training)
        loss = loss_fn(logits, y)
        self.opt.zero_grad()
        loss.backward()
        gnorm = torch.nn.utils.clip_grad_norm_(list(self.leaves.values()), CLIP_NORM)
        self.opt.step()
        return loss.detach(), logits.detach(), gnorm.detach()

    def params(self) -> Dict[str, torch.Tensor]:
        return {k: v.detach() for k, v in self.leaves.items()}
