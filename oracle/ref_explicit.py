"""Oracle (TEST INFRASTRUCTURE ONLY): explicit-math restatement in any dtype.

No ``conv2d`` / ``nn.GRU`` / ``layer_norm`` library ops: convolutions are
written as shifted multiply-adds, the GRU as a per-step loop with the gate
equations of SURVEY.md section 3.2-5, the pooling and head by hand.  Run in
float64 it is the "truth" that bounds how far fp32 implementations (ATen's and
ours) may sit from the real-number answer; it also exposes every intermediate
(input projections, per-layer sequences, gates) so a failing CUDA stage can be
localised in one GPU run.

Follows train_model_official.py:279-310 (forward), 212-229 (CNN), 240-248
(pool), 271-277 (head); gate order (r, z, n) and packed-sequence semantics as
in torch.nn.GRU (reverse direction starts at t = len-1, h0 = 0, zeros past len).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


def _conv3x3(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """x [N,Ci,H,W], w [Co,Ci,3,3] -> [N,Co,H,W]; zero padding 1, cross-correlation."""
    N, Ci, H, W = x.shape
    xp = torch.zeros(N, Ci, H + 2, W + 2, dtype=x.dtype)
    xp[:, :, 1:-1, 1:-1] = x
    out = b.view(1, -1, 1, 1).expand(N, w.shape[0], H, W).clone()
    for dy in range(3):
        for dx in range(3):
            patch = xp[:, :, dy:dy + H, dx:dx + W]                      # [N,Ci,H,W]
            out += torch.einsum("nchw,oc->nohw", patch, w[:, :, dy, dx])
    return out


def _maxpool2(x: torch.Tensor) -> torch.Tensor:
    N, C, H, W = x.shape
    return x.reshape(N, C, H // 2, 2, W // 2, 2).amax(dim=(3, 5))


def roi_embed(p: Dict[str, torch.Tensor], R: torch.Tensor, standardize: bool, dtype) -> torch.Tensor:
    B, T, H, W = R.shape
    r = R.to(dtype) / 255.0
    r = r.reshape(B * T, 1, H, W)
    if standardize:
        n = H * W
        mu = r.sum(dim=(1, 2, 3), keepdim=True) / n
        var = ((r - mu) ** 2).sum(dim=(1, 2, 3), keepdim=True) / (n - 1)
        r = (r - mu) / var.sqrt().clamp_min(1e-6)
    g = lambda k: p[k].to(dtype)
    a = _maxpool2(_conv3x3(r, g("roi_cnn.net.0.weight"), g("roi_cnn.net.0.bias")).clamp_min(0))
    a = _maxpool2(_conv3x3(a, g("roi_cnn.net.3.weight"), g("roi_cnn.net.3.bias")).clamp_min(0))
    a = _conv3x3(a, g("roi_cnn.net.6.weight"), g("roi_cnn.net.6.bias")).clamp_min(0)
    m = a.mean(dim=(2, 3))
    e = m @ g("roi_cnn.fc.weight").T + g("roi_cnn.fc.bias")
    return e.reshape(B, T, -1)


def gru_direction(gi: torch.Tensor, w_hh: torch.Tensor, b_hh: torch.Tensor,
                  lengths: torch.Tensor, reverse: bool):
    """gi [B,T,3H] (input projection incl. b_ih) -> h sequence [B,T,H], zeros past len."""
    B, T, H3 = gi.shape
    H = H3 // 3
    out = torch.zeros(B, T, H, dtype=gi.dtype)
    for b in range(B):
        n = int(lengths[b])
        h = torch.zeros(H, dtype=gi.dtype)
        for s in range(n):
            t = n - 1 - s if reverse else s
            gh = w_hh @ h + b_hh
            r = torch.sigmoid(gi[b, t, :H] + gh[:H])
            z = torch.sigmoid(gi[b, t, H:2 * H] + gh[H:2 * H])
            c = torch.tanh(gi[b, t, 2 * H:] + r * gh[2 * H:])
            h = (1 - z) * c + z * h
            out[b, t] = h
    return out


def bigru(p: Dict[str, torch.Tensor], Z: torch.Tensor, lengths: torch.Tensor, dtype):
    seq = Z.to(dtype)
    inter = {}
    for layer in (0, 1):
        halves = []
        for sfx, rev in (("", False), ("_reverse", True)):
            g = lambda k: p[f"gru.{k}_l{layer}{sfx}"].to(dtype)
            gi = seq @ g("weight_ih").T + g("bias_ih")
            inter[f"gi_l{layer}{sfx}"] = gi
            halves.append(gru_direction(gi, g("weight_hh"), g("bias_hh"), lengths, rev))
        seq = torch.cat(halves, dim=2)
        inter[f"seq_l{layer}"] = seq
    return seq, inter


def pool_and_head(p: Dict[str, torch.Tensor], seq: torch.Tensor, lengths: torch.Tensor, dtype):
    g = lambda k: p[k].to(dtype)
    B, T, D = seq.shape
    pooled = torch.zeros(B, D, dtype=dtype)
    for b in range(B):
        n = int(lengths[b])
        s = seq[b, :n] @ g("pool.score.weight")[0] + g("pool.score.bias")[0]
        a = torch.softmax(s, dim=0)
        pooled[b] = (a.unsqueeze(1) * seq[b, :n]).sum(0)
    mu = pooled.mean(dim=1, keepdim=True)
    var = ((pooled - mu) ** 2).mean(dim=1, keepdim=True)               # biased, eps 1e-5
    ln = (pooled - mu) / (var + 1e-5).sqrt() * g("head.0.weight") + g("head.0.bias")
    hid = (ln @ g("head.1.weight").T + g("head.1.bias")).clamp_min(0)
    logits = hid @ g("head.4.weight").T + g("head.4.bias")
    return pooled, logits


def forward(p: Dict[str, torch.Tensor], X: torch.Tensor, lengths: torch.Tensor,
            R: Optional[torch.Tensor] = None, standardize_roi: bool = True,
            dtype=torch.float64):
    """Returns a dict of every stage: roi_emb, gi_l*, seq_l0, seq_l1 (= gru_out),
    pooled, logits.  Eval semantics (no dropout)."""
    out = {}
    Z = X.to(dtype)
    if R is not None and "roi_cnn.fc.weight" in p:
        out["roi_emb"] = roi_embed(p, R, standardize_roi, dtype)
        Z = torch.cat([Z, out["roi_emb"]], dim=2)
    seq, inter = bigru(p, Z, lengths, dtype)
    out.update(inter)
    out["gru_out"] = seq
    out["pooled"], out["logits"] = pool_and_head(p, seq, lengths, dtype)
    return out
