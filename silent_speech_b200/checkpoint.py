"""Checkpoint I/O in the reference's exact format (train_model_official.py:489-500; read back by
live_infer_official.py:198-221): a torch.save'd dict with keys model, x_dim, max_t, use_roi, roi_w,
roi_h, labels, label_to_id, id_to_label, seed.  Files written here load in the reference's
``load_classifier`` and vice versa; optimiser state, if any, goes under a NEW key that the reference ignores."""
from __future__ import annotations

import torch


def save(path, model, labels, max_t=90, seed=42, roi_w=96, roi_h=48, trainer=None):
    label_to_id = {lab: i for i, lab in enumerate(labels)}
    ckpt = {"model": {k: v.detach().cpu().clone() for k, v in model.state_dict().items()},
            "x_dim": model.x_dim, "max_t": max_t, "use_roi": model.use_roi, "roi_w": roi_w, "roi_h": roi_h,
            "labels": list(labels), "label_to_id": label_to_id,
            "id_to_label": {i: lab for lab, i in label_to_id.items()}, "seed": seed}
    if trainer is not None:
        ckpt["ssb200_optimizer"] = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in trainer.state_dict().items()}
    torch.save(ckpt, path)
    return ckpt


def load(path, device="cuda", standardize_roi=True):
    from .model import BiGRUClassifier
    ckpt = torch.load(path, map_location="cpu")
    model = BiGRUClassifier(int(ckpt["x_dim"]), len(ckpt["labels"]), use_roi=bool(ckpt.get("use_roi", False)),
                            gru_layers=int(ckpt.get("gru_layers", 2)), standardize_roi=standardize_roi)
    model.load_state_dict(ckpt["model"])
    return model.to(device), ckpt
