"""Logit error of the tc3 inference mode as a function of which stages round their ACTIVATIONS to fp16 once instead of
splitting them (weights are always split): the two GRU input projections (ssb_proj_gemm_wsplit) and the ROI CNN (K2 passes=2).
    python tests/probes/proj_probe.py        (B200; the oracle is the checker)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref_model
from silent_speech_b200 import BiGRUClassifier, synth
torch.set_num_threads(os.cpu_count() or 1)
dev = torch.device("cuda")
def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
for use_roi, B, T, seed in ((True, 150, 90, 7), (False, 260, 64, 7), (True, 64, 90, 42)):
    p = ref_model.synth_params(seed, use_roi=use_roi)
    d = synth.clip_batch(B, T, mode="uniform", seed=9, use_roi=use_roi)
    X, L = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"])
    R = torch.from_numpy(d["R"]) if use_roi else None
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    m = BiGRUClassifier(180, 10, use_roi=use_roi, precision="tc3")
    m.load_state_dict(p); m = m.to(dev).eval()
    print("X absmax", float(X.abs().max()), "X rms", float(X.pow(2).mean().sqrt()))
    for full in (True, (0,), (1,), False):      # which projection layers keep the activation split
        for cnn_full in (False, True):
            m.proj_full_split = full; m.cnn_full_split = cnn_full
            with torch.no_grad():
                lg = m(X.to(dev), L.to(dev), None if R is None else R.to(dev))
            print(f"roi={use_roi} seed={seed} proj_full={full} cnn_full={cnn_full}: rel_err {rel_err(lg, ref):.3e}")

# the ROI CNN alone in single-pass fp16 (weights rounded too) under split projections / recurrences ("cnn16")
for use_roi, B, T, seed in ((True, 150, 90, 7), (True, 64, 90, 42), (True, 96, 90, 3)):
    p = ref_model.synth_params(seed, use_roi=use_roi)
    d = synth.clip_batch(B, T, mode="uniform", seed=9, use_roi=use_roi)
    X, L, R = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"]), torch.from_numpy(d["R"])
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
        emb_ref = ref_model.roi_cnn(p, ref_model.roi_prepare(R, True))
    for prec in ("tc3", "cnn16", "fp16"):
        m = BiGRUClassifier(180, 10, use_roi=True, precision=prec)
        m.load_state_dict(p); m = m.to(dev).eval()
        with torch.no_grad():
            lg = m(X.to(dev), L.to(dev), R.to(dev))
        print(f"seed={seed} precision={prec}: logit rel_err {rel_err(lg, ref):.3e}  (|emb| max {float(emb_ref.abs().max()):.2f}, |logit| max {float(ref.abs().max()):.2f})")
