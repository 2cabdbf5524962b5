"""Where does the single-pass (fp16) logit error come from?  Logit error of every precision mode against the CPU
oracle on two seeded batches (GPU only).  Measured: tc3 1.6e-6, fp16 3.6e-4, cnn16 (CNN single-pass, GRU split)
0.7-2.0e-4, gru16 (GRU single-pass, CNN split) 3.3e-4 -- the recurrence dominates."""
import os, sys
sys.path.insert(0, "/root/repo")
import torch
from oracle import ref_model
from silent_speech_b200 import BiGRUClassifier, synth
dev = torch.device("cuda")
torch.set_num_threads(os.cpu_count())
for seed, scale in ((7, 1.0), (11, 1.5)):
    p = ref_model.synth_params(seed, scale=scale)
    d = synth.clip_batch(192, 90, mode="uniform", seed=9 + seed)
    X, L, R = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"]), torch.from_numpy(d["R"])
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    for prec in ("tc3", "fp16", "cnn16", "gru16"):
        m = BiGRUClassifier(180, 10, use_roi=True, precision=prec); m.load_state_dict(p); m = m.to(dev).eval()
        with torch.no_grad():
            out = m(X.to(dev), L.to(dev), R.to(dev)).cpu()
        e = float((out - ref).abs().max() / ref.abs().max())
        print(f"seed {seed} scale {scale} {prec:6s} rel_err {e:.3e} argmax agree {(out.argmax(1) == ref.argmax(1)).float().mean():.4f}")
