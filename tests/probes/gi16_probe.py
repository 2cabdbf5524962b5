"""What would fp16 storage of gi cost in logit error?  Rounds gi through fp16 between the projection GEMM and the recurrence."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import ref_model
from silent_speech_b200 import BiGRUClassifier, synth, ops
import silent_speech_b200.functional as F

torch.set_num_threads(os.cpu_count() or 1)
orig = ops.proj_gemm
for use_roi, B, T in ((True, 150, 90), (False, 260, 64)):
    p = ref_model.synth_params(7, use_roi=use_roi)
    d = synth.clip_batch(B, T, mode="uniform", seed=9, use_roi=use_roi)
    X, L = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"])
    R = torch.from_numpy(d["R"]) if use_roi else None
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    m = BiGRUClassifier(180, 10, use_roi=use_roi, precision="fp16")
    m.load_state_dict(p); m = m.cuda().eval()
    for mode in ("fp32 gi", "fp16 gi"):
        def patched(A1, packed, bias, C2d, N, A2=None, lengths=None, T=0):
            out = orig(A1, packed, bias, C2d, N, A2=A2, lengths=lengths, T=T)
            if mode == "fp16 gi":
                C2d.copy_(C2d.half().float())
            return out
        ops.proj_gemm = patched
        with torch.no_grad():
            lg = m(X.cuda(), L.cuda(), None if R is None else R.cuda()).cpu()
        e = float((lg - ref).abs().max() / ref.abs().max())
        print(f"use_roi={use_roi} {mode}: rel err {e:.3e} argmax equal {bool(torch.equal(lg.argmax(1), ref.argmax(1)))}")
ops.proj_gemm = orig
