"""Small-shape launches of the shared-memory / cluster kernels for compute-sanitizer (racecheck, memcheck):

    compute-sanitizer --tool racecheck python tools/sanitize_small.py

The cluster recurrence (DSMEM exchange + mbarriers), the BPTT cluster kernel, K2 (named barriers, TMA frame ring, TMEM),
the two persistent projection GEMMs.  Checks each result against the fp32 CUDA-core kernels so that a sanitizer-induced
slow-down cannot hide a wrong answer."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from silent_speech_b200 import ops, synth

dev = torch.device("cuda")
p = {k: v.to(dev).contiguous() for k, v in synth.synth_params(3).items()}
B, T, H, G3 = 6, 5, 192, 576
L = torch.tensor([5, 3, 1, 5, 2, 4], device=dev)


def rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


# K2: both tensor-core instances against the fp32 kernel
d = synth.clip_batch(B, T, mode="uniform", seed=2)
R = torch.from_numpy(d["R"]).to(dev)
names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
         "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
w = [p[n] for n in names]
ref = ops.roi_cnn_fwd(R, L, w, passes=0)
for passes, tol in ((3, 1e-4), (1, 3e-3)):
    e = rel(ops.roi_cnn_fwd(R, L, w, passes=passes), ref)
    print(f"K2 passes={passes}: rel err {e:.2e}")
    assert e < tol

# projection GEMMs
A = torch.randn(B * T, 384, device=dev)
Wa, Wb = p["gru.weight_ih_l1"], p["gru.weight_ih_l1_reverse"]
bias = torch.randn(2 * G3, device=dev)
want = A.double() @ torch.cat([Wa, Wb]).double().T + bias.double()
gi = torch.empty(B * T, 2 * G3, device=dev)
ops.proj_gemm(A, ops.proj_prepack(Wa, Wb, 384, 0, split=True), bias, gi, 2 * G3, split=True)
print(f"proj split: rel err {rel(gi.double(), want):.2e}")
assert rel(gi.double(), want) < 2e-6
gi1 = torch.empty_like(gi)
ops.proj_gemm(A, ops.proj_prepack(Wa, Wb, 384, 0), bias, gi1, 2 * G3)
assert rel(gi1.double(), want) < 3e-3

# cluster recurrence (forward) and the BPTT kernels
gi = (torch.randn(B, T, 2 * G3, device=dev) * 0.5).contiguous()
wh = [p["gru.weight_hh_l1"], p["gru.bias_hh_l1"], p["gru.weight_hh_l1_reverse"], p["gru.bias_hh_l1_reverse"]]
zero = torch.zeros(G3, device=dev)
ref = ops.gru_layer_fwd(gi, L, wh[0], wh[1], wh[2], wh[3])
for passes, tol in ((3, 1e-5), (1, 3e-3)):
    gates = torch.empty(B, T, 2, 4, H, device=dev) if passes == 3 else None
    out = ops.gru_layer_fwd_tc(gi + torch.cat([torch.cat([wh[1][:2 * H], zero[:H]]), torch.cat([wh[3][:2 * H], zero[:H]])]),
                               L, wh[0], wh[1], wh[2], wh[3], passes=passes, gates=gates)
    print(f"recurrence passes={passes}: rel err {rel(out, ref):.2e}")
    assert rel(out, ref) < tol
gates = torch.empty(B, T, 2, 4, H, device=dev)
out = ops.gru_layer_fwd(gi, L, wh[0], wh[1], wh[2], wh[3], gates=gates)
dgi, dghn, hprev = ops.gru_layer_bwd(torch.randn_like(out), out, gates, ops.transpose(wh[0]), ops.transpose(wh[2]), L)
assert bool(torch.isfinite(dgi).all())
torch.cuda.synchronize()
print("sanitize_small: ok")
