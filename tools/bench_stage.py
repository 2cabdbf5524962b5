"""Time one stage of the hot path alone (CUDA events, inputs larger than L2 or L2 flushed between runs).
    python tools/bench_stage.py {cnn|proj|gru|head} [--precision fp16|tc3] [--batch 4096] [--layer 0|1] [--reps 5]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from silent_speech_b200 import ops, synth

ap = argparse.ArgumentParser()
ap.add_argument("stage")
ap.add_argument("--precision", default="fp16")
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--layer", type=int, default=1)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--lengths", default="full")
ap.add_argument("--gi-half", action="store_true", help="proj/gru: fp16 gi (GRU-blocked, double-buffered staging)")
ap.add_argument("--standard-gi", action="store_true", help="gru: feed the single-pass kernel gi in nn.GRU column order (per-thread bulk copies) instead of the GRU-blocked order")
a = ap.parse_args()
dev = torch.device("cuda")
B, T, G3, H = a.batch, 90, 576, 192
passes = {"fp16": 1, "tc2": 2, "tc3": 3}[a.precision]   # tc2: K2 only (split weights, fp16 activations)
p = {k: v.to(dev).contiguous() for k, v in synth.synth_params(42).items()}
L = torch.from_numpy(synth.lengths(B, T, a.lengths)).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sum(ts) / len(ts)


if a.stage == "cnn":
    R = torch.randint(0, 256, (B, T, 48, 96), dtype=torch.uint8, device=dev)
    names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
             "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
    w = [p[n] for n in names]
    best, mean = timeit(lambda: ops.roi_cnn_fwd(R, L, w, passes=passes))
    fl = 5_309_952 * int(L.sum())
    print(f"cnn {a.precision} B={B}: best {best:.3f} ms mean {mean:.3f} ms  {fl / best / 1e9:.1f} TFLOP/s algorithmic")
elif a.stage == "features":
    from silent_speech_b200 import extract_features
    lm_h, present_h = synth.landmarks(min(B, 64), T, seed=1)
    reps = (B + lm_h.shape[0] - 1) // lm_h.shape[0]
    lm = torch.from_numpy(lm_h).to(dev).repeat(reps, 1, 1, 1)[:B].contiguous()
    present = torch.from_numpy(present_h).to(dev).repeat(reps, 1)[:B].contiguous()
    best, mean = timeit(lambda: extract_features(lm, synth.FRAME_W, synth.FRAME_H, present=present))
    fr = B * T
    print(f"features B={B}: best {best:.3f} ms mean {mean:.3f} ms  {fr * 1424 / best / 1e6:.0f} GB/s algorithmic (1,424 B/frame), "
          f"{fr * (478 * 12 + 720 + 13) / best / 1e6:.0f} GB/s of touched rows (5,736 B landmark row + outputs)")
elif a.stage == "crop":
    from silent_speech_b200 import crop_rois
    N = min(B, 2048)
    frames = torch.randint(0, 256, (N, 720, 1280, 3), dtype=torch.uint8, device=dev)
    g = torch.Generator().manual_seed(0)
    centers = torch.stack([torch.rand(N, generator=g) * 900 + 190, torch.rand(N, generator=g) * 400 + 160], 1).to(dev)
    mw = (torch.rand(N, generator=g, dtype=torch.float64) * 90 + 60).to(dev)
    best, mean = timeit(lambda: crop_rois(frames, centers, mw, flavor="live"))
    by = float((2.4 * mw * 2.0 * mw * 3).sum()) + N * 4608
    print(f"crop N={N}: best {best:.3f} ms mean {mean:.3f} ms  {by / best / 1e6:.0f} GB/s algorithmic (3 x crop area in + 4,608 B out), {N / best * 1e3:.0f} crops/s")
elif a.stage == "proj":
    K1, K2 = (180, 32) if a.layer == 0 else (384, 0)
    A1 = torch.randn(B * T, K1, device=dev)
    A2 = torch.randn(B * T, K2, device=dev) if K2 else None
    sfx = f"_l{a.layer}"
    Wa, Wb = p["gru.weight_ih" + sfx], p["gru.weight_ih" + sfx + "_reverse"]
    bias = torch.randn(2 * G3, device=dev)
    gi = torch.empty(B * T, 2 * G3, device=dev, dtype=torch.float16 if a.gi_half else torch.float32)
    if passes == 1:
        packed = ops.proj_prepack(Wa, Wb, K1, K2)
        fn = lambda: ops.proj_gemm(A1, packed, bias, gi, 2 * G3, A2=A2, lengths=L, T=T)
    else:
        packed = ops.proj_prepack(Wa, Wb, K1, K2, split=True)
        fn = lambda: ops.proj_gemm(A1, packed, bias, gi, 2 * G3, A2=A2, lengths=L, T=T, split=True, act_split=(a.precision != 'tc2'))
    best, mean = timeit(fn)
    by = B * T * (2 * G3 + K1 + K2) * 4
    print(f"proj layer {a.layer} {a.precision} B={B}: best {best:.3f} ms mean {mean:.3f} ms  {by / best / 1e6:.0f} GB/s algorithmic (A read + C write)")
elif a.stage == "gru":
    gi = torch.randn(B, T, 2 * G3, device=dev) * 0.5
    if a.gi_half:
        gi = gi.half()
    sfx = f"_l{a.layer}"
    w = [p["gru.weight_hh" + sfx], p["gru.bias_hh" + sfx], p["gru.weight_hh" + sfx + "_reverse"], p["gru.bias_hh" + sfx + "_reverse"]]
    order = torch.argsort(L, descending=True).to(torch.int32)
    out = torch.empty(B, T, 2 * H, device=dev)
    best, mean = timeit(lambda: ops.gru_layer_fwd_tc(gi, L, *w, passes=passes, order=order, out=out, gi_blocked=(passes == 1 and not a.standard_gi)))
    print(f"gru layer {a.precision} B={B}: best {best:.3f} ms mean {mean:.3f} ms  {best * 1e3 / T:.2f} us per step")
