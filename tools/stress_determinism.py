"""Repeat each tensor-core stage on fixed inputs and report run-to-run deviations (race hunting; GPU only).
compute-sanitizer is closed on this GPU pool, so bit-identity of repeated launches under allocator churn, together with
the oracle comparisons of the test-suite, is the race check: every stage below is deterministic by construction
(fixed reduction orders), so ANY differing bit is a synchronisation bug.  usage: python tools/stress_determinism.py [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from silent_speech_b200 import ops, synth

dev = torch.device("cuda")
p = synth.synth_params(7, use_roi=False)
B, T = 260, 64
d = synth.clip_batch(B, T, mode="uniform", seed=9, use_roi=False)
X = torch.from_numpy(d["X"]).to(dev); L = torch.from_numpy(d["lengths"]).to(dev)
G3 = 576
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
mask = (torch.arange(T, device=dev)[None, :] < L[:, None])


def dev_of(name):
    return p[name].to(dev).contiguous()


wih = [dev_of("gru.weight_ih_l0"), dev_of("gru.weight_ih_l0_reverse")]
whh = [dev_of("gru.weight_hh_l0"), dev_of("gru.weight_hh_l0_reverse")]
bih = [dev_of("gru.bias_ih_l0"), dev_of("gru.bias_ih_l0_reverse")]
bhh = [dev_of("gru.bias_hh_l0"), dev_of("gru.bias_hh_l0_reverse")]
fold = ops.gru_fold_bias(bih[0], bhh[0], bih[1], bhh[1])
order = torch.argsort(L, descending=True).to(torch.int32)
for passes in (3, 1):
    gi_ref = None
    out_ref = None
    bad_g = bad_r = 0
    for it in range(reps):
        # churn the allocator so buffers move around / contain garbage
        junk = torch.randn(1 << 20, device=dev) * 1e3
        gi = torch.empty(B, T, 2 * G3, device=dev)
        gi.fill_(float("nan"))
        gi2d = gi.view(B * T, 2 * G3)
        for dd in (0, 1):
            ops.gemm_tc(X.view(B * T, 180), wih[dd], fold[dd * G3:(dd + 1) * G3], gi2d, passes=passes, lengths=L, T=T, c_col0=dd * G3)
        g = torch.where(mask[..., None], gi, torch.zeros_like(gi))
        if gi_ref is None:
            gi_ref = g.clone()
        elif not torch.equal(g, gi_ref):
            bad_g += 1
            print(f"passes={passes} it={it}: gemm_tc differs, max abs {float((g - gi_ref).abs().max()):.3e}, nan={bool(torch.isnan(g).any())}")
        out = ops.gru_layer_fwd_tc(gi_ref.clone() if False else gi, L, whh[0], bhh[0], whh[1], bhh[1], passes=passes, order=order)
        if out_ref is None:
            out_ref = out.clone()
        elif not torch.equal(out, out_ref):
            bad_r += 1
            print(f"passes={passes} it={it}: gru_tc differs, max abs {float((out - out_ref).abs().max()):.3e}")
        del junk
    torch.cuda.synchronize()
    print(f"passes={passes}: {reps} reps, gemm mismatches {bad_g}, gru mismatches {bad_r}")

# single-pass recurrence fed through the GRU-blocked gi layout (bulk-copy staging): run-to-run AND against the standard layout
perm = torch.arange(2 * G3, device=dev).view(2, 3, 4, 48).permute(0, 2, 1, 3).reshape(-1)
for Bs, Ts in ((128, 3), (129, 5), (260, 64), (700, 40)):
    g = torch.Generator().manual_seed(Bs)
    Ls = torch.randint(1, Ts + 1, (Bs,), generator=g); Ls[0] = Ts
    Ls = Ls.to(dev)
    gi_s = (torch.randn(Bs, Ts, 2 * G3, generator=g) * 0.7).to(dev)
    gi_b = gi_s[:, :, perm].contiguous()
    od = torch.argsort(Ls, descending=True).to(torch.int32)
    bad = 0
    for o in (None, od):
        std = ops.gru_layer_fwd_tc(gi_s, Ls, whh[0], bhh[0], whh[1], bhh[1], passes=1, order=o)
        for it in range(reps):
            which = it & 1
            out = ops.gru_layer_fwd_tc(gi_b if which else gi_s, Ls, whh[0], bhh[0], whh[1], bhh[1], passes=1, order=o, gi_blocked=bool(which))
            if not torch.equal(out, std):
                bad += 1
    torch.cuda.synchronize()
    print(f"gi staging B={Bs} T={Ts}: {2 * reps} runs (standard / blocked alternating), mismatches {bad}")


# K2 (TMA frame ring, named barriers per slot, TMEM ping-pong, dynamic frame hand-out) and the persistent projection GEMMs
pr = {k: v.to(dev).contiguous() for k, v in synth.synth_params(13).items()}
names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
         "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
wc = [pr[n] for n in names]
dr = synth.clip_batch(200, 90, mode="uniform", seed=4)
Rd, Ld = torch.from_numpy(dr["R"]).to(dev), torch.from_numpy(dr["lengths"]).to(dev)
for passes in (3, 2, 1):      # 2: the tc3 inference form (four narrow slots, hi|lo stacked, conv2 tiles one after the other)
    first, bad = None, 0
    for it in range(reps):
        junk = torch.randn(1 << 20, device=dev) * 1e3
        e = ops.roi_cnn_fwd(Rd, Ld, wc, passes=passes)
        if first is None:
            first = e.clone()
        elif not torch.equal(e, first):
            bad += 1
            print(f"K2 passes={passes} it={it}: differs, max abs {float((e - first).abs().max()):.3e}")
        del junk
    torch.cuda.synchronize()
    print(f"K2 passes={passes}: {reps} reps over {int(Ld.sum())} frames, mismatches {bad}")
A = torch.randn(260 * 64, 384, device=dev)
Wa, Wb = pr["gru.weight_ih_l1"], pr["gru.weight_ih_l1_reverse"]
bias = torch.randn(2 * G3, device=dev)
for split, act_split in ((True, True), (True, False), (False, True)):
    packed = ops.proj_prepack(Wa, Wb, 384, 0, split=split)
    first, bad = None, 0
    for it in range(reps):
        junk = torch.randn(1 << 20, device=dev) * 1e3
        c = torch.full((260 * 64, 2 * G3), float("nan"), device=dev)
        ops.proj_gemm(A, packed, bias, c, 2 * G3, lengths=L, T=T, split=split, act_split=act_split)
        c = torch.where(mask.reshape(-1, 1), c, torch.zeros_like(c))
        if first is None:
            first = c.clone()
        elif not torch.equal(c, first):
            bad += 1
            print(f"proj_gemm split={split} act_split={act_split} it={it}: differs")
        del junk
    torch.cuda.synchronize()
    print(f"proj_gemm split={split} act_split={act_split}: {reps} reps, mismatches {bad}")
