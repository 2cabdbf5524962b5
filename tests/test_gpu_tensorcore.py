"""GPU parity of the tcgen05 (tensor-core) kernels against fp64 matmuls and the fp32 kernels."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("M,N,K1,K2", [(128, 192, 64, 0), (300, 1152, 180, 32), (1000, 1152, 384, 0), (77, 200, 20, 0)])
def test_gemm_tc_vs_fp64(M, N, K1, K2):
    dev = _cuda()
    from silent_speech_b200 import ops
    g = torch.Generator().manual_seed(M + N)
    A1 = torch.randn(M, K1, generator=g)
    A2 = torch.randn(M, K2, generator=g) if K2 else None
    W = torch.randn(N, K1 + K2, generator=g) * 0.1
    b = torch.randn(N, generator=g)
    A = A1 if A2 is None else torch.cat([A1, A2], 1)
    ref = A.double() @ W.double().T + b.double()
    for passes, tol in ((3, 2e-5), (1, 3e-3)):
        C = torch.full((M, N + 8), -7.0, device=dev)
        ops.gemm_tc(A1.to(dev), W.to(dev), b.to(dev), C, A2=None if A2 is None else A2.to(dev), passes=passes, c_col0=8)
        e = rel_err(C[:, 8:], ref)
        assert e < tol, (passes, e)
        assert torch.all(C[:, :8] == -7.0)
    # accumulate + row mask
    T = 10
    Mr = (M // T) * T
    if Mr >= T:
        L = torch.full((Mr // T,), T, dtype=torch.int64)
        C = torch.ones(Mr, N, device=dev)
        ops.gemm_tc(A1[:Mr].to(dev), W.to(dev), None, C, A2=None if A2 is None else A2[:Mr].to(dev), passes=3,
                    accumulate=True, lengths=L.to(dev), T=T)
        assert rel_err(C, A[:Mr].double() @ W.double().T + 1.0) < 2e-5


@pytest.mark.parametrize("B,T", [(5, 7), (200, 23), (700, 40)])
def test_gru_tc_recurrence_vs_fp32_kernel(B, T):
    dev = _cuda()
    from oracle import ref_model
    from silent_speech_b200 import ops
    p = ref_model.synth_params(21, scale=1.5)
    g = torch.Generator().manual_seed(B)
    L = torch.randint(1, T + 1, (B,), generator=g); L[0] = T; L[min(1, B - 1)] = 1
    gi = (torch.randn(B, T, 1152, generator=g) * 0.7).to(dev)
    w = [p[f"gru.{n}"].to(dev).contiguous() for n in ("weight_hh_l0", "bias_hh_l0", "weight_hh_l0_reverse", "bias_hh_l0_reverse")]
    Ld = L.to(dev)
    gates_ref = torch.zeros(B, T, 2, 4, 192, device=dev)
    ref = ops.gru_layer_fwd(gi, Ld, *w, gates=gates_ref)
    # the TC kernel expects b_hr, b_hz folded into gi
    fold = torch.zeros(1152, device=dev)
    fold[:384] = w[1][:384]; fold[576:960] = w[3][:384]
    gi_f = (gi + fold).contiguous()
    order = torch.argsort(Ld, descending=True).to(torch.int32)
    mask = (torch.arange(T)[None, :] < L[:, None]).to(dev)
    for passes, tol in ((3, 2e-5), (1, 5e-3)):
        for od in (None, order):
            gates = torch.zeros(B, T, 2, 4, 192, device=dev)
            out = ops.gru_layer_fwd_tc(gi_f, Ld, *w, passes=passes, order=od, gates=gates)
            torch.cuda.synchronize()
            e = rel_err(out, ref)
            assert e < tol, (passes, od is not None, e)
            assert torch.all(out[~mask] == 0)
            eg = rel_err(gates * mask[..., None, None, None], gates_ref * mask[..., None, None, None])
            assert eg < tol * 5, (passes, eg)
    # GRU-blocked gi (columns [dir][rank][gate][48], fed by the producer warp): same arithmetic on the same values,
    # so the result must equal the standard-layout single-pass result bit for bit
    perm = torch.arange(1152).view(2, 3, 4, 48).permute(0, 2, 1, 3).reshape(-1).to(dev)     # blocked col -> standard col
    gi_b = gi_f[:, :, perm].contiguous()
    for od in (None, order):
        std = ops.gru_layer_fwd_tc(gi_f, Ld, *w, passes=1, order=od)
        blk = ops.gru_layer_fwd_tc(gi_b, Ld, *w, passes=1, order=od, gi_blocked=True)
        torch.cuda.synchronize()
        assert torch.equal(std, blk)
    # fp16 storage of the blocked gi (double-buffered staging): identical to the standard kernel fed the fp16-rounded values
    gi_r = gi_f.half().float().contiguous()
    gi_h = gi_r[:, :, perm].half().contiguous()
    for od in (None, order):
        std = ops.gru_layer_fwd_tc(gi_r, Ld, *w, passes=1, order=od)
        blk = ops.gru_layer_fwd_tc(gi_h, Ld, *w, passes=1, order=od, gi_blocked=True)
        torch.cuda.synchronize()
        assert torch.equal(std, blk)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("tc3", 1e-4), ("fp16", 1e-3)])
@pytest.mark.parametrize("use_roi,B,T,mode", [(True, 150, 90, "uniform"), (False, 260, 64, "uniform")])
def test_model_logits_by_precision(precision, tol, use_roi, B, T, mode):
    """fp32 and tc3 (split-bf16 on the tensor cores) must meet the 1e-4 logits bar with identical argmax;
    fp16 (single pass, the inference bench default) has its own stated tolerance: 1e-3 of max |logit| and identical
    argmax on this batch (measured ~3e-4)."""
    import os
    dev = _cuda()
    from oracle import ref_model
    from silent_speech_b200 import BiGRUClassifier, synth
    torch.set_num_threads(os.cpu_count() or 1)
    p = ref_model.synth_params(7, use_roi=use_roi)
    d = synth.clip_batch(B, T, mode=mode, seed=9, use_roi=use_roi)
    X, L = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"])
    R = torch.from_numpy(d["R"]) if use_roi else None
    with torch.no_grad():
        ref = ref_model.forward(p, X, L, R)
    m = BiGRUClassifier(180, 10, use_roi=use_roi, precision=precision)
    m.load_state_dict(p)
    m = m.to(dev).eval()
    with torch.no_grad():
        logits = m(X.to(dev), L.to(dev), None if R is None else R.to(dev))
    e = rel_err(logits, ref)
    if precision == "fp16":          # opt-in fp16 storage of gi: same tolerance, same argmax
        m.gi_half = True
        with torch.no_grad():
            lg16 = m(X.to(dev), L.to(dev), None if R is None else R.to(dev))
        m.gi_half = False
        assert rel_err(lg16, ref) < tol and torch.equal(lg16.argmax(1).cpu(), ref.argmax(1))
    agree = (logits.argmax(1).cpu() == ref.argmax(1)).float().mean().item()
    print(f"precision={precision} roi={use_roi} rel_err={e:.3e} argmax_agree={agree:.4f}")
    assert e < tol, e
    assert agree == 1.0


@pytest.mark.parametrize("passes,tol", [(3, 1e-4), (2, 1e-4), (1, 3e-3)])
@pytest.mark.parametrize("standardize", [True, False])
def test_roi_cnn_tc_vs_oracle(passes, tol, standardize):
    import numpy as np
    dev = _cuda()
    from oracle import ref_model
    from silent_speech_b200 import ops, synth
    p = ref_model.synth_params(11)
    B, T = 7, 9
    L = np.array([9, 4, 1, 7, 9, 9, 2], np.int64)
    R = synth.rois(B, T, L, seed=3)
    R[0, 2] = 0
    R[4, 3, 10, 20] = 255; R[4, 3, 0, 0] = 255; R[4, 3, 47, 95] = 0      # corners / outliers
    names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
             "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
    w = [p[n].to(dev).contiguous() for n in names]
    Rd, Ld = torch.from_numpy(R).to(dev), torch.from_numpy(L).to(dev)
    emb = ops.roi_cnn_fwd(Rd, Ld, w, standardize=standardize, passes=passes)
    torch.cuda.synchronize()
    ref32 = ops.roi_cnn_fwd(Rd, Ld, w, standardize=standardize)
    ref = ref_model.roi_cnn(p, ref_model.roi_prepare(torch.from_numpy(R), standardize))
    mask = torch.arange(T)[None, :] < torch.from_numpy(L)[:, None]
    ref = ref * mask[..., None]
    print(f"cnn passes={passes} std={standardize} vs oracle {rel_err(emb, ref):.3e} vs fp32 kernel {rel_err(emb, ref32):.3e}")
    assert rel_err(emb, ref) < tol
    assert torch.all(emb.cpu()[~mask] == 0)


@pytest.mark.parametrize("passes,tol", [(3, 1e-4), (2, 1e-4), (1, 3e-3)])
def test_roi_cnn_tc3_many_frames_dynamic_handout(passes, tol):
    """More frames than resident slots, ragged lengths: exercises the atomic frame hand-out, the register
    prefetch across padded frames and slot re-use; checked against the fp32 CUDA-core kernel, and repeated
    launches must be bit-identical (no race in the hand-out)."""
    dev = _cuda()
    from oracle import ref_model
    from silent_speech_b200 import ops, synth
    p = ref_model.synth_params(13)
    B, T = 96, 90
    d = synth.clip_batch(B, T, mode="uniform", seed=4)
    names = ["roi_cnn.net.0.weight", "roi_cnn.net.0.bias", "roi_cnn.net.3.weight", "roi_cnn.net.3.bias",
             "roi_cnn.net.6.weight", "roi_cnn.net.6.bias", "roi_cnn.fc.weight", "roi_cnn.fc.bias"]
    w = [p[n].to(dev).contiguous() for n in names]
    Rd, Ld = torch.from_numpy(d["R"]).to(dev), torch.from_numpy(d["lengths"]).to(dev)
    ref32 = ops.roi_cnn_fwd(Rd, Ld, w, passes=0)
    a = ops.roi_cnn_fwd(Rd, Ld, w, passes=passes)
    b = ops.roi_cnn_fwd(Rd, Ld, w, passes=passes)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert rel_err(a, ref32) < tol
    mask = (torch.arange(T)[None, :] < torch.from_numpy(d["lengths"])[:, None])
    assert torch.all(a.cpu()[~mask] == 0)


@pytest.mark.parametrize("M,n_each,K1,K2,T", [(128, 576, 180, 32, 0), (1000, 576, 384, 0, 10), (77, 576, 180, 0, 0),
                                              (39960, 576, 180, 32, 90), (300, 200, 64, 0, 0)])
def test_proj_gemm_vs_fp64(M, n_each, K1, K2, T):
    """Persistent both-direction projection GEMM (packed fp16 weights, bulk-copy weight ring) against an fp64
    matmul of the fp16-rounded operands (exact products, fp32 accumulation error only) and of the raw operands."""
    dev = _cuda()
    from silent_speech_b200 import ops
    g = torch.Generator().manual_seed(M + K1)
    A1 = torch.randn(M, K1, generator=g)
    A2 = torch.randn(M, K2, generator=g) if K2 else None
    Wa = torch.randn(n_each, K1 + K2, generator=g) * 0.1
    Wb = torch.randn(n_each, K1 + K2, generator=g) * 0.1
    b = torch.randn(2 * n_each, generator=g)
    A = A1 if A2 is None else torch.cat([A1, A2], 1)
    W = torch.cat([Wa, Wb], 0)
    ref16 = A.half().double() @ W.half().double().T + b.double()
    ref = A.double() @ W.double().T + b.double()
    N = 2 * n_each
    C = torch.full((M, N + 4), -7.0, device=dev)
    Cv = C[:, :N]
    L = None
    if T:
        Lh = torch.randint(1, T + 1, (M // T,), generator=g)
        Lh[0] = T
        L = Lh.to(dev)
    packed = ops.proj_prepack(Wa.to(dev), Wb.to(dev), K1, K2)
    ops.proj_gemm(A1.to(dev), packed, b.to(dev), Cv, N, A2=None if A2 is None else A2.to(dev), lengths=L, T=T)
    torch.cuda.synchronize()
    out = C.cpu()
    assert torch.all(out[:, N:] == -7.0)
    if T:
        Mr = (M // T) * T
        mask = (torch.arange(T)[None, :] < Lh[:, None]).reshape(-1)
        live = torch.zeros(M, dtype=torch.bool); live[:Mr] = mask
        assert torch.all(out[~live][:, :N] == -7.0)          # rows of padded frames are not written
    else:
        live = torch.ones(M, dtype=torch.bool)
    assert rel_err(out[live][:, :N], ref16[live]) < 2e-6
    assert rel_err(out[live][:, :N], ref[live]) < 3e-3
    if n_each % 12 == 0:
        # GRU-blocked column order [dir][rank][gate][n_each/12]: the same numbers, permuted columns (bit for bit)
        u4 = n_each // 12
        perm = torch.arange(N).view(2, 3, 4, u4).permute(0, 2, 1, 3).reshape(-1)           # blocked col -> standard col
        C2 = torch.full((M, N), -7.0, device=dev)
        packed_b = ops.proj_prepack(Wa.to(dev), Wb.to(dev), K1, K2, gru_blocked=True)
        ops.proj_gemm(A1.to(dev), packed_b, b[perm].to(dev), C2, N, A2=None if A2 is None else A2.to(dev), lengths=L, T=T)
        torch.cuda.synchronize()
        assert torch.equal(C2.cpu()[live], out[live][:, :N][:, perm])
        # fp16 output: the same fp32 accumulators rounded once to fp16; dead rows untouched
        C3 = torch.full((M, N + 8), -7.0, device=dev, dtype=torch.float16)
        ops.proj_gemm(A1.to(dev), packed_b, b[perm].to(dev), C3[:, :N], N, A2=None if A2 is None else A2.to(dev), lengths=L, T=T)
        torch.cuda.synchronize()
        h = C3.cpu()
        assert torch.all(h[:, N:] == -7.0) and torch.all(h[~live][:, :N] == -7.0)
        assert torch.equal(h[live][:, :N], C2.cpu()[live].half())


@pytest.mark.parametrize("M,n_each,K1,K2,T", [(64, 576, 180, 32, 0), (1000, 576, 384, 0, 10), (77, 576, 180, 0, 0),
                                              (39960, 576, 180, 32, 90), (300, 200, 64, 0, 0), (130, 576, 384, 0, 0)])
@pytest.mark.parametrize("act_split", [True, False])
def test_proj_gemm_split_vs_fp64(M, n_each, K1, K2, T, act_split):
    """The fp32-parity projection GEMM (hi/lo split of both operands, weights as the M = 128 operand, transposed
    accumulator): within 2e-6 of the fp64 product of the raw fp32 operands (tolerance of the 1e-4 parity budget),
    rows of padded frames and the columns past N untouched.  act_split=False is the inference form (activations rounded
    to fp16 once: 128-row tiles, two MMAs per k-step): within 3e-4 of fp64 at the GEMM (the activations' rounding, 2^-12
    per element, averaged over K), and it is the model-level tests and bench.py's parity block that hold the logits to 1e-4."""
    dev = _cuda()
    from silent_speech_b200 import ops
    g = torch.Generator().manual_seed(M + K1 + 1)
    A1 = torch.randn(M, K1, generator=g)
    A2 = torch.randn(M, K2, generator=g) if K2 else None
    Wa = torch.randn(n_each, K1 + K2, generator=g) * 0.1
    Wb = torch.randn(n_each, K1 + K2, generator=g) * 0.1
    b = torch.randn(2 * n_each, generator=g)
    A = A1 if A2 is None else torch.cat([A1, A2], 1)
    ref = A.double() @ torch.cat([Wa, Wb], 0).double().T + b.double()
    N = 2 * n_each
    C = torch.full((M, N + 4), -7.0, device=dev)
    L = None
    if T:
        Lh = torch.randint(1, T + 1, (M // T,), generator=g)
        Lh[0] = T
        L = Lh.to(dev)
    packed = ops.proj_prepack(Wa.to(dev), Wb.to(dev), K1, K2, split=True)
    ops.proj_gemm(A1.to(dev), packed, b.to(dev), C[:, :N], N, A2=None if A2 is None else A2.to(dev), lengths=L, T=T, split=True,
                  act_split=act_split)
    torch.cuda.synchronize()
    out = C.cpu()
    assert torch.all(out[:, N:] == -7.0)
    live = torch.ones(M, dtype=torch.bool)
    if T:
        Mr = (M // T) * T
        live = torch.zeros(M, dtype=torch.bool)
        live[:Mr] = (torch.arange(T)[None, :] < Lh[:, None]).reshape(-1)
        assert torch.all(out[~live][:, :N] == -7.0)
    e = rel_err(out[live][:, :N], ref[live])
    print(f"proj split M={M} K={K1}+{K2} act_split={act_split}: rel err {e:.2e}")
    assert e < (2e-6 if act_split else 3e-4)
    # repeated launches are bit-identical (ring phases, accumulator hand-over)
    C2 = torch.full((M, N + 4), -7.0, device=dev)
    ops.proj_gemm(A1.to(dev), packed, b.to(dev), C2[:, :N], N, A2=None if A2 is None else A2.to(dev), lengths=L, T=T, split=True,
                  act_split=act_split)
    assert torch.equal(C2.cpu(), out)


@pytest.mark.parametrize("use_roi,B,chunk,pinned", [(True, 70, 16, True), (True, 37, 512, False), (False, 90, 8, True)])
def test_infer_host_matches_device_forward(use_roi, B, chunk, pinned):
    """The host-buffer entry (bench.py's e2e arm): chunked, copy/compute-overlapped inference from host tensors returns
    exactly what one device-resident forward call returns -- ramped chunk schedule, ragged last chunk, pageable inputs."""
    dev = _cuda()
    from oracle import ref_model
    from silent_speech_b200 import BiGRUClassifier, synth
    p = ref_model.synth_params(3, use_roi=use_roi)
    d = synth.clip_batch(B, 20, mode="uniform", seed=4, use_roi=use_roi)
    X, L = torch.from_numpy(d["X"]), torch.from_numpy(d["lengths"])
    R = torch.from_numpy(d["R"]) if use_roi else None
    if pinned:
        X, L, R = X.pin_memory(), L.pin_memory(), None if R is None else R.pin_memory()
    m = BiGRUClassifier(180, 10, use_roi=use_roi, precision="tc3")
    m.load_state_dict(p)
    m = m.to(dev).eval()
    with torch.no_grad():
        want = m(X.to(dev), L.to(dev), None if R is None else R.to(dev)).cpu()
    got = m.infer_host(X, L, R, chunk=chunk)
    assert got.shape == (B, 10) and not got.is_cuda
    # a clip's logits do not depend on which other clips share its launch (tile-local arithmetic, fixed-order sums)
    assert torch.equal(got, want)
    assert torch.equal(m.infer_host(X, L, R, chunk=chunk), got)          # buffers are reused across calls
