"""ssb_proj_gemm_static (the no-allocation entry point: packed weights in a file-scope static device array) against the
existing ssb_proj_prepack + ssb_proj_gemm pair, through the C ABI, on two calls in a row.

Collected last on purpose.  Both entry points launch the same two kernels on the same
operands, so the outputs are compared bit for bit (tighter than the 3e-3 the fp16 projection is held to elsewhere)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda")


def _operands(dev, M, n_each, K1, K2, two_dirs, seed):
    g = torch.Generator().manual_seed(seed)
    A1 = torch.randn(M, K1, generator=g).to(dev)
    A2 = torch.randn(M, K2, generator=g).to(dev) if K2 else None
    Wa = (torch.randn(n_each, K1 + K2, generator=g) * 0.1).to(dev)
    Wb = (torch.randn(n_each, K1 + K2, generator=g) * 0.1).to(dev) if two_dirs else None
    N = n_each * (2 if two_dirs else 1)
    bias = torch.randn(N, generator=g).to(dev)
    return A1, A2, Wa, Wb, bias, N


def _existing(A1, A2, Wa, Wb, bias, N, K1, K2, blocked, lengths, T, half):
    from silent_speech_b200 import ops
    packed = ops.proj_prepack(Wa, Wb, K1, K2, gru_blocked=blocked)
    C = torch.full((A1.shape[0], N), -3.0, device=A1.device, dtype=torch.float16 if half else torch.float32)
    ops.proj_gemm(A1, packed, bias, C, N, A2=A2, lengths=lengths, T=T)
    return C


def _static(A1, A2, Wa, Wb, bias, N, K1, K2, blocked, lengths, T, half):
    """Straight through the C ABI: nothing but the caller's inputs and outputs is passed."""
    from silent_speech_b200 import _lib
    from silent_speech_b200._lib import i32, i64, ptr, stream_ptr
    L = _lib.lib()
    C = torch.full((A1.shape[0], N), -3.0, device=A1.device, dtype=torch.float16 if half else torch.float32)
    free0 = torch.cuda.mem_get_info()[0]
    alloc0 = torch.cuda.memory_allocated()
    rc = L.ssb_proj_gemm_static(ptr(A1), i64(A1.stride(0)), i64(K1), ptr(A2), i64(0 if A2 is None else A2.stride(0)), i64(K2),
                                ptr(Wa), ptr(Wb), i64(Wa.stride(0)), i64(Wa.shape[0]), i32(1 if blocked else 0),
                                ptr(bias), ptr(C), i64(C.stride(0)), i64(A1.shape[0]),
                                ptr(lengths), i64(T), i32(1 if half else 0), stream_ptr())
    torch.cuda.synchronize()
    return rc, C, (free0, torch.cuda.mem_get_info()[0], alloc0, torch.cuda.memory_allocated())


def test_static_projection_matches_existing_entry_point_twice():
    dev = _cuda()
    from silent_speech_b200 import _lib
    L = _lib.lib()
    L.ssb_proj_static_max_bytes.restype = ctypes.c_int64
    assert L.ssb_proj_static_max_bytes() == 1152 * 384 * 2

    # call 1: the largest shape the workspace holds (layer-1 projection of the classifier: 384 -> 2 x 576, GRU-blocked
    # columns, fp16 gi, row mask); call 2: a smaller, differently shaped problem on top of call 1's leftovers
    cases = [dict(M=64 * 90, n_each=576, K1=384, K2=0, two=True, blocked=True, T=90, half=True, seed=1),
             dict(M=300, n_each=200, K1=180, K2=32, two=False, blocked=False, T=0, half=False, seed=2),
             dict(M=64 * 90, n_each=576, K1=180, K2=32, two=True, blocked=False, T=90, half=False, seed=3)]
    mem = []
    for c in cases:
        A1, A2, Wa, Wb, bias, N = _operands(dev, c["M"], c["n_each"], c["K1"], c["K2"], c["two"], c["seed"])
        lengths = None
        if c["T"]:
            g = torch.Generator().manual_seed(c["seed"])
            lengths = torch.randint(1, c["T"] + 1, (c["M"] // c["T"],), generator=g).to(dev)
        want = _existing(A1, A2, Wa, Wb, bias, N, c["K1"], c["K2"], c["blocked"], lengths, c["T"], c["half"])
        rc, got, m = _static(A1, A2, Wa, Wb, bias, N, c["K1"], c["K2"], c["blocked"], lengths, c["T"], c["half"])
        assert rc == 0, L.ssb_last_error()
        assert torch.equal(got, want), (c, float((got.float() - want.float()).abs().max()))
        # and both sit inside the fp16 projection's stated tolerance of an fp64 product
        if not c["blocked"] and lengths is None:
            A = A1 if A2 is None else torch.cat([A1, A2], 1)
            W = Wa if Wb is None else torch.cat([Wa, Wb], 0)
            ref = A.double() @ W.double().T + bias.double()
            assert float((got.double() - ref).abs().max() / ref.abs().max()) < 3e-3
        mem.append(m)
    # no allocation: once the kernels are loaded (first call), neither the driver's free memory nor torch's pool moves
    for free0, free1, a0, a1 in mem[1:]:
        assert free1 == free0 and a1 == a0, mem


def test_static_projection_rejects_what_does_not_fit():
    dev = _cuda()
    from silent_speech_b200 import _lib
    L = _lib.lib()
    L.ssb_last_error.restype = ctypes.c_char_p
    for n_each, K1, two in ((577, 64, True), (1160, 64, False), (64, 392, False)):
        A1, A2, Wa, Wb, bias, N = _operands(dev, 128, n_each, K1, 0, two, 5)
        rc, C, _ = _static(A1, A2, Wa, Wb, bias, N, K1, 0, False, None, 0, False)
        assert rc == 1, (n_each, K1, rc)                               # SSB_ERR_ARG
        assert b"static workspace" in L.ssb_last_error()
        assert torch.all(C == -3.0)                                    # nothing ran
