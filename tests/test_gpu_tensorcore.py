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
    for passes, tol in ((3, 2e-5), (1, 2e-2)):
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
