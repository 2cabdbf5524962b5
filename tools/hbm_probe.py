"""Write-only / copy / read-only HBM bandwidth at the projection GEMM's output size (torch fill, copy, sum), to tell
whether a write-dominated kernel can be HBM-bound below the copy figure in MEASURED_PEAKS.json.  B200: fill 7.5 TB/s,
copy 6.5 TB/s, read 5.9 TB/s ("What the profile says about K3a").
    python tools/hbm_probe.py"""
import torch
dev=torch.device('cuda')
def t(fn, n=10):
    for _ in range(3): fn()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/n
N=368640*1152
a=torch.empty(N,dtype=torch.float32,device=dev); b=torch.empty(N,dtype=torch.float32,device=dev)
ms=t(lambda: a.zero_()); print(f"memset 1.7GB: {ms:.3f} ms  {N*4/ms/1e6:.0f} GB/s write-only")
ms=t(lambda: a.fill_(1.5)); print(f"fill   1.7GB: {ms:.3f} ms  {N*4/ms/1e6:.0f} GB/s write-only")
ms=t(lambda: b.copy_(a)); print(f"copy   1.7GB: {ms:.3f} ms  {2*N*4/ms/1e6:.0f} GB/s r+w")
s=torch.empty(368640*384,dtype=torch.float32,device=dev)
ms=t(lambda: torch.sum(s)); print(f"sum 566MB read-only: {ms:.3f} ms {s.numel()*4/ms/1e6:.0f} GB/s")
