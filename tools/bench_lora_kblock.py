import sys, os, json
sys.path.insert(0, "/root/repo")
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit
M = 16384
for N, K in ((28672, 4096), (4096, 14336), (6144, 4096)):
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    t = timeit(lambda: ops.gemm(a, w, out))
    print(json.dumps(dict(N=N, K=K, K2=0, tflops=round(fl / t / 1e9))))
    for K2 in (16, 32, 48, 64, 128):
        xa = torch.randn(M, K2, device="cuda").bfloat16()
        lb = torch.randn(N, K2, device="cuda").bfloat16()
        t = timeit(lambda: ops.gemm(a, w, out, a2=xa, b2=lb))
        # padded storage: logical K2 columns inside a 64-wide zero-padded buffer
        xap = torch.zeros(M, 64 * ((K2 + 63) // 64), device="cuda", dtype=torch.bfloat16); xap[:, :K2] = xa
        lbp = torch.zeros(N, 64 * ((K2 + 63) // 64), device="cuda", dtype=torch.bfloat16); lbp[:, :K2] = lb
        t2 = timeit(lambda: ops.gemm(a, w, out, a2=xap, b2=lbp))
        print(json.dumps(dict(N=N, K=K, K2=K2, tflops=round(fl / t / 1e9), tflops_padded64=round(fl / t2 / 1e9))))
