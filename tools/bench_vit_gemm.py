"""ViT-shaped GEMMs (M = 64 frames x 577 tokens, K = 1024) with and without the bias / activation / residual epilogue."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit
for M in (36928, 36864):
    for N, K in ((4096, 1024), (1024, 1024), (3072, 1024), (1024, 4096)):
        a = torch.randn(M, K, device="cuda").bfloat16()
        w = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
        b = torch.randn(N, device="cuda").bfloat16()
        res = torch.randn(M, N, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        for name, kw in (("plain", {}), ("bias", dict(bias=b)), ("bias+qgelu", dict(bias=b, act=ops.ACT_QUICK_GELU)),
                         ("bias+res", dict(bias=b, residual=res))):
            t = timeit(lambda: ops.gemm(a, w, out, **kw), iters=20)
            print(json.dumps(dict(M=M, N=N, K=K, epi=name, us=round(t * 1e3, 1), tflops=round(2.0 * M * N * K / t / 1e9, 1))))
        t = timeit(lambda: torch.matmul(a, w.t(), out=out), iters=20)
        print(json.dumps(dict(M=M, N=N, K=K, epi="cublas plain", us=round(t * 1e3, 1), tflops=round(2.0 * M * N * K / t / 1e9, 1))))
