"""Micro-benchmark of the tcgen05 GEMM on the decoder shapes (M = B*2048) against torch.matmul (cuBLASLt)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    shapes = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
    rows = []
    for N, K in shapes:
        a = torch.randn(M, K, device="cuda").bfloat16()
        w = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
        xa = torch.randn(M, 48, device="cuda").bfloat16()
        lb = torch.randn(N, 48, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        for bn in (128, 256):
            t = timeit(lambda: ops.gemm(a, w, out, block_n=bn))
            t2 = timeit(lambda: ops.gemm(a, w, out, a2=xa, b2=lb, block_n=bn))
            rows.append(dict(kind="fwd", M=M, N=N, K=K, bn=bn, ms=t, tflops=fl / t / 1e9, ms_lora=t2,
                             tflops_lora=fl / t2 / 1e9))
        tt = timeit(lambda: torch.matmul(a, w.t(), out=out))
        rows.append(dict(kind="torch", M=M, N=N, K=K, ms=tt, tflops=fl / tt / 1e9))
        # backward dX = dY @ W  (W stored [N,K] -> MN-major B operand)
        dy = torch.randn(M, N, device="cuda").bfloat16()
        dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
        for bn in (128, 256):
            t = timeit(lambda: ops.gemm(dy, w, dx, b_mn=True, block_n=bn))
            rows.append(dict(kind="bwd_dx", M=M, N=K, K=N, bn=bn, ms=t, tflops=fl / t / 1e9))
        tt = timeit(lambda: torch.matmul(dy, w, out=dx))
        rows.append(dict(kind="torch_dx", M=M, N=K, K=N, ms=tt, tflops=fl / tt / 1e9))
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
