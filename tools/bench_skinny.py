"""Skinny LoRA-side GEMMs of one decoder layer (M = B*2048 tokens) against their HBM floor and torch.matmul."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit

M = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)


def run(kind, fn, nbytes):
    def f():
        flush.zero_()          # inputs must come from HBM, as in the step (they were written many kernels earlier)
        fn()
    t_all = timeit(f, iters=10)
    t_flush = timeit(lambda: flush.zero_(), iters=10)
    t = t_all - t_flush
    print(json.dumps(dict(kind=kind, M=M, us=round(t * 1e3, 1), GBs=round(nbytes / t / 1e6), frac_hbm=round(nbytes / t / 1e6 / 6447.8, 3))))


for K, N in ((4096, 48), (4096, 16), (14336, 16)):
    x = torch.randn(M, K, device="cuda").bfloat16(); A = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    run("xA K=%d N=%d" % (K, N), lambda: ops.gemm(x, A, out), 2 * M * K)
    run("torch xA K=%d N=%d" % (K, N), lambda: torch.matmul(x, A.t(), out=out), 2 * M * K)
for K, N in ((28672, 32), (6144, 48)):
    dy = torch.randn(M, K, device="cuda").bfloat16(); Bc = torch.randn(K, N, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    run("dY.B K=%d N=%d" % (K, N), lambda: ops.gemm(dy, Bc, out, b_mn=True), 2 * M * K)
for N, r in ((14336, 16), (4096, 48)):
    t = torch.randn(M, r, device="cuda").bfloat16(); x = torch.randn(M, N, device="cuda").bfloat16()
    g = torch.zeros(r, N, device="cuda")
    run("grad r=%d N=%d" % (r, N), lambda: ops.gemm(t, x, g, a_mn=True, b_mn=True, out_f32=True, accumulate=True), 2 * M * N)
