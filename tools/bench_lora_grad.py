"""Rank-r gradient GEMM g[r, N] += t[M, r]^T x[M, N] (M = 16384 tokens) across tile widths / split-K, against the HBM floor.
x is a column slice of a wider buffer when ld > N (d(gate|up) halves)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit
M = 16384
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
t_flush = timeit(lambda: flush.zero_(), iters=10)
for N, ld, r in ((14336, 28672, 16), (14336, 14336, 16), (4096, 4096, 16), (4096, 6144, 16)):
    t = torch.randn(M, r, device="cuda").bfloat16()
    xw = torch.randn(M, ld, device="cuda").bfloat16()
    x = xw[:, :N]
    g = torch.zeros(r, N, device="cuda")
    for bn in (0, 64, 128, 256):
        for sk in (0, 2, 4, 8):
            def f():
                flush.zero_()
                ops.gemm(t, x, g, a_mn=True, b_mn=True, out_f32=True, accumulate=True, block_n=bn, split_k=sk)
            try:
                tt = timeit(f, iters=10) - t_flush
            except Exception as e:
                print("N=%d ld=%d bn=%d sk=%d failed: %s" % (N, ld, bn, sk, str(e)[:80])); continue
            nb = 2 * M * N
            print(json.dumps(dict(N=N, ld=ld, bn=bn, split_k=sk, us=round(tt * 1e3, 1), frac_hbm=round(nb / tt / 1e6 / 6548.8, 3))))
