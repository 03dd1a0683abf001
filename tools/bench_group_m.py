"""Sustained timing (10 back-to-back launches) of the K = 4096 decoder GEMM shapes for several M-group sizes of the tile order
(VLB_GEMM_GROUP_M toggled between launches)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops

M = 16384
for name, N, K, bmn in (("qkv", 6144, 4096, False), ("gu", 28672, 4096, False), ("d_d", 14336, 4096, True), ("o", 4096, 4096, False)):
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn((K, N) if bmn else (N, K), device="cuda") * 0.02).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for gm in (16, 32, 64, 16, 32, 64):
        os.environ["VLB_GEMM_GROUP_M"] = str(gm)
        ops.gemm(a, w, out, b_mn=bmn)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.gemm(a, w, out, b_mn=bmn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("%s N=%d K=%d group_m=%d: %.3f ms  %.0f TF/s" % (name, N, K, gm, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
    del a, w, out
