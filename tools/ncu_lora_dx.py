"""One launch set of the dX pair GEMM with the masked adapter epilogue (LORA_EPI) at the decoder's shapes, for ncu.
usage: python tools/ncu_lora_dx.py [qkv|o|gu|down][,...] [reps]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops

shapes = (sys.argv[1] if len(sys.argv) > 1 else "qkv").split(",")
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for which in shapes:
    M = 16384
    N, K, na = {"qkv": (4096, 6144, 3), "o": (4096, 4096, 1), "gu": (4096, 28672, 2), "down": (14336, 4096, 1)}[which]
    g = torch.Generator(device="cuda").manual_seed(0)
    dy = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    W = (torch.randn(K, N, device="cuda", generator=g) * K ** -0.5).bfloat16()
    t = torch.randn(M, 16 * na, device="cuda", generator=g).bfloat16()
    A = (torch.randn(16 * na, N, device="cuda", generator=g) * 0.25).bfloat16()
    bits = torch.empty((na, N // 64, M), device="cuda", dtype=torch.int64)
    for a in range(na):
        ops.dropout_bits(M, N, 0.1, 5 + a, bits[a])
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if which == "down":
        gu = torch.randn(M, 2 * N, device="cuda", generator=g).bfloat16()
        dgu = torch.empty_like(gu)
        fn = lambda: ops.gemm_down_dx_swiglu_bwd_masked(dy, W, gu, dgu, t, A, bits[0], 1 / 0.9)
        fn0 = lambda: ops.gemm_down_dx_swiglu_bwd(dy, W, gu, dgu, a2=t, b2=A)
    else:
        fn = lambda: ops.gemm_dx_lora_masked(dy, W, out, t, A, [bits[a] for a in range(na)], 1 / 0.9)
        fn0 = lambda: ops.gemm(dy, W, out, a2=t, b2=A, b_mn=True)
    for f, name in ((fn0, "fused k-block (no dropout)"), (fn, "masked epilogue")):
        for _ in range(2):
            f()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            f()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print("%s %s: %.3f ms  %.0f TF/s" % (which, name, ms, 2.0 * M * N * K / ms / 1e9))
