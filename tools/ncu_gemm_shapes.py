"""Two launches of the GEMM per decoder shape (M = B*2048), for `ncu -k regex:gemm` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
M = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
only = sys.argv[2].split(",") if len(sys.argv) > 2 else None
for name, N, K, bmn in [s for s in [("qkv", 6144, 4096, False), ("gu", 28672, 4096, False), ("d", 4096, 14336, False),
                        ("d_gu", 4096, 28672, True), ("d_d", 14336, 4096, True)] if only is None or s[0] in only]:
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn((K, N) if bmn else (N, K), device="cuda") * 0.02).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for _ in range(2):
        ops.gemm(a, w, out, b_mn=bmn)
    torch.cuda.synchronize()
    del a, w, out
