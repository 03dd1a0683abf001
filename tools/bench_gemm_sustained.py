"""Sustained (power-capped) throughput of the decoder GEMM shapes: each config runs back to back for ~1 s."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops

M = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
shapes = [("qkv", 6144, 4096, False), ("o", 4096, 4096, False), ("gu", 28672, 4096, False), ("d", 4096, 14336, False),
          ("d_gu", 4096, 28672, True), ("d_d", 14336, 4096, True)]
bufs = {}
for name, N, K, bmn in shapes:
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn((K, N) if bmn else (N, K), device="cuda") * 0.02).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    bufs[name] = (a, w, out, bmn, 2.0 * M * N * K)


def sustained(fn, secs=1.0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.time(); n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(5):
            fn()
        n += 5
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


which = os.environ.get("VLB_GEMM_GROUP_M", "8")
for name, (a, w, out, bmn, fl) in bufs.items():
    t = sustained(lambda: ops.gemm(a, w, out, b_mn=bmn))
    tt = sustained(lambda: torch.matmul(a, w if bmn else w.t(), out=out))
    print(json.dumps(dict(shape=name, group_m=which, ms=round(t, 3), tflops=round(fl / t / 1e9), cublas=round(fl / tt / 1e9))))
