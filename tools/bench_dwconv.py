"""STC depthwise 3x3 at the step's sizes (B=8 clips: 64 frames of 24x24, 40 of 13x13, C=4096) against its HBM floor."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit

flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
for F, Hh, Ww, C in ((64, 24, 24, 4096), (40, 13, 13, 4096)):
    x = torch.randn(F * Hh * Ww, C, device="cuda").bfloat16(); w = torch.randn(9, C, device="cuda").bfloat16()
    y = torch.empty_like(x)
    def f():
        flush.zero_(); ops.dwconv3x3(x, w, F, Hh, Ww, y)
    t = timeit(f, iters=20) - timeit(lambda: flush.zero_(), iters=20)
    nbytes = 2 * x.numel() * 2
    print(json.dumps(dict(kind="dwconv3x3", F=F, H=Hh, W=Ww, C=C, us=round(t * 1e3, 1), GBs=round(nbytes / t / 1e6))))
