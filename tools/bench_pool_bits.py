"""Timings of the pooled-head kernels (B=8, S=2048, H=4096, nch=64, ~60 % of the rows without HRF weight) and of the keep-mask
generator at the decoder's two site widths; CUDA events, 20 iterations after warm-up."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import lib as _l
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit

dev = "cuda"
B, S, H, nch = 8, 2048, 4096, 64
x = torch.randn(B, S, H, device=dev).bfloat16()
wm = torch.rand(B, S, device=dev)
wm = torch.where(wm < 0.6, torch.zeros_like(wm), wm)
live = float((wm != 0).float().mean())
w = [torch.ones(H, device=dev).bfloat16() for _ in range(3)]
b = [torch.zeros(H, device=dev).bfloat16() for _ in range(2)]
partial = torch.empty(B, nch, H, device=dev)
pooled, hrf = torch.empty(B, H, device=dev).bfloat16(), torch.empty(B, H, device=dev).bfloat16()
dpooled = torch.randn(B, H, device=dev)
dx = torch.empty_like(x)
dg1, db1 = torch.zeros(H, device=dev), torch.zeros(H, device=dev)
st = _l.cur_stream()
t = timeit(lambda: _l.call("vlb_hrf_pool_fwd", _l.ptr(x), _l.ptr(w[0]), _l.ptr(w[1]), _l.ptr(b[0]), _l.ptr(wm), _l.ptr(partial),
                           B, S, H, nch, 1e-5, 1e-5, st))
gb = live * B * S * H * 2 / 1e9
print(json.dumps(dict(kernel="hrf_pool_fwd", us=round(t * 1e3, 1), live_rows=round(live, 3), GBs=round(gb / t * 1e3, 1))), flush=True)
t = timeit(lambda: _l.call("vlb_hrf_pool_finalize", _l.ptr(partial), _l.ptr(w[2]), _l.ptr(b[1]), _l.ptr(pooled), _l.ptr(hrf), None,
                           B, H, nch, 1e-5, 0.0, 1, st))
print(json.dumps(dict(kernel="hrf_pool_finalize", us=round(t * 1e3, 1), GBs=round(B * nch * H * 4 / 1e9 / t * 1e3, 1))), flush=True)
t = timeit(lambda: _l.call("vlb_hrf_pool_bwd", _l.ptr(x), _l.ptr(w[0]), _l.ptr(w[1]), _l.ptr(wm), _l.ptr(dpooled), _l.ptr(dx),
                           _l.ptr(dg1), _l.ptr(db1), B, S, H, nch, 1e-5, 1e-5, st))
gb = (live * B * S * H * 2 + B * S * H * 2) / 1e9
print(json.dumps(dict(kernel="hrf_pool_bwd", us=round(t * 1e3, 1), GBs=round(gb / t * 1e3, 1))), flush=True)
M = 16384
for K in (4096, 14336):
    bits = torch.empty(K // 64, M, device=dev, dtype=torch.int64)
    t = timeit(lambda: ops.dropout_bits(M, K, 0.1, 1234, bits))
    keep = float(ops.unpack_bits(bits[:, :2048]).float().mean())
    print(json.dumps(dict(kernel="dropout_bits", M=M, K=K, us=round(t * 1e3, 1), Gbit_s=round(M * K / t / 1e6, 1), keep=round(keep, 5))), flush=True)
