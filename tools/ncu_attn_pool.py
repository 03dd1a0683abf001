"""Launches for one `ncu --set full -k regex:"attn|hrf_pool"` pass: decoder attention forward / backward (B=8, S=2048, 32/8
heads, d=128, causal), ViT attention (64 frames x 577 tokens, 16 heads, d=64, non-causal) and the pooled head forward /
backward (B=8, S=2048, H=4096) through the same C-ABI calls the step makes."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import lib as _l
from phantom_vlb_b200 import ops

dev = "cuda"
B, S, Hq, Hkv, D = 8, 2048, 32, 8, 128
qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device=dev).bfloat16()
do = torch.randn(B * S, Hq * D, device=dev).bfloat16()
for _ in range(2):
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
    ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True)
qv = torch.randn(64 * 577, 3 * 16 * 64, device=dev).bfloat16()
for _ in range(2):
    ops.attn_fwd(qv, 64, 577, 16, 16, 64, False, need_lse=False)

H, nch = 4096, 64
x = torch.randn(B, S, H, device=dev).bfloat16()
wm = torch.rand(B, S, device=dev)
wm = torch.where(wm < 0.6, torch.zeros_like(wm), wm)                 # ~60 % of the tokens carry no HRF weight
w = [torch.ones(H, device=dev).bfloat16() for _ in range(3)]
b = [torch.zeros(H, device=dev).bfloat16() for _ in range(2)]
partial = torch.empty(B, nch, H, device=dev)
pooled, hrf = torch.empty(B, H, device=dev).bfloat16(), torch.empty(B, H, device=dev).bfloat16()
dpooled = torch.randn(B, H, device=dev)
dx = torch.empty_like(x)
dg1, db1 = torch.zeros(H, device=dev), torch.zeros(H, device=dev)
st = _l.cur_stream()
for _ in range(2):
    _l.call("vlb_hrf_pool_fwd", _l.ptr(x), _l.ptr(w[0]), _l.ptr(w[1]), _l.ptr(b[0]), _l.ptr(wm), _l.ptr(partial), B, S, H, nch,
            1e-5, 1e-5, st)
    _l.call("vlb_hrf_pool_finalize", _l.ptr(partial), _l.ptr(w[2]), _l.ptr(b[1]), _l.ptr(pooled), _l.ptr(hrf), None, B, H, nch,
            1e-5, 0.0, 1, st)
    _l.call("vlb_hrf_pool_bwd", _l.ptr(x), _l.ptr(w[0]), _l.ptr(w[1]), _l.ptr(wm), _l.ptr(dpooled), _l.ptr(dx), _l.ptr(dg1),
            _l.ptr(db1), B, S, H, nch, 1e-5, 1e-5, st)
torch.cuda.synchronize()
print("ok")
