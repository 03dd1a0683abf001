"""Time the decoder attention backward alone (B=4, S=2048, 32/8 heads, d=128, causal). usage: python tools/bench_attn_bwd.py [B]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
S, Hq, Hkv, D = 2048, 32, 8, 128
qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
do = torch.randn(B * S, Hq * D, device="cuda").bfloat16()
o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
dqkv = torch.empty_like(qkv)
dq = torch.empty(B * S, Hq * D, device="cuda")
delta = torch.empty(B, Hq, S, device="cuda")
t = timeit(lambda: ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True, dqkv=dqkv, dq_f32=dq, delta=delta))
fl = 2.5 * 4.0 * B * Hq * S * S * D / 2
print(json.dumps(dict(kind="attn_bwd_causal", dbg=os.environ.get("VLB_BW_DBG", "0"), B=B, S=S, ms=t, tflops=fl / t / 1e9)))
