import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit

B, S, Hq, Hkv, D = 4, 2048, 32, 8, 128
qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
t = timeit(lambda: ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True))
fl = 4.0 * B * Hq * S * S * D / 2
print(json.dumps(dict(kind="attn_fwd_causal", B=B, S=S, ms=t, tflops=fl / t / 1e9)))
q = qkv[:, :Hq * D].view(B, S, Hq, D).transpose(1, 2)
k = qkv[:, Hq * D:(Hq + Hkv) * D].view(B, S, Hkv, D).transpose(1, 2)
v = qkv[:, (Hq + Hkv) * D:].view(B, S, Hkv, D).transpose(1, 2)
t2 = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True))
print(json.dumps(dict(kind="torch_sdpa", ms=t2, tflops=fl / t2 / 1e9)))
N, T = 32, 577
qkv2 = torch.randn(N * T, 3 * 1024, device="cuda").bfloat16()
t = timeit(lambda: ops.attn_fwd(qkv2, N, T, 16, 16, 64, False, need_lse=False))
fl = 4.0 * N * 16 * T * T * 64
print(json.dumps(dict(kind="attn_fwd_vit", N=N, ms=t, tflops=fl / t / 1e9)))
