"""Attention forward / backward, ours vs the libraries the reference can reach on this GPU (torch SDPA, flash_attn 2):
decoder shape (B=8, S=2048, 32/8 heads, d=128, causal) and ViT shape (64 frames x 577 tokens, 16 heads, d=64, non-causal)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
from tools.bench_gemm import timeit


def bench(name, B, S, Hq, Hkv, D, causal):
    qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
    fl = 4.0 * B * Hq * S * S * D / (2 if causal else 1)
    t = timeit(lambda: ops.attn_fwd(qkv, B, S, Hq, Hkv, D, causal))
    print(json.dumps(dict(kind=name + "_fwd", impl="ours", ms=round(t, 4), tflops=round(fl / t / 1e9, 1))), flush=True)
    q = qkv[:, :Hq * D].view(B, S, Hq, D)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].view(B, S, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(B, S, Hkv, D)
    qt, kt, vt = (x.transpose(1, 2).contiguous().requires_grad_() for x in (q, k, v))
    f = lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, enable_gqa=Hq != Hkv)
    t2 = timeit(f)
    print(json.dumps(dict(kind=name + "_fwd", impl="torch_sdpa", ms=round(t2, 4), tflops=round(fl / t2 / 1e9, 1))), flush=True)
    o_t = f()
    g_t = torch.randn_like(o_t)
    tb = timeit(lambda: torch.autograd.grad(o_t, (qt, kt, vt), g_t, retain_graph=True))
    print(json.dumps(dict(kind=name + "_bwd", impl="torch_sdpa", ms=round(tb, 4), tflops=round(2.5 * fl / tb / 1e9, 1))), flush=True)
    try:
        from flash_attn import flash_attn_func
        qf, kf, vf = (x.contiguous().requires_grad_() for x in (q, k, v))
        ff = lambda: flash_attn_func(qf, kf, vf, causal=causal)
        t3 = timeit(ff)
        print(json.dumps(dict(kind=name + "_fwd", impl="flash_attn2", ms=round(t3, 4), tflops=round(fl / t3 / 1e9, 1))), flush=True)
        o_f = ff()
        g_f = torch.randn_like(o_f)
        t4 = timeit(lambda: torch.autograd.grad(o_f, (qf, kf, vf), g_f, retain_graph=True))
        print(json.dumps(dict(kind=name + "_bwd", impl="flash_attn2", ms=round(t4, 4), tflops=round(2.5 * fl / t4 / 1e9, 1))), flush=True)
    except Exception as e:
        print(json.dumps(dict(kind=name, impl="flash_attn2", error=repr(e)[:200])), flush=True)
    if D == 128 and causal:
        do = torch.randn(B * S, Hq * D, device="cuda").bfloat16()
        o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
        dqkv = torch.empty_like(qkv)
        dq = torch.empty(B * S, Hq * D, device="cuda")
        delta = torch.empty(B, Hq, S, device="cuda")
        t5 = timeit(lambda: ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True, dqkv=dqkv, dq_f32=dq, delta=delta))
        print(json.dumps(dict(kind=name + "_bwd", impl="ours", ms=round(t5, 4), tflops=round(2.5 * fl / t5 / 1e9, 1))), flush=True)


bench("decoder_causal_d128", 8, 2048, 32, 8, 128, True)
bench("vit_d64", 64, 577, 16, 16, 64, False)
