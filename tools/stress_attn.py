"""Full-size repeated runs of both attention kernels: NaN / run-to-run differences reveal ordering bugs that small tests miss."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from phantom_vlb_b200 import ops
B, S, Hq, Hkv, D = 8, 2048, 32, 8, 128
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda", generator=g).bfloat16()
do = torch.randn(B * S, Hq * D, device="cuda", generator=g).bfloat16()
q = qkv[:, :Hq * D].view(B, S, Hq, D).transpose(1, 2)
k = qkv[:, Hq * D:(Hq + Hkv) * D].view(B, S, Hkv, D).transpose(1, 2).repeat_interleave(Hq // Hkv, dim=1)
v = qkv[:, (Hq + Hkv) * D:].view(B, S, Hkv, D).transpose(1, 2).repeat_interleave(Hq // Hkv, dim=1)
ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B * S, Hq * D)
o0 = None
for i in range(30):
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
    if o0 is None:
        o0 = o.clone()
        print("fwd max err vs sdpa", float((o.float() - ref.float()).abs().max()), "nan", bool(torch.isnan(o).any()))
    elif not torch.equal(o, o0):
        d = (o.float() - o0.float()).abs()
        print("fwd run", i, "differs: max", float(d.max()), "count", int((d > 0).sum()), "nan", bool(torch.isnan(o).any()))
dq0 = dkv0 = None
for i in range(30):
    dq, dqkv = ops.attn_bwd(qkv, o0, do, lse, B, S, Hq, Hkv, D, True)
    dkv = dqkv[:, Hq * D:].clone()
    if dq0 is None:
        dq0, dkv0 = dq.clone(), dkv
        print("bwd nan", bool(torch.isnan(dq).any()), bool(torch.isnan(dkv).any()), "|dq|max", float(dq.abs().max()))
    else:
        e1 = float((dq - dq0).abs().max()); e2 = float((dkv.float() - dkv0.float()).abs().max())
        if e1 > 1e-3 or e2 > 0 or torch.isnan(dq).any():
            print("bwd run", i, "dq diff", e1, "dkv diff", e2, "nan", bool(torch.isnan(dq).any()), bool(torch.isnan(dkv).any()))
print("done")
