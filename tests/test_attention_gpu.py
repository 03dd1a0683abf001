"""tcgen05 flash attention vs fp32 softmax attention on the same bf16-rounded inputs (tolerance 2e-2 abs on O(1) outputs:
bf16 P and bf16 output rounding)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(qkv, B, S, Hq, Hkv, D, causal):
    q, k, v = qkv.float().split([Hq * D, Hkv * D, Hkv * D], dim=1)
    q = q.view(B, S, Hq, D).transpose(1, 2)
    k = k.view(B, S, Hkv, D).transpose(1, 2).repeat_interleave(Hq // Hkv, dim=1)
    v = v.view(B, S, Hkv, D).transpose(1, 2).repeat_interleave(Hq // Hkv, dim=1)
    s = q @ k.transpose(-1, -2) / math.sqrt(D)
    if causal:
        s = s.masked_fill(torch.ones(S, S, device=s.device, dtype=torch.bool).triu(1), float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    o = torch.softmax(s, dim=-1) @ v
    return o.transpose(1, 2).reshape(B * S, Hq * D), lse


@pytest.mark.parametrize("B,S,Hq,Hkv,D,causal", [
    (1, 128, 1, 1, 128, True), (1, 256, 2, 1, 128, True), (2, 512, 8, 2, 128, True), (1, 2048, 4, 1, 128, True),
    (1, 128, 1, 1, 64, False), (3, 577, 4, 4, 64, False), (2, 200, 2, 2, 128, False), (1, 320, 2, 1, 64, True),
])
def test_attn_fwd(B, S, Hq, Hkv, D, causal):
    from phantom_vlb_b200 import ops
    torch.manual_seed(S + Hq)
    qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, causal)
    torch.cuda.synchronize()
    ro, rlse = _ref(qkv, B, S, Hq, Hkv, D, causal)
    err = (o.float() - ro).abs().max().item()
    assert err < 2e-2, "attention output max err %g" % err
    lerr = (lse * math.log(2.0) - rlse).abs().max().item()
    assert lerr < 2e-2, "lse max err %g" % lerr


def test_attn_fwd_large_logits():
    """rows whose running max grows by more than 2^8 between blocks exercise the lazy TMEM rescale."""
    from phantom_vlb_b200 import ops
    torch.manual_seed(5)
    B, S, Hq, Hkv, D = 1, 512, 2, 2, 128
    qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda")
    qkv[:, :Hq * D] *= 6.0
    qkv[S // 2:, Hq * D:(Hq + Hkv) * D] *= 4.0          # later keys give much larger logits
    qkv = qkv.bfloat16()
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
    torch.cuda.synchronize()
    ro, _ = _ref(qkv, B, S, Hq, Hkv, D, True)
    assert (o.float() - ro).abs().max().item() < 3e-2


@pytest.mark.parametrize("B,S,Hq,Hkv,causal", [(1, 128, 1, 1, True), (1, 256, 2, 1, True), (2, 512, 8, 2, True),
                                               (1, 1024, 4, 1, True), (1, 256, 2, 2, False)])
def test_attn_bwd(B, S, Hq, Hkv, causal):
    from phantom_vlb_b200 import ops
    D = 128
    torch.manual_seed(S * 3 + Hq)
    qkv = torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
    do = torch.randn(B * S, Hq * D, device="cuda").bfloat16()
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, causal)
    dq, dqkv = ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, causal)
    torch.cuda.synchronize()
    x = qkv.float().requires_grad_(True)
    ro, _ = _ref(x, B, S, Hq, Hkv, D, causal)
    (ro * do.float()).sum().backward()
    g = x.grad
    rq, rk, rv = g.split([Hq * D, Hkv * D, Hkv * D], dim=1)
    dk = dqkv[:, Hq * D:(Hq + Hkv) * D].float()
    dv = dqkv[:, (Hq + Hkv) * D:].float()
    for name, a, r in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv)):
        err = (a - r).abs().max().item()
        scale = r.abs().max().item()
        assert err < 2e-2 * scale + 1e-3, "%s max err %g (scale %g)" % (name, err, scale)
