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


@pytest.mark.parametrize("B,S,Hq,Hkv,D,causal", [(4, 2048, 32, 8, 128, True), (32, 577, 16, 16, 64, False)])
def test_attn_fwd_repeatable_when_warps_run_ahead(B, S, Hq, Hkv, D, causal):
    """The forward has no atomics: every repetition must be bit-identical, also on inputs whose row maxima keep growing so
    that some softmax warps take the O-rescale path while their siblings run a key block ahead. (r01d regression: a single
    p_ready barrier let a fast warp's arrival for block j+1 complete block j's phase; only full-size, many-CTA runs showed it.)"""
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda", generator=g) * 3).bfloat16()
    qkv[:, :Hq * D] *= (torch.rand(B * S, 1, device="cuda", generator=g) * 2).bfloat16()      # rows with very different score ranges
    o0, lse0 = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, causal)
    o0, lse0 = o0.clone(), lse0.clone()
    assert not torch.isnan(o0).any()
    ro, _ = _ref(qkv[:S], 1, S, Hq, Hkv, D, causal)
    assert (o0[:S].float() - ro).abs().max().item() < 0.15 * ro.abs().max().item()
    for i in range(60):
        o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, causal)
        assert torch.equal(o, o0) and torch.equal(lse, lse0), "run %d differs from run 0" % i


def test_attn_bwd_repeatable_full_size():
    """dK/dV are accumulated inside one CTA (deterministic); dQ is reduced across key-block CTAs by TMA reduce-adds, so it may
    differ in the last bits only. Full decoder shape, hard inputs, 20 repetitions."""
    from phantom_vlb_b200 import ops
    B, S, Hq, Hkv, D = 4, 2048, 32, 8, 128
    g = torch.Generator(device="cuda").manual_seed(12)
    qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda", generator=g) * 2).bfloat16()
    do = torch.randn(B * S, Hq * D, device="cuda", generator=g).bfloat16()
    o, lse = ops.attn_fwd(qkv, B, S, Hq, Hkv, D, True)
    dq0, dqkv0 = ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True)
    dq0, dkv0 = dq0.clone(), dqkv0[:, Hq * D:].clone()
    assert not torch.isnan(dq0).any() and not torch.isnan(dkv0.float()).any()
    scale = float(dq0.abs().max())
    for i in range(20):
        dq, dqkv = ops.attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, True)
        assert torch.equal(dqkv[:, Hq * D:], dkv0), "dK/dV of run %d differ" % i
        assert float((dq - dq0).abs().max()) < 1e-4 * scale, "dQ of run %d differs beyond reduction-order noise" % i
