"""Parity of the HBM-bound head / scoring / optimiser kernels with the oracle, through the C ABI.

  Pearson r (src/utils.py:101-106 -> torchmetrics PearsonCorrCoef after nan_to_num): fp32 <= 1e-4 vs the numpy fp64
      oracle (north_star), incl. bf16 inputs, NaN/Inf cleaning, ragged P (scalar fallback), constant columns, row shards.
  paint_labels (make_acc_brainmaps.py:50-58): integer label gather, bit-exact.
  clip + AdamW (litmodule.py:345-379, gradient_clip_val 1): vs torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW, 5 steps.
  row kernels (RMSNorm / RoPE / SwiGLU fwd+bwd): vs the fp32 torch formula on bf16-rounded inputs.
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,P,dtype", [(700, 1000, torch.float32), (700, 1000, torch.bfloat16), (64, 1003, torch.float32),
                                       (33, 7, torch.bfloat16), (5632, 4096, torch.bfloat16), (1, 8, torch.float32)])
def test_pearson_matches_fp64_oracle(N, P, dtype):
    from oracle.ref_model import pearson_r
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(N + P)
    x = torch.randn(N, P, generator=g)
    y = (0.3 * x + torch.randn(N, P, generator=g))
    x, y = x.to(dtype), y.to(dtype)
    r, stats = ops.pearson(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    ref = pearson_r(x.float().numpy(), y.float().numpy())
    got = r.cpu().numpy().astype(np.float64)
    if N == 1:                                           # zero variance -> NaN in both
        assert np.isnan(got).all() and np.isnan(ref).all()
        return
    assert np.abs(got - ref).max() < 1e-4


def test_pearson_cleans_nan_inf_and_flags_constant_columns():
    from oracle.ref_model import pearson_r
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(5)
    x = torch.randn(256, 64, generator=g)
    y = torch.randn(256, 64, generator=g) + x
    x[3, 5] = float("nan"); y[7, 9] = float("nan"); x[:, 11] = 2.5          # constant column -> r undefined
    r, _ = ops.pearson(x.cuda(), y.cuda())
    ref = pearson_r(x.numpy(), y.numpy())
    got = r.cpu().numpy()
    assert np.isnan(got[11]) and np.isnan(ref[11])
    keep = np.arange(64) != 11
    assert np.abs(got[keep] - ref[keep]).max() < 1e-4
    # +-inf become +-dtype-max like torch.nan_to_num: the moments must stay finite in fp64
    xb = x.bfloat16(); xb[0, 0] = float("inf"); yb = y.bfloat16()
    _, stats = ops.pearson(xb.cuda(), yb.cuda(), finalize=False)
    assert torch.isfinite(stats).all()


def test_pearson_row_shards_merge_to_single_pass():
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(9)
    x = torch.randn(1000, 520, generator=g).cuda(); y = (torch.randn(1000, 520, generator=g)).cuda() + 0.5 * x
    r_all, _ = ops.pearson(x, y)
    stats = torch.zeros((5, 520), device="cuda", dtype=torch.float64)
    for lo, hi in ((0, 137), (137, 600), (600, 1000)):          # uneven shards accumulate into the same moments
        ops.pearson(x[lo:hi].contiguous(), y[lo:hi].contiguous(), stats=stats, finalize=False)
    r_sh = ops.pearson_finalize(stats, 1000)
    assert (r_all - r_sh).abs().max().item() < 1e-6


def test_paint_labels_bit_exact():
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(2)
    P = 1000
    r = (torch.rand(P, generator=g) * 2 - 1)
    labels = torch.randint(0, P + 1, (37, 41, 29), generator=g, dtype=torch.int32)
    vol = ops.paint_labels(r.cuda(), labels.cuda(), square=True).cpu()
    r2 = r * r
    exp = torch.where(labels > 0, r2[(labels - 1).clamp(min=0).long()], torch.zeros(()))
    assert torch.equal(vol, exp)


def test_clip_adamw_matches_torch_over_steps():
    from phantom_vlb_b200 import lib as _l
    n = 4096 * 3 + 8
    g = torch.Generator().manual_seed(11)
    p0 = torch.randn(n, generator=g)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    p = p0.clone().cuda(); m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    pb = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    sumsq = torch.zeros(1, device="cuda")
    for step in range(1, 6):
        grad = torch.randn(n, generator=g) * (3.0 if step % 2 else 0.001)     # alternately clipped / not clipped
        ref.grad = grad.clone()
        torch.nn.utils.clip_grad_norm_([ref], 1.0)
        opt.step()
        gd = grad.cuda()
        sumsq.zero_()
        _l.call("vlb_grad_sumsq", _l.ptr(gd), n, _l.ptr(sumsq), _l.cur_stream())
        _l.call("vlb_adamw_step", _l.ptr(p), _l.ptr(gd), _l.ptr(m), _l.ptr(v), _l.ptr(pb), n, _l.ptr(sumsq), 1e-3, 0.9, 0.999,
                1e-8, 1e-2, step, 1.0, 1.0, _l.cur_stream())
        torch.cuda.synchronize()
        assert (p.cpu() - ref.detach()).abs().max().item() < 2e-6, "step %d" % step
    assert torch.equal(pb.cpu(), p.cpu().bfloat16())


def test_row_kernels_match_fp32_formulas():
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(4)
    M, H = 300, 4096
    x = torch.randn(M, H, generator=g).bfloat16().cuda()
    w = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16().cuda()
    rstd = torch.empty(M, device="cuda")
    y = ops.rmsnorm_fwd(x, w, 1e-5, rstd=rstd)
    xf = x.float()
    r = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)
    yr = (xf * r).bfloat16().float() * w.float()
    assert (y.float() - yr).abs().max().item() < 4e-2
    assert torch.allclose(rstd, r[:, 0], rtol=1e-4)
    # SwiGLU forward / backward
    I = 1024
    gu = torch.randn(M, 2 * I, generator=g).bfloat16().cuda()
    act = ops.swiglu_fwd(gu)
    gg, uu = gu.float()[:, :I], gu.float()[:, I:]
    def close(a, b):                                     # bf16 output rounding: 2^-8 relative
        return bool(((a.float() - b).abs() <= 8e-3 * b.abs() + 1e-2).all())
    assert close(act, torch.nn.functional.silu(gg) * uu)
    dact = torch.randn(M, I, generator=g).bfloat16().cuda()
    dgu = ops.swiglu_bwd(gu, dact)
    sg = torch.sigmoid(gg)
    dg_ref = dact.float() * uu * sg * (1 + gg * (1 - sg))
    du_ref = dact.float() * gg * sg
    assert close(dgu[:, :I], dg_ref) and close(dgu[:, I:], du_ref)
    # RoPE then inverse RoPE is the identity up to bf16 rounding (rotation is orthogonal)
    S, nh, D = 100, 3, 128
    inv = 1.0 / (1e6 ** (torch.arange(0, D, 2).float() / D))
    ang = torch.arange(S).float()[:, None] * inv[None]
    cos, sin = ang.cos().bfloat16().cuda(), ang.sin().bfloat16().cuda()          # [S, D/2] bf16 (model.rope_tables)
    q = torch.randn(3 * S, nh * D, generator=g).bfloat16().cuda()
    q0 = q.clone()
    qf = q0.float().view(3, S, nh, D)
    rot = torch.cat([-qf[..., D // 2:], qf[..., :D // 2]], -1)
    c2, s2 = torch.cat([cos, cos], 1).float(), torch.cat([sin, sin], 1).float()
    exp = qf * c2.view(1, S, 1, D) + rot * s2.view(1, S, 1, D)
    ops.rope_(q, cos, sin, S, nh, D)
    assert (q.float().view(3, S, nh, D) - exp).abs().max().item() < 4e-2
    ops.rope_(q, cos, sin, S, nh, D, inverse=True)
    assert (q.float() - q0.float()).abs().max().item() < 6e-2


def test_uint8_frames_normalised_in_gather_bit_exact():
    """f-1: raw uint8 frames + fused CLIP rescale/normalise give the same im2col matrix, bit for bit, as the reference
    route (CLIPImageProcessor arithmetic in fp32 on the host, fp32 pixels in the batch, .to(bf16) on the device)."""
    from phantom_vlb_b200 import ops
    g = torch.Generator().manual_seed(8)
    frames = torch.randint(0, 256, (3, 3, 56, 56), generator=g, dtype=torch.uint8)
    mean = torch.tensor(ops.CLIP_MEAN, dtype=torch.float32).view(1, 3, 1, 1)
    std = torch.tensor(ops.CLIP_STD, dtype=torch.float32).view(1, 3, 1, 1)
    ref_pixels = ((frames.float() * (1.0 / 255.0)) - mean) / std          # rescale then normalise, fp32 (image_processing_clip)
    kpad = (3 * 14 * 14 + 7) // 8 * 8
    a = ops.patch_im2col(ref_pixels.cuda().contiguous(), 14, kpad)
    b = ops.patch_im2col(frames.cuda().contiguous(), 14, kpad)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


@pytest.mark.parametrize("F,Hh,Ww,C", [(3, 24, 24, 256), (2, 13, 13, 64), (1, 1, 1, 8), (2, 1, 5, 16), (2, 5, 1, 16),
                                        (1, 2, 2, 24)])
def test_dwconv3x3_matches_torch_grouped_conv(F, Hh, Ww, C):
    """STC RegStage depthwise 3x3 (timm Bottleneck conv2 with group width 1, projector.py:153-161): NHWC rows, tap-major
    weights, zero padding. fp32 torch conv on the same bf16 inputs; the only difference allowed is the bf16 rounding of the
    output (and fp32 summation order)."""
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(F * Hh * Ww, C, device="cuda", generator=g).bfloat16()
    w = torch.randn(9, C, device="cuda", generator=g).bfloat16()
    y = ops.dwconv3x3(x, w, F, Hh, Ww)
    xt = x.float().view(F, Hh, Ww, C).permute(0, 3, 1, 2)
    wt = w.float().t().reshape(C, 1, 3, 3)
    ref = torch.nn.functional.conv2d(xt, wt, padding=1, groups=C).permute(0, 2, 3, 1).reshape(F * Hh * Ww, C)
    err = (y.float() - ref).abs()
    tol = ref.abs() * 2.0 ** -7 + 1e-5          # one bf16 ulp of the result + fp32 reassociation noise
    assert bool((err <= tol).all()), float((err - tol).max())


@pytest.mark.parametrize("F,HW,C", [(3, 576, 4096), (5, 169, 1024), (2, 1, 8), (2, 7, 264)])
def test_se_pool_and_scale_match_fp32(F, HW, C):
    """SE squeeze / excite of the STC RegStage bottleneck (timm SEModule: mean over H, W; x * gate) on NHWC rows."""
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(F * HW, C, device="cuda", generator=g).bfloat16()
    pool = ops.se_pool(x, F, HW)
    ref = x.float().view(F, HW, C).mean(1)
    assert float((pool.float() - ref).abs().max()) <= float(ref.abs().max()) * 2.0 ** -7 + 1e-6
    gate = torch.rand(F, C, device="cuda", generator=g).bfloat16()
    y = ops.se_scale(x, gate, F, HW)
    ref2 = (x.float().view(F, HW, C) * gate.float()[:, None, :]).reshape(F * HW, C)
    assert float((y.float() - ref2).abs().max()) <= float(ref2.abs().max()) * 2.0 ** -7
