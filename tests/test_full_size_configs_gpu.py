"""BASELINE configs[3] and configs[4] at their FULL sizes (the CPU oracle cannot run a 7B backbone, so these use the
size-independent checks the domain offers, on the same kernels and shapes bench.py / tools/bench_configs.py time):

  configs[4]  per-voxel Pearson r over a held-out season [5632, 250 000] (make_acc_brainmaps path, src/utils.py:101-106):
              r on sampled voxels vs the numpy fp64 oracle <= 1e-4 (north_star), row-sharded moments == one pass.
  configs[3]  frozen 7B backbone, 16-frame clips (T'=9, 1521 visual tokens, L_text=528), 250 000-voxel readout head
              (VLB_vllama2_friends_baseline.yaml:26-27): extract_features == the pooled vector of the full step
              (pooled = g1*u + b1*sw), readout on sampled voxels vs fp32 torch on the same operands, and the fused
              rank-B clip + AdamW update vs the closed-form AdamW step from a snapshot of (W, m, v).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_config4_pearson_full_size_matches_fp64_oracle():
    from oracle.ref_model import pearson_r
    from phantom_vlb_b200 import ops
    N, P = 5632, 250000
    g = torch.Generator(device="cuda").manual_seed(3)
    cols = torch.arange(0, P, P // 256, device="cuda")[:256]
    for dt in (torch.float32, torch.bfloat16):
        x = torch.randn((N, P), device="cuda", generator=g).to(dt)
        y = (0.5 * x.float() + torch.randn((N, P), device="cuda", generator=g)).to(dt)
        x[17, cols[3]] = float("nan")                 # the callback nan_to_num's both tensors (src/utils.py:101-102)
        r, stats = ops.pearson(x, y)
        # two uneven row shards accumulated into the same moments (the 8-GPU sharding of configs[4], one rank's view)
        st2 = torch.zeros_like(stats)
        ops.pearson(x[:2000], y[:2000], stats=st2, finalize=False)
        ops.pearson(x[2000:], y[2000:], stats=st2, finalize=False)
        r2 = ops.pearson_finalize(st2, N)
        torch.cuda.synchronize()
        r_ref = pearson_r(x[:, cols].float().cpu().numpy(), y[:, cols].float().cpu().numpy())
        err = float(np.abs(r[cols].double().cpu().numpy() - r_ref).max())
        err2 = float((r2 - r).abs().max())
        print("configs[4] Pearson [%d, %d] %s: max |r - fp64 oracle| = %.3g on %d sampled voxels, shard merge diff %.3g"
              % (N, P, str(dt).split(".")[-1], err, cols.numel(), err2))
        assert err <= 1e-4 and err2 <= 1e-6
        assert bool(torch.isfinite(r).all())
        del x, y


def test_config3_frozen_backbone_250k_voxel_head_full_size():
    from bench import host_batches
    from phantom_vlb_b200.engine import Engine
    from phantom_vlb_b200.model import ModelConfig, Trainables, frozen_random
    dev = torch.device("cuda", 0)
    P, B = 250000, 2
    mc = ModelConfig(num_frames=16, num_target=P, use_lora=False)
    assert (mc.t_out, mc.n_vis, mc.l_text) == (9, 1521, 528)
    T = Trainables(mc, dev)
    T.init_default(1234)
    with torch.no_grad():
        gg = torch.Generator(device=dev).manual_seed(11)
        for n in ("ln1_w", "ln1_b", "ln2_w", "ln2_b"):
            T.p(n).add_(0.1 * torch.randn(mc.hidden, device=dev, generator=gg))
    T.sync_working()
    eng = Engine(mc, frozen_random(mc, dev, 1234), T, device=dev)
    assert eng.fused_head_opt
    batch = {k: v.to(dev) for k, v in host_batches(mc, B, 1, seed=1234)[0].items()}
    # ---- (1) cached features == what the full step pools
    u, sw = eng.extract_features(batch)
    u, sw = u.clone(), sw.clone()
    loss_full, pred_full = eng.loss_and_grads(batch)
    pooled_full = eng.head_saved["pooled"].float().clone()
    loss_full, pred_full = loss_full.clone(), pred_full.clone()
    pooled_cached = T.w("ln1_w").float() * u + T.w("ln1_b").float() * sw[:, None]
    e_pool = float((pooled_cached - pooled_full).norm() / pooled_full.norm())
    loss_c, pred_c = eng.cached_loss_and_grads(u, sw, batch["timeseries"])
    e_loss = abs(float(loss_c[0]) - float(loss_full[0])) / abs(float(loss_full[0]))
    e_pred = float((pred_c.float() - pred_full.float()).norm() / pred_full.float().norm())
    print("configs[3] T=16 P=250k: pooled(cached) vs pooled(full) %.3g, loss %.3g, readout %.3g" % (e_pool, e_loss, e_pred))
    assert torch.isfinite(pred_full.float()).all() and e_pool < 5e-3 and e_loss < 1e-2 and e_pred < 1e-2
    # ---- (2) readout + fused optimiser at N=64 cached samples (the all-gathered global batch of an 8-GPU run)
    N = 64
    u = torch.randn((N, mc.hidden), device=dev, generator=gg)
    sw = torch.rand((N,), device=dev, generator=gg) + 0.5
    y = torch.randn((N, P), device=dev, generator=gg)
    idx = torch.randint(0, P, (512,), device=dev, generator=gg)
    eng.cached_loss_and_grads(u, sw, y)
    hs = eng.head_saved
    pred_ref = hs["hrf"].float() @ T.w("ridge_w")[idx].float().t() + T.w("ridge_b")[idx].float()
    err = float((hs["pred"][:, idx].float() - pred_ref).norm() / pred_ref.norm())
    w0, m0, v0 = T.p("ridge_w")[idx].clone(), T.view(T.m, "ridge_w")[idx].clone(), T.view(T.v, "ridge_w")[idx].clone()
    gw = hs["dpred"][:, idx].float().t() @ hs["hrf"].float() + 2 * mc.l2_lambda * T.w("ridge_w")[idx].float()
    lr = 1e-4
    eng.optimizer_step(lr)
    clip = min(1.0, 1.0 / (float(T.sumsq) ** 0.5 + 1e-6))
    t = T.step_count
    gw = gw * clip
    m1 = 0.9 * m0 + 0.1 * gw
    v1 = 0.999 * v0 + 0.001 * gw * gw
    w1 = w0 * (1 - lr * 1e-2) - (lr / (1 - 0.9 ** t)) * m1 / (v1.sqrt() / (1 - 0.999 ** t) ** 0.5 + 1e-8)
    uerr = float((T.p("ridge_w")[idx] - w1).abs().max() / (w1 - w0).abs().max())
    print("configs[3] head step at P=250k, N=64: readout rel err %.3g on 512 sampled voxels, AdamW update rel err %.3g, "
          "grad norm %.4g" % (err, uerr, float(T.sumsq) ** 0.5))
    assert err < 1e-2 and uerr < 1e-3
