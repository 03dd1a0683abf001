"""Frozen-backbone baseline (reference freeze_backbone=True / use_lora=False, litmodule.py:86-90; BASELINE configs[3]):
head-only training step, the cached-feature reformulation and the voxel-sharded readout head, each
against the CPU fp32 oracle on identical weights.  Tolerances are the bf16 bounds of north_star (1e-2 loss/readout)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


def build_frozen(B, seed=1234, target_slice=None, **cfg_kw):
    from oracle.ref_model import RefVLB, round_weights_to_bf16, synthetic_batch, tiny_config
    from phantom_vlb_b200.engine import Engine
    from phantom_vlb_b200.model import ModelConfig, Trainables, frozen_from_oracle
    cfg_ref = tiny_config(**cfg_kw)
    ref = RefVLB(cfg_ref, seed=seed, lora_b_std=0.0)         # B = 0: the wrapped linears equal the frozen base linears
    for n, p in ref.named_parameters():                       # baseline experiment: only the brain head trains
        p.requires_grad_(n.startswith(("layer_norm1", "layer_norm2", "ridge")))
    with torch.no_grad():
        g = torch.Generator().manual_seed(seed + 7)
        for t in (ref.layer_norm1.weight, ref.layer_norm1.bias, ref.layer_norm2.weight, ref.layer_norm2.bias):
            t.add_(0.1 * torch.randn(cfg_ref.hidden, generator=g))
    round_weights_to_bf16(ref)
    batch = synthetic_batch(cfg_ref, B, seed=seed)
    cfg = ModelConfig.from_any(cfg_ref)
    cfg.use_lora = False
    W = frozen_from_oracle(ref, cfg, "cuda")
    T = Trainables(cfg, torch.device("cuda"), target_slice=target_slice)
    T.load_from_oracle(ref)
    return ref, batch, cfg, Engine(cfg, W, T)


def ref_grads(ref, batch):
    out = ref.step(batch)
    out["loss"].backward()
    g = dict(ridge_w=ref.ridge.weight.grad, ridge_b=ref.ridge.bias.grad, ln1_w=ref.layer_norm1.weight.grad,
             ln1_b=ref.layer_norm1.bias.grad, ln2_w=ref.layer_norm2.weight.grad, ln2_b=ref.layer_norm2.bias.grad)
    return out, g


def test_frozen_backbone_step_matches_oracle():
    ref, batch, cfg, eng = build_frozen(3)
    out, g = ref_grads(ref, batch)
    assert len(eng.T.spec) == 6 and eng.T.lora_layers == 0       # no adapter tensors exist in this mode
    loss3, pred = eng.loss_and_grads({k: v.cuda() for k, v in batch.items()})
    torch.cuda.synchronize()
    assert abs(float(loss3[0]) - float(out["loss"])) / abs(float(out["loss"])) < 1e-2
    assert rel(pred, out["pred"]) < 1e-2
    for k, v in g.items():
        assert rel(eng.T.g(k), v) < 6e-2, k


def test_cached_features_step_matches_oracle_and_full_step():
    ref, batch, cfg, eng = build_frozen(3)
    out, g = ref_grads(ref, batch)
    dev = {k: v.cuda() for k, v in batch.items()}
    u, sw = eng.extract_features(dev)
    torch.cuda.synchronize()
    # u / sw against the oracle: LN1 without affine, pooled with the HRF weights
    h = torch.nn.functional.layer_norm(out["hidden"].detach(), (cfg.hidden,))
    u_ref = torch.einsum("bse,bs->be", h, out["weight_mask"])
    assert rel(u, u_ref) < 2e-2
    assert torch.allclose(sw.cpu(), out["weight_mask"].sum(1), rtol=1e-5, atol=1e-6)
    loss3, pred = eng.cached_loss_and_grads(u, sw, dev["timeseries"])
    torch.cuda.synchronize()
    assert abs(float(loss3[0]) - float(out["loss"])) / abs(float(out["loss"])) < 1e-2
    assert rel(pred, out["pred"]) < 1e-2
    for k, v in g.items():
        assert rel(eng.T.g(k), v) < 6e-2, k
    # and an optimiser step from the cached path moves the same parameters as torch AdamW on the oracle
    eng.optimizer_step(1e-3)
    params = [p for _, p in ref.named_parameters() if p.requires_grad]
    torch.nn.utils.clip_grad_norm_(params, 1.0)
    before = ref.ridge.weight.detach().clone()
    torch.optim.AdamW(params, lr=1e-3, weight_decay=1e-2).step()
    torch.cuda.synchronize()
    d_ref = ref.ridge.weight.detach() - before
    d_gpu = eng.T.p("ridge_w").cpu() - before
    # the first AdamW step is lr * sign(g) (+ decay): compare where the oracle gradient is not within bf16 noise of zero
    big = g["ridge_w"].abs() > g["ridge_w"].abs().median()
    assert rel(d_gpu[big], d_ref[big]) < 2e-2


def test_voxel_sharded_head_equals_unsharded():
    """Two voxel shards run one after the other on one GPU with a recording shard_reduce reproduce the unsharded loss,
    gradients and clipped update (the collective itself is covered by tests/test_dp_gloo_cpu.py)."""
    from phantom_vlb_b200.dp import shard_columns
    ref, batch, cfg, full = build_frozen(2)
    dev = {k: v.cuda() for k, v in batch.items()}
    u, sw = full.extract_features(dev)
    loss_full, pred_full = full.cached_loss_and_grads(u, sw, dev["timeseries"])
    loss_full = loss_full.clone()
    g_full = {k: full.T.g(k).clone() for k in ("ridge_w", "ridge_b", "ln1_w", "ln2_w", "ln2_b")}
    P = cfg.num_target
    slices = [shard_columns(P, r, 2) for r in range(2)]
    assert slices[0][1] == slices[1][0] and slices[1][1] == P
    engs = [build_frozen(2, target_slice=s)[3] for s in slices]

    # pass 1 records every tensor each shard would contribute; pass 2 replays with the cross-shard sums
    class Rec:
        def __init__(self):
            self.log, self.replay, self.i = [], None, 0

        def __call__(self, t):
            if self.replay is None:
                self.log.append(t.clone())
            else:
                t.copy_(self.replay[self.i])
                self.i += 1

    recs = [Rec(), Rec()]
    for e, r, (lo, hi) in zip(engs, recs, slices):
        e.shard_reduce = r
        e.cached_loss_and_grads(u, sw, dev["timeseries"][:, lo:hi].contiguous())
    sums = [a + b for a, b in zip(recs[0].log, recs[1].log)]
    outs = []
    for e, r, (lo, hi) in zip(engs, recs, slices):
        r.replay, r.i = sums, 0
        outs.append(e.cached_loss_and_grads(u, sw, dev["timeseries"][:, lo:hi].contiguous()))
    torch.cuda.synchronize()
    for (l3, pred), e, (lo, hi) in zip(outs, engs, slices):
        assert abs(float(l3[0]) - float(loss_full[0])) < 1e-5 * abs(float(loss_full[0])) + 1e-6
        assert torch.equal(pred, pred_full[:, lo:hi])
        assert rel(e.T.g("ridge_w"), g_full["ridge_w"][lo:hi]) < 1e-5
        for k in ("ln1_w", "ln2_w", "ln2_b"):
            assert rel(e.T.g(k), g_full[k]) < 1e-3, k


@pytest.mark.parametrize("lr_scale", [1.0])
def test_fused_head_optimizer_matches_materialised_gradient(lr_scale):
    """Wide-readout path (vlb_head_gradnorm + vlb_head_adamw_fused): the clip norm from Gram matrices and the AdamW update
    in the gradient GEMM's epilogue must reproduce the materialised-gradient path (dW GEMM + vlb_grad_sumsq +
    vlb_adamw_step) over several steps, with and without the clip being active."""
    _, batch, cfg, e_ref = build_frozen(4)
    _, _, _, e_fus = build_frozen(4)
    e_fus.fused_head_opt = True
    assert not e_ref.fused_head_opt
    dev = {k: v.cuda() for k, v in batch.items()}
    u, sw = e_ref.extract_features(dev)
    g = torch.Generator(device="cuda").manual_seed(1)
    for step in range(4):
        y = dev["timeseries"] * (30.0 if step % 2 == 0 else 0.01) + torch.randn(dev["timeseries"].shape, device="cuda", generator=g)
        for e in (e_ref, e_fus):
            e.cached_loss_and_grads(u, sw, y)
            e.optimizer_step(1e-3 * lr_scale)
        torch.cuda.synchronize()
        assert abs(float(e_ref.T.sumsq) - float(e_fus.T.sumsq)) <= 1e-4 * float(e_ref.T.sumsq), "clip norm, step %d" % step
        for name in ("ridge_w", "ridge_b", "ln1_w", "ln2_b"):
            a, b = e_fus.T.p(name), e_ref.T.p(name)
            assert float((a - b).abs().max()) <= 2e-6 + 1e-5 * float(b.abs().max()), "%s step %d" % (name, step)
        assert torch.equal(e_fus.T.w("ridge_w"), e_fus.T.p("ridge_w").bfloat16())
    assert float(e_fus.T.g("ridge_w").abs().max()) == 0.0          # the ridge gradient was never written
