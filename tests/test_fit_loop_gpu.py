"""The plugin driven the way lightning.Trainer.fit drives it (Lightning is not installed here, so the loop is replayed by
hand in Lightning's order): configure_model -> configure_optimizers -> per step, under bf16 autocast (trainer.precision
"bf16-mixed"), optimizer.step(closure) with closure = training_step(batch) -> zero_grad -> backward -> clip_grad_norm_(1.0)
(trainer.gradient_clip_val 1) -> scheduler.step. Plus: checkpoint round trip through the hooks, model_path = local
checkpoint directory, forward() with the reference's positional arguments, and the two exported layers called alone.
Reference: train.py:41-56, config/experiment/VLB_vllama2_friends_lora.yaml:42-51, litmodule.py:58-122, 229-256; src/utils.py:40-73."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(**over):
    from phantom_vlb_b200.litmodule import VLBLitModuleConfig
    kw = dict(model_path="random:tiny", freeze_backbone=False, use_lora=True, lora_r=16, lora_alpha=32, lora_dropout=0.0,
              dropout_rate=0.0, num_target=40, l2_lambda=1e-3, lr=1e-3, betas=[0.9, 0.999], eps=1e-8, weight_decay=1e-2,
              lr_scheduler_name="CosineAnnealingLR", last_epoch=-1, t_max=50)
    kw.update(over)
    return VLBLitModuleConfig(**kw)


def _module(**over):
    from phantom_vlb_b200.litmodule import VLBLitModule
    m = VLBLitModule(_cfg(**over))
    m.configure_model(seed=1234, lora_b_std=0.02)
    return m


def _batches(mc, n, B=2):
    from bench import collate
    from phantom_vlb_b200.datamodule import SyntheticVLBDataset
    ds = SyntheticVLBDataset(mc, length=n * B, seed=7)
    return [collate([ds[i * B + j] for j in range(B)]) for i in range(n)]


def _lightning_step(module, opt, sched, batch):
    def closure():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = module.training_step(batch)
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(module.parameters(), 1.0)          # Trainer(gradient_clip_val=1), before the optimiser
        return loss
    loss = opt.step(closure)
    sched.step()
    return float(loss)


def test_fit_loop_in_lightning_order_equals_the_engine_steps():
    from phantom_vlb_b200.engine import cosine_lr
    a, b = _module(), _module()
    assert [n for n, _ in a.named_parameters()] == ["flat"]
    (opt,), (sch,) = a.configure_optimizers()
    assert sch["interval"] == "step" and sch["frequency"] == 1
    batches = _batches(a.model_config, 3)
    for i, batch in enumerate(batches):
        la = _lightning_step(a, opt, sch["scheduler"], batch)
        dev = {k: v.cuda() for k, v in batch.items()}
        lb = float(b.engine.train_step(dev, cosine_lr(1e-3, i, 50))[0])
        assert abs(la - lb) <= 1e-5 * abs(lb) + 1e-7, (i, la, lb)
    torch.cuda.synchronize()
    # Lightning's clip followed by the fused clip: the second one sees a norm <= 1 and is a no-op
    d = (a.engine.T.param - b.engine.T.param).abs().max()
    moved = (b.engine.T.param - _module().engine.T.param).abs().max()
    print("fit loop vs engine after 3 steps: max |dparam| %.3g (parameters moved by up to %.3g)" % (float(d), float(moved)))
    assert float(d) <= 2e-6 and float(moved) > 1e-4
    assert abs(opt.param_groups[0]["lr"] - cosine_lr(1e-3, 3, 50)) < 1e-12


def test_external_write_to_flat_reaches_the_kernels():
    m = _module()
    batch = _batches(m.model_config, 1)[0]
    p0 = m.validation_step(batch)["brain_preds"].float().clone()
    with torch.no_grad():
        m.flat.mul_(0.5)                        # what a plain torch optimiser does: an in-place write to the parameter
    p1 = m.validation_step(batch)["brain_preds"].float()
    assert not torch.allclose(p0, p1)           # the bf16 working copy and the staged LoRA B were refreshed


def test_checkpoint_round_trip_through_the_lightning_hooks():
    a = _module()
    (opt,), (sch,) = a.configure_optimizers()
    batches = _batches(a.model_config, 3)
    for batch in batches[:2]:
        _lightning_step(a, opt, sch["scheduler"], batch)
    ckpt = {"state_dict": a.state_dict(), "optimizer_states": [opt.state_dict()], "lr_schedulers": [sch["scheduler"].state_dict()]}
    a.on_save_checkpoint(ckpt)
    assert "flat" not in ckpt["state_dict"] and "ridge_layer.linear.weight" in ckpt["state_dict"]
    assert any(k.endswith("self_attn.q_proj.lora_A.default.weight") for k in ckpt["state_dict"])
    # ---- resume into a fresh module, the order Lightning uses: on_load_checkpoint -> load_state_dict(strict) -> optimizers
    from phantom_vlb_b200.litmodule import VLBLitModule
    b = VLBLitModule(_cfg())
    b.on_load_checkpoint(ckpt)
    res = b.load_state_dict(ckpt["state_dict"], strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    (opt_b,), (sch_b,) = b.configure_optimizers()
    opt_b.load_state_dict(ckpt["optimizer_states"][0])
    sch_b["scheduler"].load_state_dict(ckpt["lr_schedulers"][0])
    assert b.engine.T.step_count == a.engine.T.step_count == 2
    assert torch.equal(b.engine.T.param, a.engine.T.param) and torch.equal(b.engine.T.m, a.engine.T.m)
    la = _lightning_step(a, opt, sch["scheduler"], batches[2])
    lb = _lightning_step(b, opt_b, sch_b["scheduler"], batches[2])
    torch.cuda.synchronize()
    assert abs(la - lb) <= 1e-6 * abs(la)
    assert float((a.engine.T.param - b.engine.T.param).abs().max()) <= 1e-6
    with pytest.raises(KeyError):
        b.load_state_dict({**ckpt["state_dict"], "bogus.weight": torch.zeros(1)}, strict=True)


def test_model_path_local_checkpoint_directory(tmp_path):
    """load_pretrained_vllama2 (litmodule.py:58-122): config.json + sharded safetensors under the VideoLLaMA2 key names, the
    CLIP tower in its own directory (config.mm_vision_tower, encoder.py:21-29)."""
    from safetensors.torch import save_file
    from oracle.ref_model import RefVLB, canonical_params, round_weights_to_bf16, tiny_config
    from phantom_vlb_b200.litmodule import VLBLitModule
    from phantom_vlb_b200.model import ModelConfig, _flatten, frozen_from_oracle
    rc = tiny_config()
    ref = round_weights_to_bf16(RefVLB(rc, seed=3, lora_b_std=0.0))
    sd = {k: v.detach().clone().contiguous() for k, v in canonical_params(ref).items()}
    sd["lm_head.weight"] = torch.zeros(rc.vocab, rc.hidden)                    # present in real checkpoints, never read
    clip_dir, ck = tmp_path / "clip-vit-tiny", tmp_path / "VideoLLaMA2-tiny"
    clip_dir.mkdir(); ck.mkdir()
    vt = "model.vision_tower.vision_tower."
    save_file({k[len(vt):]: v for k, v in sd.items() if k.startswith(vt)}, str(clip_dir / "model.safetensors"))
    json.dump({"vision_config": {"hidden_size": rc.vit_hidden, "num_hidden_layers": rc.vit_layers, "num_attention_heads": rc.vit_heads,
                                 "intermediate_size": rc.vit_mlp, "image_size": rc.image_size, "patch_size": rc.patch_size,
                                 "layer_norm_eps": 1e-5}}, open(clip_dir / "config.json", "w"))
    rest = sorted(k for k in sd if not k.startswith(vt))
    half = len(rest) // 2
    save_file({"base_model.model." + k: sd[k] for k in rest[:half]}, str(ck / "model-00001-of-00002.safetensors"))   # peft-prefixed shard
    save_file({k: sd[k] for k in rest[half:]}, str(ck / "model-00002-of-00002.safetensors"))
    json.dump({"hidden_size": rc.hidden, "intermediate_size": rc.mlp, "num_hidden_layers": rc.layers,
               "num_attention_heads": rc.heads, "num_key_value_heads": rc.kv_heads, "vocab_size": rc.vocab,
               "rms_norm_eps": rc.rms_eps, "rope_theta": rc.rope_theta, "tokenizer_model_max_length": rc.max_len,
               "mm_vision_tower": str(clip_dir), "mm_vision_select_layer": -2, "mm_projector_type": "stc_connector",
               "mm_hidden_size": rc.vit_hidden, "num_frames": 12}, open(ck / "config.json", "w"))
    m = VLBLitModule(_cfg(model_path=str(ck)))
    m.configure_model()
    mc = m.model_config
    assert (mc.hidden, mc.layers, mc.kv_heads, mc.vit_layers, mc.stc_depth, mc.max_len, mc.image_size) == \
           (rc.hidden, rc.layers, rc.kv_heads, rc.vit_layers, rc.stc_depth, rc.max_len, rc.image_size)
    want = frozen_from_oracle(ref, ModelConfig.from_any(rc), "cuda")
    flat = lambda W: [t for d in [W.vit, W.stc, W.llm] + W.vit_layers + W.llm_layers for t in _flatten(d)]
    got, exp = flat(m.engine.W), flat(want)
    assert len(got) == len(exp) and all(torch.equal(x, y) for x, y in zip(got, exp))
    batch = _batches(mc, 1)[0]
    assert torch.isfinite(m.validation_step(batch)["brain_preds"].float()).all()
    with pytest.raises(FileNotFoundError):
        VLBLitModule(_cfg(model_path="DAMO-NLP-SG/VideoLLaMA2-7B")).configure_model()


def test_forward_takes_the_reference_arguments():
    m = _module().eval()
    batch = _batches(m.model_config, 1)[0]
    out = m.validation_step(batch)
    x_video = [(batch["vision"][i].to(torch.bfloat16), "video") for i in range(batch["vision"].shape[0])]      # litmodule.py:267
    x_lang = batch["language"].long()
    wm = m.make_weight_mask(batch["padvals"], batch["vis_weights"], batch["lang_weights"], x_lang.shape[1], m.model_config.max_len)
    pred, l2 = m.forward(x_video, x_lang, wm, attention_mask=x_lang.ne(0).long())
    # validation_step feeds fp32 pixels, this call bf16-rounded ones (as the reference does for both): compare on equal inputs
    b2 = dict(batch, vision=batch["vision"].to(torch.bfloat16).float())
    ref = m.validation_step(b2)["brain_preds"]
    assert torch.equal(pred, ref)
    w = m.engine.T.w("ridge_w").float()
    assert abs(float(l2) - 1e-3 * float((w * w).sum())) <= 1e-3 * float(l2)
    assert out["brain_preds"].shape == pred.shape


def test_exported_layers_are_callable_and_differentiable():
    from phantom_vlb_b200.utils import HRFConvolveLayer, RidgeRegressionLayer
    g = torch.Generator(device="cuda").manual_seed(0)
    B, S, H, P = 3, 200, 512, 72
    emb = torch.randn(B, S, H, device="cuda", generator=g).bfloat16().requires_grad_()
    w = torch.rand(B, S, device="cuda", generator=g)
    w[:, :50] = 0
    w = w.requires_grad_()
    out = HRFConvolveLayer()(emb, w)
    ref = torch.einsum("bse,bs->be", emb.detach().float(), w.detach())
    assert out.shape == (B, H) and float((out.float() - ref).norm() / ref.norm()) < 5e-3
    gout = torch.randn(B, H, device="cuda", generator=g)
    out.backward(gout.to(out.dtype))
    go = gout.to(out.dtype).float()
    assert float((emb.grad.float() - w.detach()[:, :, None] * go[:, None, :]).norm()) / float((w.detach()[:, :, None] * go[:, None, :]).norm()) < 5e-3
    dw_ref = torch.einsum("bse,be->bs", emb.detach().float(), go)
    assert float((w.grad - dw_ref).norm() / dw_ref.norm()) < 1e-3
    # ridge: (output, l2_reg) like src/utils.py:66-73, and output only with add_regularization=False
    ridge = RidgeRegressionLayer(H, P, l2_lambda=0.01, dtype=torch.bfloat16).cuda()
    x = torch.randn(B, H, device="cuda", generator=g).bfloat16().requires_grad_()
    y, l2 = ridge(x)
    W, b = ridge.linear.weight.detach().float(), ridge.linear.bias.detach().float()
    y_ref = x.detach().float() @ W.t() + b
    assert float((y.float() - y_ref).norm() / y_ref.norm()) < 5e-3
    assert abs(float(l2) - 0.01 * float((W * W).sum())) <= 5e-3 * float(l2)
    gy = torch.randn(B, P, device="cuda", generator=g).bfloat16()
    (y.float() * gy.float()).sum().add(l2.float()).backward()
    assert float((x.grad.float() - gy.float() @ W).norm() / (gy.float() @ W).norm()) < 1e-2
    dW_ref = gy.float().t() @ x.detach().float() + 2 * 0.01 * W
    assert float((ridge.linear.weight.grad.float() - dW_ref).norm() / dW_ref.norm()) < 1e-2
    assert float((ridge.linear.bias.grad.float() - gy.float().sum(0)).norm() / gy.float().sum(0).norm()) < 1e-2
    assert torch.equal(ridge(x, add_regularization=False), y)
