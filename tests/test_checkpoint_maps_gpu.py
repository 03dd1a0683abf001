"""Adapter-only checkpoints (train.py:58-60 TODO), the validation callback's metrics.csv and the r^2 brain maps
(make_acc_brainmaps.py:33-58), end to end on the tiny model through the public plugin classes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make_module(P=40, seed=1234, lora_b_std=0.02):
    from phantom_vlb_b200.litmodule import VLBLitModule, VLBLitModuleConfig
    cfg = VLBLitModuleConfig(model_path="random:tiny", freeze_backbone=False, use_lora=True, lora_r=16, lora_alpha=32,
                             lora_dropout=0.0, dropout_rate=0.0, num_target=P, l2_lambda=1e-3, lr=1e-3, betas=[0.9, 0.999],
                             eps=1e-8, weight_decay=1e-2, lr_scheduler_name="CosineAnnealingLR", last_epoch=-1, t_max=100)
    m = VLBLitModule(cfg)
    m.configure_model(device="cuda", seed=seed, lora_b_std=lora_b_std)
    return m


def tiny_batches(m, B, n, seed=0):
    from bench import collate
    from phantom_vlb_b200.datamodule import SyntheticVLBDataset
    ds = SyntheticVLBDataset(m.model_config, length=B * n, seed=seed)
    return [collate([ds[i * B + j] for j in range(B)]) for i in range(n)]


def test_adapter_checkpoint_round_trip(tmp_path):
    m = make_module()
    (opt,), _ = m.configure_optimizers()
    batches = tiny_batches(m, 2, 2)
    for b in batches:                                    # two real optimiser steps so A, B, head and Adam state are non-trivial
        loss = m.training_step(b)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
    sd = m.adapter_state_dict()
    L, r, mc = m.model_config.layers, m.model_config.lora_r, m.model_config
    assert len(sd) == L * 14 + 6                         # 7 projections x (A, B) per layer + ridge W/b + LN1/LN2
    k = "nnmodule.base_model.model.model.layers.1.mlp.down_proj.lora_B.default.weight"
    assert sd[k].shape == (mc.hidden, r) and sd[k].abs().sum() > 0
    assert sd["nnmodule.base_model.model.model.layers.0.self_attn.k_proj.lora_A.default.weight"].shape == (r, mc.hidden)
    assert sd["ridge_layer.linear.weight"].shape == (40, mc.hidden)
    nbytes = sum(v.numel() * 4 for v in sd.values())
    assert nbytes == sum(int(np.prod(s)) for _, s in m.engine.T.spec) * 4      # nothing but the trainables is stored
    path = str(tmp_path / "adapters.pt")
    m.save_adapters(path)
    ck = {}
    m.on_save_checkpoint(ck)                             # the Lightning hook stores the same adapter-only dict
    assert set(ck["state_dict"]) == set(sd)
    val = tiny_batches(m, 2, 1, seed=9)[0]
    ref_out = m.validation_step(val)
    m2 = make_module(seed=77, lora_b_std=0.0)            # different init: everything must come from the file
    m2.engine.W = m.engine.W                             # same frozen backbone (the checkpoint does not carry it)
    m2.load_adapters(path)
    out2 = m2.validation_step(val)
    assert torch.equal(out2["brain_preds"], ref_out["brain_preds"])
    assert float(out2["loss"]) == float(ref_out["loss"])
    assert torch.equal(m2.engine.T.m, m.engine.T.m) and m2.engine.T.step_count == 2
    sd_bad = dict(sd)
    sd_bad.pop(k)
    with pytest.raises(KeyError):
        m2.load_adapter_state_dict(sd_bad)


def test_validation_callback_writes_metrics_and_maps(tmp_path):
    from oracle.ref_model import pearson_r
    from phantom_vlb_b200.postprocessing import make_brainmaps, read_validation_rois
    from phantom_vlb_b200.utils import LogValAccuracyCallback
    m = make_module()
    cb = LogValAccuracyCallback(csv_dir=str(tmp_path))

    class Trainer:
        current_epoch, global_step = 0, 7

    preds, vals = [], []
    for epoch in range(2):
        Trainer.current_epoch = epoch
        cb.on_validation_epoch_start(Trainer, m)
        for i, b in enumerate(tiny_batches(m, 3, 4, seed=20 + epoch)):
            out = m.validation_step(b)
            cb.on_validation_batch_end(Trainer, m, out, b, i)
            if epoch == 1:
                preds.append(out["brain_preds"].float().cpu().numpy()); vals.append(out["brain_vals"].float().cpu().numpy())
        cb.on_validation_epoch_end(Trainer, m)
    rois = read_validation_rois(str(tmp_path))
    assert rois.shape == (2, 40)
    ref = pearson_r(np.concatenate(preds), np.concatenate(vals))           # fp64 oracle on the same bf16-rounded tensors
    assert np.abs(rois[1] - ref).max() < 1e-4
    assert np.abs(cb.last_correlations - ref).max() < 1e-4
    labels = np.random.RandomState(1).randint(0, 41, size=(9, 10, 11)).astype(np.int32)
    vols = make_brainmaps(str(tmp_path), labels, str(tmp_path / "acc"), export_nii=True)
    assert len(vols) == 2 and (tmp_path / "acc_val-1.nii").exists() and (tmp_path / "acc_val-0.npy").exists()
    r2 = (rois[1].astype(np.float32)) ** 2
    exp = np.where(labels > 0, r2[np.clip(labels - 1, 0, None)], 0.0).astype(np.float32)
    assert np.array_equal(vols[1], exp)
