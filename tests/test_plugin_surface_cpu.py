"""The reference-facing plugin surface checked on CPU: Hydra config keys, hook signatures, batch contract.

Reference: src/litmodule/videollama2_vlb_litmodule.py:126-157 (config keys, __post_init__), :259,:309 (step signatures
without batch_idx); src/datamodule/videollama2_vlb_datamodule.py:31-62 (config), :98-109 (item contract), :130-153 (split);
src/__init__.py:3-13 (re-exports).
"""
import dataclasses
import inspect
import os

import numpy as np
import torch

from oracle import indexing
from phantom_vlb_b200 import utils
from phantom_vlb_b200.datamodule import (SyntheticVLBDataset, VLBDataModule, VLBDataModuleConfig, season_files, split_train_val)
from phantom_vlb_b200.litmodule import VLBLitModule, VLBLitModuleConfig, model_config_for

LIT_KEYS = ["model_path", "freeze_backbone", "use_lora", "lora_r", "lora_alpha", "lora_dropout", "dropout_rate", "num_target",
            "l2_lambda", "lr", "betas", "eps", "weight_decay", "lr_scheduler_name", "last_epoch", "t_max"]
DATA_KEYS = ["lazyload_path", "subject", "seasons", "delay", "window", "random_state", "shuffle_val_data", "batch_size",
             "num_workers"]


def lit_config(**over):
    kw = dict(model_path="random:7b", freeze_backbone=False, use_lora=True, lora_r=16, lora_alpha=32, lora_dropout=0.1,
              dropout_rate=0.1, num_target=1000, l2_lambda=1e-3, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2,
              lr_scheduler_name="CosineAnnealingLR", last_epoch=-1, t_max=50000)
    kw.update(over)
    return VLBLitModuleConfig(**kw)


def test_litmodule_config_keys_and_post_init():
    assert [f.name for f in dataclasses.fields(VLBLitModuleConfig)] == LIT_KEYS
    cfg = lit_config()
    assert cfg.dtype == torch.bfloat16 and cfg.device_map == "auto"          # litmodule.py:152-157
    mc = model_config_for(cfg, num_frames=8)
    assert mc.num_target == 1000 and mc.lora_r == 16 and mc.use_lora
    # token budget of the spliced sequence (extractfeatures.py:198-212): always 2048
    for frames, l_text in ((8, 1204), (12, 866), (16, 528)):
        m = model_config_for(cfg, num_frames=frames)
        assert m.l_text == l_text == indexing.text_budget(frames)
        assert m.t_out * 169 + m.l_text - 1 == 2048


def test_step_hooks_have_the_reference_signatures():
    # Lightning calls training_step(batch) / validation_step(batch): the reference declares NO batch_idx (litmodule.py:259,309)
    assert list(inspect.signature(VLBLitModule.training_step).parameters) == ["self", "batch"]
    assert list(inspect.signature(VLBLitModule.validation_step).parameters) == ["self", "batch"]
    for name in ("configure_model", "configure_optimizers", "make_weight_mask", "forward", "on_save_checkpoint"):
        assert callable(getattr(VLBLitModule, name))
    for name in ("HRFConvolveLayer", "RidgeRegressionLayer", "get_hrf_weight", "LogValAccuracyCallback"):   # src/__init__.py
        assert hasattr(utils, name)
    cb = utils.LogValAccuracyCallback
    for hook in ("on_validation_epoch_start", "on_validation_batch_end", "on_validation_epoch_end"):       # utils.py:85-110
        assert callable(getattr(cb, hook))


def test_datamodule_config_keys_and_defaults():
    fields = dataclasses.fields(VLBDataModuleConfig)
    assert [f.name for f in fields] == DATA_KEYS
    d = {f.name: f.default for f in fields}
    assert d["batch_size"] == 1 and d["num_workers"] == 0                    # datamodule.py:40-41
    for name in ("train_dataloader", "val_dataloader"):
        assert callable(getattr(VLBDataModule, name))


def test_file_split_draws_one_validation_file(tmp_path):
    # datamodule.py:130-153: glob per season, ONE validation file from RandomState(random_state), the rest train
    for s in ("s1", "s2"):
        for k in range(3):
            (tmp_path / f"sub-01_{s}_chunk{k}.h5").write_bytes(b"")
    cfg = VLBDataModuleConfig(lazyload_path=str(tmp_path / "sub-01_s*_chunk*.h5"), subject="sub-01", seasons=["s1", "s2"],
                              delay=3, window=3, random_state=1234, shuffle_val_data=False)
    files = season_files(cfg)
    assert files == sorted(str(p) for p in tmp_path.iterdir())                # per season, sorted
    train, val = split_train_val(files, cfg.random_state)
    assert len(val) == 1 and len(train) == 5 and val[0] not in train
    for seed in (0, 1, 7, 1234, 99999):                                       # the SAME held-out file the reference draws
        assert split_train_val(files, seed)[1] == np.random.RandomState(seed).choice(files, 1).tolist()
    assert split_train_val(files, cfg.random_state) == (train, val)           # deterministic
    assert os.path.basename(val[0]).startswith("sub-01_")


def test_synthetic_items_obey_the_batch_contract():
    cfg = lit_config()
    mc = model_config_for(cfg, num_frames=8)
    ds = SyntheticVLBDataset(mc, length=4, seed=1234)
    a, b = ds[1], ds[1]
    for k in a:                                                               # deterministic per index
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k]))
    it = ds[2]
    assert it["timeseries"].dtype == torch.float32 and tuple(it["timeseries"].shape) == (1000,)
    assert it["vision"].dtype == torch.float32 and tuple(it["vision"].shape) == (8, 3, 336, 336)
    assert it["language"].dtype == torch.float32 and tuple(it["language"].shape) == (mc.l_text,)   # ids stored as floats
    assert it["padvals"].dtype == np.int64 and it["padvals"].shape == (3,)
    assert it["vis_weights"].dtype == np.float64 and it["vis_weights"].shape == (mc.t_out,)
    assert it["lang_weights"].dtype == np.float64 and it["lang_weights"].shape == (64,)
    pad_len, inst_len, dialog_len = (int(v) for v in it["padvals"])
    ids = it["language"].long()
    assert int(ids[0]) == 1 and int((ids == -201).sum()) == 1                 # BOS, exactly one <video> (constants.py:18)
    assert bool((ids[mc.l_text - pad_len:] == 0).all()) and bool((ids[:mc.l_text - pad_len] != 0).all())
    assert np.all(it["lang_weights"][dialog_len:] == 0)
    # the weight row the litmodule builds from this item is exactly 2048 long (litmodule.py:178-203)
    v = int((ids == -201).nonzero()[0])
    rest = mc.t_out * 169 + (2 + inst_len) + dialog_len + (4 + pad_len)
    assert rest <= 2048 and mc.l_text - v - 1 == (2 + inst_len) + dialog_len + 4 + pad_len
    # visual lag grid of lazyloading.py:108-115 at delay=3, window=3 (SURVEY 8a known answers)
    lags = 1.49 * (5.5 - np.arange(mc.t_out) * (3 / (mc.t_out - 1)))
    assert np.allclose(lags, [8.195, 7.0775, 5.96, 4.8425, 3.725])
    assert np.allclose(it["vis_weights"], [utils.get_hrf_weight(t) for t in lags])
