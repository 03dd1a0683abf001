"""The drop-in boundary the way train.py reaches it (train.py:32-56): Hydra resolves the `_target_` dotted paths of the
experiment YAML and instantiates config objects from the YAML's own keys. Hydra / OmegaConf are not installed here, so a
20-line `instantiate` does what they do (import the target, recurse into nested `_target_`s, resolve ${...}); the YAML
sections are copies of the reference's files as parsed (tests/golden/*.yaml)."""
import importlib
import os
import re

import numpy as np
import pytest
import torch
import yaml

from oracle import fake_h5py

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BASE = {"subject": "sub-01", "random_state": 1234}             # config/base.yaml:1-2 + the CLI's subject=sub-01


def _resolve(v):
    if isinstance(v, str):
        v = re.sub(r"\$\{(\w+)\}", lambda m: str(BASE[m.group(1)]), v)
        if re.fullmatch(r"[-+]?\d+", v):
            return int(v)
        if re.fullmatch(r"[-+]?(\d+\.?\d*|\.\d+)([eE][-+]?\d+)?", v):      # OmegaConf reads 1e-4 as a float, PyYAML as a str
            return float(v)
    return v


def instantiate(node, **extra):
    if isinstance(node, dict) and "_target_" in node:
        mod, _, name = node["_target_"].rpartition(".")
        target = getattr(importlib.import_module(mod), name)
        kw = {k: instantiate(v) for k, v in node.items() if k != "_target_"}
        kw.update(extra)
        return target(**kw)
    if isinstance(node, dict):
        return {k: instantiate(v) for k, v in node.items()}
    if isinstance(node, list):
        return [instantiate(v) for v in node]
    return _resolve(node)


@pytest.mark.parametrize("name", ["VLB_vllama2_friends_lora", "VLB_vllama2_friends_baseline"])
def test_yaml_targets_resolve_to_this_implementation(name, tmp_path, monkeypatch):
    import phantom_vlb_b200.datamodule as dm_mod
    import phantom_vlb_b200.litmodule as lit_mod
    cfg = yaml.safe_load(open(os.path.join(G, name + ".yaml")))
    assert cfg["litmodule"]["_target_"] == "src.litmodule.VLBLitModule"
    assert cfg["datamodule"]["_target_"] == "src.datamodule.VLBDataModule"
    # ---- litmodule: every YAML key is a constructor key, values arrive unmodified
    lit = instantiate(cfg["litmodule"])
    assert type(lit) is lit_mod.VLBLitModule and type(lit.config) is lit_mod.VLBLitModuleConfig
    for k, v in cfg["litmodule"]["config"].items():
        if k != "_target_":
            assert getattr(lit.config, k) == _resolve(v) or getattr(lit.config, k) == v, k
    assert lit.config.dtype == torch.bfloat16 and lit.config.device_map == "auto"
    assert lit.config.model_path == "DAMO-NLP-SG/VideoLLaMA2-7B"
    with pytest.raises(FileNotFoundError):                      # a hub id cannot be resolved offline: loud, not random weights
        lit_mod.model_config_for(lit.config)
    # ---- datamodule: lazy-load files found through $SCRATCH_PATH + the season glob, one held out (datamodule.py:130-153)
    fake_h5py.reset()
    monkeypatch.setenv("SCRATCH_PATH", str(tmp_path))
    monkeypatch.setattr(dm_mod, "_open_h5", lambda p: fake_h5py.File(p, "r"))
    pattern = _resolve(cfg["datamodule"]["config"]["lazyload_path"]).replace("$SCRATCH_PATH", str(tmp_path))
    for season in cfg["datamodule"]["config"]["seasons"]:
        for k in range(2):
            path = pattern.replace("s*", season).replace("_n*", "_n%d" % k).replace("*", "chunk%d" % k)
            open(path, "wb").close()
            with fake_h5py.File(path, "w") as f:
                g = f.create_group("0")
                for key, val in (("timeseries", np.zeros(4, np.float32)), ("vision", np.zeros((2, 3, 2, 2), np.float32)),
                                 ("language", np.zeros(5, np.float32)), ("padvals", np.zeros(3, np.int64)),
                                 ("vis_weights", np.zeros(2)), ("lang_weights", np.zeros(8))):
                    g.create_dataset("0_" + key, data=val)
                f.create_dataset("dset_len", data=[1])
    import glob
    f_list = []                                                 # the reference's own listing (datamodule.py:135-139)
    for season in cfg["datamodule"]["config"]["seasons"]:
        f_list += sorted(glob.glob(pattern.replace("s*", season)))
    n_files = len(f_list)       # NB: the LoRA yaml's pattern has no "s*", so every season lists every file (duplicates kept)
    val_ref = np.random.RandomState(BASE["random_state"]).choice(f_list, 1).tolist()
    data = instantiate(cfg["datamodule"])
    assert type(data) is dm_mod.VLBDataModule and type(data.config) is dm_mod.VLBDataModuleConfig
    names = data.datasets.dset_names                            # train.py:50 logs these
    n_train = len([x for x in f_list if x not in val_ref])
    assert names["val_set"] == [os.path.basename(val_ref[0])] and len(names["train_set"]) == n_train
    assert len(data.datasets.train) == n_train and len(data.datasets.val) == 1
    assert data.config.batch_size == cfg["datamodule"]["config"]["batch_size"]
    fake_h5py.reset()


def test_src_package_reexports_the_reference_names():
    import src
    import src.datamodule
    import src.litmodule
    from phantom_vlb_b200 import datamodule, litmodule, utils
    assert src.__all__ == ["HRFConvolveLayer", "RidgeRegressionLayer", "get_hrf_weight", "LogValAccuracyCallback"]   # src/__init__.py:10-13
    for n in src.__all__:
        assert getattr(src, n) is getattr(utils, n)
    assert src.litmodule.VLBLitModule is litmodule.VLBLitModule and src.litmodule.VLBLitModuleConfig is litmodule.VLBLitModuleConfig
    for n in ("VLBDataModule", "VLBDataModuleConfig", "VLB_Dataset", "VLBDatasets"):
        assert getattr(src.datamodule, n) is getattr(datamodule, n)


def test_trainer_section_is_what_the_fit_loop_test_replays():
    t = yaml.safe_load(open(os.path.join(G, "VLB_vllama2_friends_lora.yaml")))["trainer"]
    assert t["precision"] == "bf16-mixed" and t["gradient_clip_val"] == 1       # replayed in tests/test_fit_loop_gpu.py
