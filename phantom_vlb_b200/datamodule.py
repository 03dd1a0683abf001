"""Host-side mirror of src/datamodule/videollama2_vlb_datamodule.py (batch contract kept verbatim) plus the synthetic
"Friends-shaped" generator used by tests and benchmarks.

Batch contract (VLB_Dataset.__getitem__, datamodule.py:98-109, default collate):
  timeseries float32 [B,P] | vision float32 [B,T,3,336,336] | language float32 [B,L_text] (ids incl. -201, right 0-pad)
  padvals int64 [B,3] = (pad_len, inst_len, dialog_len) | vis_weights float64 [B,T'] | lang_weights float64 [B,64]
"""
import glob
import os
from dataclasses import dataclass

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset

try:
    from lightning import LightningDataModule
except Exception:                                   # Lightning is not installed in this image
    class LightningDataModule:
        def __init__(self):
            pass

from .model import VIDEO_TOKEN_ID, ModelConfig
from .utils import get_hrf_weight


def get_idx(ranges, val):
    for i, (min_val, max_val) in enumerate(ranges):
        if min_val <= val < max_val:
            return i
    return -1


@dataclass
class VLBDataModuleConfig:
    lazyload_path: str
    subject: str
    seasons: list[str]
    delay: int
    window: int
    random_state: int
    shuffle_val_data: bool
    batch_size: int = 1
    num_workers: int = 0


class VLB_Dataset(Dataset):
    """Lazy HDF5 reader with the reference's file layout (datamodule.py:66-109). Needs h5py (absent in this image)."""

    def __init__(self, ds_paths):
        import h5py
        self.ds_files, self.length, self.ranges = {}, 0, []
        for i, ds_path in enumerate(ds_paths):
            self.ds_files[i] = {"ds_file": h5py.File(ds_path, "r"), "idx_from": self.length}
            ds_length = int(np.array(self.ds_files[i]["ds_file"]["dset_len"])[0])
            self.ranges.append((self.length, self.length + ds_length))
            self.length += ds_length

    def __len__(self):
        return self.length

    def __getitem__(self, idx):
        i = get_idx(self.ranges, idx)
        ds_file = self.ds_files[i]["ds_file"]
        set_idx = idx - self.ds_files[i]["idx_from"]
        item = {}
        for mod in ["timeseries", "vision", "language"]:
            item[mod] = torch.from_numpy(np.array(ds_file[f"{set_idx}"][f"{set_idx}_{mod}"])).float()
        for mod in ["padvals", "vis_weights", "lang_weights"]:
            item[mod] = np.array(ds_file[f"{set_idx}"][f"{set_idx}_{mod}"])
        return item


def split_files(config: VLBDataModuleConfig):
    """File-level train/val split (datamodule.py:130-153): one validation file drawn with RandomState(random_state)."""
    f_list = []
    for s in config.seasons:
        f_list += sorted(glob.glob(config.lazyload_path.replace("$SCRATCH_PATH", os.environ.get("SCRATCH_PATH", ""))
                                   .replace("s*", f"{s}")))
    r = np.random.RandomState(config.random_state)
    val_file = r.choice(f_list, 1).tolist()
    train_files = [x for x in f_list if x not in val_file]
    return train_files, val_file


class VLBDatasets:
    def __init__(self, config: VLBDataModuleConfig):
        self.config = config
        train_files, val_file = split_files(config)
        self.dset_names = {"val_set": [os.path.basename(x) for x in val_file],
                           "train_set": [os.path.basename(x) for x in train_files]}
        self.val = VLB_Dataset(val_file)
        self.train = VLB_Dataset(train_files)


class SyntheticVLBDataset(Dataset):
    """Deterministic synthetic clips obeying the batch contract: N(0,1) pixels, token rows
    [BOS, sys..., <video>, 2+inst, dialog, 4, 0-pad], Glover-HRF weights of the reference's lag grid, N(0,1) BOLD."""

    def __init__(self, mc: ModelConfig, length=64, seed=1234, delay=3, window=3):
        self.mc, self.length, self.seed, self.delay, self.window = mc, length, seed, delay, window

    def __len__(self):
        return self.length

    def __getitem__(self, idx):
        mc = self.mc
        g = torch.Generator().manual_seed(self.seed * 1000003 + idx)
        L = mc.l_text
        inst_len = min(9, max(1, L // 16))
        max_dialog = min(mc.lang_w_len - 6, L // 4)
        dialog_len = int(torch.randint(2, max(3, max_dialog), (1,), generator=g))
        pad_len = int(torch.randint(0, max(1, L // 2 - dialog_len - inst_len - 8), (1,), generator=g))
        n_sys = L - (1 + 2 + inst_len + dialog_len + 4 + pad_len) - 1
        toks = torch.randint(3, mc.vocab, (L,), generator=g).float()
        toks[0] = 1
        toks[1 + n_sys] = VIDEO_TOKEN_ID
        if pad_len > 0:
            toks[L - pad_len:] = 0
        tp = mc.t_out
        # visual token lags 1.49*(abs_tr_delay - arange(0, window+step, step)) (lazyloading.py:108-115)
        lags = 1.49 * ((self.window - 1) + self.delay + 0.5 - np.arange(tp) * (self.window / max(tp - 1, 1)))
        vis_w = np.array([get_hrf_weight(t) for t in lags], dtype=np.float64)
        lang_w = np.zeros(mc.lang_w_len, dtype=np.float64)
        tl = 1.5 + 7.5 * torch.rand(dialog_len, generator=g).numpy()
        lang_w[:dialog_len] = [get_hrf_weight(t) for t in tl]
        return {"timeseries": torch.randn(mc.num_target, generator=g),
                "vision": torch.randn(mc.num_frames, 3, mc.image_size, mc.image_size, generator=g),
                "language": toks,
                "padvals": np.array([pad_len, inst_len, dialog_len], dtype=np.int64),
                "vis_weights": vis_w, "lang_weights": lang_w}


class VLBDataModule(LightningDataModule):
    def __init__(self, config: VLBDataModuleConfig, datasets=None) -> None:
        super().__init__()
        self.config = config
        self.datasets = datasets if datasets is not None else VLBDatasets(self.config)

    def x_dataloader(self, dataset, shuffle=True):
        if dataset is None:
            raise AttributeError
        return DataLoader(dataset=dataset, batch_size=self.config.batch_size, shuffle=shuffle,
                          num_workers=self.config.num_workers)

    def train_dataloader(self):
        return self.x_dataloader(dataset=self.datasets.train)

    def val_dataloader(self):
        return self.x_dataloader(dataset=self.datasets.val, shuffle=self.config.shuffle_val_data)
