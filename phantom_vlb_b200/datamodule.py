"""Data side of the plugin: the reference's batch contract, fed the B200 way.

Batch contract (reference VLB_Dataset.__getitem__, src/datamodule/videollama2_vlb_datamodule.py:98-109, default collate):
  timeseries float32 [B,P] | vision float32 [B,T,3,336,336] (CLIP-normalised) | language float32 [B,L_text] (ids incl. -201,
  right 0-pad) | padvals int64 [B,3] = (pad_len, inst_len, dialog_len) | vis_weights float64 [B,T'] | lang_weights float64 [B,64]
`vision` may instead be uint8 raw frames: the CLIP rescale + normalise then runs inside the patch gather (vlb_patch_im2col_u8),
4x less data from disk, in host memory and over PCIe.

What is here, and the reference lines each piece stands in for:
  align_run / build_season     the TR alignment of make_lazy_loading_dsets (src/preprocessing/videollama2_vlb_lazyloading.py:
                               97-137) as array slicing + one vectorised HRF call per run, producing stacked arrays in memory
                               (checked against the reference function executed over an h5py double: tests/golden/
                               lazyload_align.npz).
  ArrayVLBDataset              those arrays as a Dataset: __getitem__ is a row view (no HDF5, no per-sample file I/O).
  VLB_Dataset                  reader of the reference's lazy-load .h5 layout (one group per sample, `dset_len`), for data
                               already written by the reference; sample -> file lookup by bisection over cumulative lengths.
  DevicePrefetcher             pinned, double-buffered staging + H2D copies on a side stream, overlapped with the previous
                               training step; the authors call this transfer their bottleneck (train_run_lora.sh:34).
  VLBDataModule / VLBDatasets  the Lightning-facing objects train.py instantiates (train.py:46-50).
  SyntheticVLBDataset          deterministic "Friends-shaped" clips for tests and benchmarks.
"""
import bisect
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
from .utils import hrf_weights

TR = 1.49                                           # seconds per BOLD volume (lazyloading.py:99,112)
FLOAT_KEYS = ("timeseries", "vision", "language")   # returned as float32 tensors by the reference reader
ARRAY_KEYS = ("padvals", "vis_weights", "lang_weights")


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


# ------------------------------------------------------------------------------------------------ TR alignment (in memory)
def align_run(video, language, lang_onsets, masking, bold, delay=3, window=3):
    """One run (episode half): pair every complete `window`-TR input window with the BOLD volume `delay` TRs after it.

    video [n,T,...], language [n,L], lang_onsets [n,W] (token onset seconds, first dialog_len valid), masking [n,3] =
    (pad_len, inst_len, dialog_len), bold [m,P]. Follows lazyloading.py:97-137: inputs lose their first window-1 rows, BOLD its
    first window-1+delay rows, n_rows = the shorter of the two; visual token groups sit at lags
    TR*((window-1)+delay+0.5 - k*window/(T'-1)), k = 0..T'-1; row i's target time is TR*((window-1)+delay+0.5+i)."""
    f0, b0 = window - 1, window - 1 + delay
    n = min(bold.shape[0] - b0, video.shape[0] - f0, language.shape[0] - f0)
    n = max(n, 0)
    masking = np.asarray(masking)[f0:f0 + n]
    t_prime = video.shape[1] // 2 + 1
    step = window / (t_prime - 1)
    lead = (window - 1) + delay + 0.5
    vis_w = hrf_weights(TR * (lead - step * np.arange(t_prime)))
    onsets = np.array(lang_onsets[f0:f0 + n], dtype=np.float64)
    valid = np.arange(onsets.shape[1])[None, :] < masking[:, 2:3]
    t_mid = TR * (lead + np.arange(n))[:, None]
    lags = np.where(valid, t_mid - onsets, 1.0)
    lang_w = np.where(valid, hrf_weights(lags), onsets)            # entries past dialog_len keep their stored value (0 padding)
    return dict(timeseries=np.asarray(bold)[b0:b0 + n], vision=np.asarray(video)[f0:f0 + n],
                language=np.asarray(language)[f0:f0 + n], padvals=masking, lang_weights=lang_w,
                vis_weights=np.broadcast_to(vis_w, (n, t_prime)).copy())


def build_season(runs, delay=3, window=3):
    """runs: iterable of dicts with keys video / language / lang_onsets / masking / bold -> one dict of stacked arrays."""
    parts = [align_run(r["video"], r["language"], r["lang_onsets"], r["masking"], r["bold"], delay, window) for r in runs]
    return {k: np.concatenate([p[k] for p in parts], 0) for k in parts[0]}


class ArrayVLBDataset(Dataset):
    """A season held as stacked tensors (optionally pinned): row i of every array is sample i. uint8 `vision` stays uint8."""

    def __init__(self, arrays, pin=False):
        self.t = {}
        for k, v in arrays.items():
            t = v if isinstance(v, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(v))
            if k in FLOAT_KEYS and t.dtype != torch.uint8:
                t = t.float()
            self.t[k] = t.pin_memory() if pin and torch.cuda.is_available() else t
        self.n = len(self.t["timeseries"])

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return {k: v[i] for k, v in self.t.items()}

    def batch(self, indices):
        """Gather a batch without going through a DataLoader (one index_select per key)."""
        idx = torch.as_tensor(indices, dtype=torch.long)
        return {k: v.index_select(0, idx) for k, v in self.t.items()}


# ------------------------------------------------------------------------------------------------ reader of the reference's files
def _open_h5(path):
    try:
        import h5py
    except ImportError as e:
        raise ImportError("reading the reference's lazy-load files needs h5py (pass opener= or use ArrayVLBDataset)") from e
    return h5py.File(path, "r")


class VLB_Dataset(Dataset):
    """Samples stored the way lazyloading.py:141-160 writes them: file[str(i)][f"{i}_{key}"], file["dset_len"] = [count]."""

    def __init__(self, ds_paths, opener=None):
        opener = opener or _open_h5
        self.files = [opener(p) for p in ds_paths]
        counts = [int(np.asarray(f["dset_len"])[0]) for f in self.files]
        self.ends = np.cumsum(counts).tolist()                     # sample i lives in the first file whose end exceeds i

    def __len__(self):
        return self.ends[-1] if self.ends else 0

    def locate(self, idx):
        if not 0 <= idx < len(self):
            raise IndexError(idx)
        f = bisect.bisect_right(self.ends, idx)
        return f, idx - (self.ends[f - 1] if f else 0)

    def __getitem__(self, idx):
        f, j = self.locate(int(idx))
        grp = self.files[f][str(j)]
        item = {}
        for k in FLOAT_KEYS:
            a = np.asarray(grp["%d_%s" % (j, k)])
            item[k] = torch.from_numpy(a) if a.dtype == np.uint8 else torch.from_numpy(a).float()
        for k in ARRAY_KEYS:
            item[k] = np.asarray(grp["%d_%s" % (j, k)])
        return item


def season_files(config: VLBDataModuleConfig):
    """Lazy-load files of every configured season, in the reference's order (per season, sorted; datamodule.py:135-139)."""
    base = config.lazyload_path.replace("$SCRATCH_PATH", os.environ.get("SCRATCH_PATH", ""))
    files = []
    for season in config.seasons:
        files.extend(sorted(glob.glob(base.replace("s*", str(season)))))
    return files


def split_train_val(files, random_state):
    """Hold out ONE file for validation, the one the reference draws: RandomState(random_state).choice(files, 1)
    (datamodule.py:141-145) -- same seed, same held-out file, so runs stay comparable."""
    if not files:
        raise FileNotFoundError("no lazy-load files matched lazyload_path / seasons")
    held = int(np.random.RandomState(random_state).choice(len(files), 1)[0])
    return [f for i, f in enumerate(files) if f != files[held]], [files[held]]


class VLBDatasets:
    def __init__(self, config: VLBDataModuleConfig, opener=None, files=None):
        self.config = config
        train_files, val_files = split_train_val(files if files is not None else season_files(config), config.random_state)
        self.dset_names = {"val_set": [os.path.basename(x) for x in val_files],
                           "train_set": [os.path.basename(x) for x in train_files]}
        self.val = VLB_Dataset(val_files, opener)
        self.train = VLB_Dataset(train_files, opener)
        self.test = None


# ------------------------------------------------------------------------------------------------ host -> device staging
def collate_contract(items):
    """Default-collate equivalent for the contract (numpy entries become tensors, dtypes preserved)."""
    out = {}
    for k in items[0]:
        out[k] = torch.stack([v if isinstance(v, torch.Tensor) else torch.from_numpy(np.asarray(v)) for v in (it[k] for it in items)], 0)
    return out


class DevicePrefetcher:
    """Iterates device-resident batches. Each host batch is copied into one of `depth` pinned staging sets and sent with
    non_blocking H2D copies on a private stream while the consumer still computes on the previous batch; the consumer's
    stream waits on the copy's event only when it takes the batch. The device buffers of staging set k are overwritten only
    after the consumer's kernels finished with them: an event recorded on the compute stream when the NEXT batch is
    requested, which the copy stream waits on (the host never blocks on the GPU)."""

    def __init__(self, batches, device, depth=2):
        self.batches, self.device, self.depth = batches, torch.device(device), depth
        self.cuda = self.device.type == "cuda"
        self.stream = torch.cuda.Stream(self.device) if self.cuda else None
        self.slots = [dict(host={}, dev={}, ready=None, free=None) for _ in range(depth)]
        self.h2d_bytes = 0

    def _stage(self, slot, batch):
        if not self.cuda:
            slot["dev"] = {k: (torch.from_numpy(np.asarray(v)) if not isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
            return
        if slot["ready"] is not None:
            slot["ready"].synchronize()                # the previous H2D out of this pinned set has landed (long ago)
        if slot["free"] is not None:
            self.stream.wait_event(slot["free"])       # device side: copies wait until the consumer's kernels left this set
        with torch.cuda.stream(self.stream):
            for k, v in batch.items():
                v = v if isinstance(v, torch.Tensor) else torch.from_numpy(np.asarray(v))
                h = slot["host"].get(k)
                if h is None or h.shape != v.shape or h.dtype != v.dtype:
                    h = torch.empty(v.shape, dtype=v.dtype).pin_memory()
                    slot["host"][k] = h
                    slot["dev"][k] = torch.empty(v.shape, dtype=v.dtype, device=self.device)
                h.copy_(v)
                slot["dev"][k].copy_(h, non_blocking=True)
                self.h2d_bytes += h.numel() * h.element_size()
            slot["ready"] = torch.cuda.Event()
            slot["ready"].record(self.stream)

    def __iter__(self):
        it = iter(self.batches)
        pending = []
        for i in range(self.depth - 1):
            b = next(it, None)
            if b is None:
                break
            self._stage(self.slots[i % self.depth], b)
            pending.append(i % self.depth)
        n = len(pending)
        while pending:
            b = next(it, None)
            if b is not None:
                s = n % self.depth
                self._stage(self.slots[s], b)
                pending.append(s)
                n += 1
            s = pending.pop(0)
            slot = self.slots[s]
            if self.cuda:
                torch.cuda.current_stream(self.device).wait_event(slot["ready"])
            yield slot["dev"]
            if self.cuda:                                  # consumer came back for the next batch: this one is finished
                slot["free"] = torch.cuda.Event()
                slot["free"].record(torch.cuda.current_stream(self.device))

    def __len__(self):
        return len(self.batches)


# ------------------------------------------------------------------------------------------------ synthetic clips
class SyntheticVLBDataset(Dataset):
    """Deterministic synthetic clips obeying the batch contract: N(0,1) pixels (or uint8 frames), token rows
    [BOS, sys..., <video>, 2+inst, dialog, 4, 0-pad], Glover-HRF weights of the reference's lag grid, N(0,1) BOLD."""

    def __init__(self, mc: ModelConfig, length=64, seed=1234, delay=3, window=3, uint8_frames=False):
        self.mc, self.length, self.seed, self.delay, self.window, self.u8 = mc, length, seed, delay, window, uint8_frames
        tp = mc.t_out
        lead = (window - 1) + delay + 0.5
        self.vis_w = hrf_weights(TR * (lead - np.arange(tp) * (window / max(tp - 1, 1))))

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
        lang_w = np.zeros(mc.lang_w_len, dtype=np.float64)
        lang_w[:dialog_len] = hrf_weights(1.5 + 7.5 * torch.rand(dialog_len, generator=g).numpy())
        shape = (mc.num_frames, 3, mc.image_size, mc.image_size)
        vision = (torch.randint(0, 256, shape, generator=g, dtype=torch.uint8) if self.u8 else torch.randn(shape, generator=g))
        return {"timeseries": torch.randn(mc.num_target, generator=g), "vision": vision, "language": toks,
                "padvals": np.array([pad_len, inst_len, dialog_len], dtype=np.int64),
                "vis_weights": self.vis_w.copy(), "lang_weights": lang_w}


# ------------------------------------------------------------------------------------------------ Lightning-facing module
class VLBDataModule(LightningDataModule):
    """`datasets` may be injected (ArrayVLBDataset / SyntheticVLBDataset holders); by default the reference's lazy-load files
    are opened. `train_dataloader` / `val_dataloader` are what Lightning calls; `device_batches` is the overlapped feed."""

    def __init__(self, config: VLBDataModuleConfig, datasets=None) -> None:
        super().__init__()
        self.config = config
        self.datasets = datasets if datasets is not None else VLBDatasets(self.config)

    def _loader(self, dataset, shuffle):
        if dataset is None:
            raise AttributeError("no dataset for this phase")
        return DataLoader(dataset, batch_size=self.config.batch_size, shuffle=shuffle, num_workers=self.config.num_workers,
                          collate_fn=collate_contract, pin_memory=torch.cuda.is_available())

    def train_dataloader(self):
        return self._loader(self.datasets.train, True)

    def val_dataloader(self):
        return self._loader(self.datasets.val, self.config.shuffle_val_data)

    def device_batches(self, phase="train", device="cuda", depth=2):
        loader = self.train_dataloader() if phase == "train" else self.val_dataloader()
        return DevicePrefetcher(loader, device, depth)
