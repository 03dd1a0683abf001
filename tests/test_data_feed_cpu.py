"""SURVEY 8 f-1 on CPU: the in-memory TR alignment against the reference's make_lazy_loading_dsets (executed over an h5py
double by oracle/make_goldens.py -> tests/golden/lazyload_align.npz), the reader of the reference's lazy-load file layout
(fake h5py), Glover HRF weights (product formulation vs the step-by-step nilearn restatement), and the prefetcher's order."""
import os

import numpy as np
import torch

from oracle import fake_h5py, glover, indexing
from phantom_vlb_b200 import utils
from phantom_vlb_b200.datamodule import (ArrayVLBDataset, DevicePrefetcher, VLB_Dataset, VLBDataModule, VLBDataModuleConfig,
                                         VLBDatasets, align_run, build_season, collate_contract)

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_glover_weights_match_the_nilearn_restatement_on_the_reference_lag_grids():
    # visual token lags at delay=3, window=3 for T = 8 / 12 / 16 frames; plus language-like lags
    for T in (8, 12, 16):
        lags = indexing.visual_token_lags(T)
        got = utils.hrf_weights(lags)
        exp = np.array([glover.get_hrf_weight(t) for t in lags])
        assert got.dtype == np.float64 and np.allclose(got, exp, rtol=0, atol=1e-13)
        assert (got > 0.1).all() and got.max() < 0.3          # 3.7 .. 8.2 s after a 1 s event: around the HRF peak
    rng = np.random.RandomState(0)
    for t in rng.uniform(0.05, 40.0, 40):
        assert abs(utils.get_hrf_weight(t) - glover.get_hrf_weight(t)) < 1e-13
    assert utils.get_hrf_weight(4.5) > utils.get_hrf_weight(2.0) > utils.get_hrf_weight(1.0) > 0       # rise
    assert utils.get_hrf_weight(15.0) < 0 and abs(utils.get_hrf_weight(33.5)) < 1e-12                   # undershoot, 32 s support
    assert utils.get_hrf_weight(0.0) == 0.0 and utils.get_hrf_weight(-2.0) == 0.0


def _golden_runs(z):
    eps = sorted({k.split(".")[1] for k in z.files if k.startswith("in.")})
    return eps, {ep: dict(video=z["in.%s.video_features" % ep], language=z["in.%s.transcript_features" % ep],
                          lang_onsets=z["in.%s.transcript_onsets" % ep], masking=z["in.%s.masking_params" % ep],
                          bold=z["in.%s.bold" % ep]) for ep in eps}


def test_in_memory_alignment_equals_the_reference_lazyload_writer():
    z = np.load(os.path.join(G, "lazyload_align.npz"))
    window, delay = int(z["window"]), int(z["delay"])
    eps, runs = _golden_runs(z)
    n_files = len(z["files"])
    chunk = np.floor(np.arange(len(eps)) / (len(eps) / n_files)).astype(int)            # lazyloading.py:89-91
    for fi in range(n_files):
        season = build_season([runs[ep] for ep in np.array(eps)[chunk == fi]], delay, window)
        assert len(season["timeseries"]) == int(z["f%d.len" % fi])
        for k in ("timeseries", "vision", "language", "padvals"):
            assert np.array_equal(season[k], z["f%d.%s" % (fi, k)]), k                    # row pairing: exact
        for k in ("vis_weights", "lang_weights"):
            assert np.allclose(season[k], z["f%d.%s" % (fi, k)], rtol=0, atol=1e-13), k
    # a run whose BOLD is too short yields no samples instead of an error
    r = runs[eps[0]]
    assert len(align_run(r["video"], r["language"], r["lang_onsets"], r["masking"], r["bold"][:4], delay, window)["timeseries"]) == 0


def _write_reference_layout(path, arrays):
    """The layout lazyloading.py:141-164 writes: file[str(i)][f"{i}_{key}"] and file["dset_len"] = [n]."""
    with fake_h5py.File(path, "w") as f:
        n = len(arrays["timeseries"])
        for i in range(n):
            g = f.create_group("%d" % i)
            for k, v in arrays.items():
                g.create_dataset("%d_%s" % (i, k), data=v[i])
        f.create_dataset("dset_len", data=[n])


def test_reader_of_the_reference_file_layout_and_batch_contract():
    fake_h5py.reset()
    z = np.load(os.path.join(G, "lazyload_align.npz"))
    eps, runs = _golden_runs(z)
    seasons = [build_season([runs[ep]]) for ep in eps]
    paths = ["/data/friends_sub-01_s1_n%d_llFile.h5" % i for i in range(len(seasons))]
    for p, a in zip(paths, seasons):
        _write_reference_layout(p, a)
    ds = VLB_Dataset(paths, opener=lambda p: fake_h5py.File(p, "r"))
    lens = [len(a["timeseries"]) for a in seasons]
    assert len(ds) == sum(lens)
    assert ds.locate(0) == (0, 0) and ds.locate(lens[0] - 1) == (0, lens[0] - 1) and ds.locate(lens[0]) == (1, 0)
    assert ds.locate(len(ds) - 1) == (len(seasons) - 1, lens[-1] - 1)
    flat = {k: np.concatenate([a[k] for a in seasons], 0) for k in seasons[0]}
    for i in (0, lens[0] - 1, lens[0], len(ds) - 1):
        it = ds[i]
        for k in ("timeseries", "vision", "language"):                       # float32 tensors (datamodule.py:104-105)
            assert isinstance(it[k], torch.Tensor) and it[k].dtype == torch.float32
            assert np.array_equal(it[k].numpy(), flat[k][i].astype(np.float32))
        for k in ("padvals", "vis_weights", "lang_weights"):                 # numpy, native dtype (datamodule.py:106-107)
            assert isinstance(it[k], np.ndarray) and np.array_equal(it[k], flat[k][i])
    # datamodule over those files: one held-out file, dset_names for the logger (train.py:50)
    cfg = VLBDataModuleConfig(lazyload_path="unused", subject="sub-01", seasons=["s1"], delay=3, window=3, random_state=1234,
                              shuffle_val_data=False, batch_size=3, num_workers=0)
    dsets = VLBDatasets(cfg, opener=lambda p: fake_h5py.File(p, "r"), files=paths)
    assert len(dsets.dset_names["val_set"]) == 1 and len(dsets.dset_names["train_set"]) == len(paths) - 1
    dm = VLBDataModule(cfg, datasets=dsets)
    batch = next(iter(dm.val_dataloader()))
    assert batch["timeseries"].dtype == torch.float32 and batch["padvals"].dtype == torch.int64
    assert batch["vis_weights"].dtype == torch.float64 and batch["lang_weights"].dtype == torch.float64
    assert batch["timeseries"].shape[0] == 3
    fake_h5py.reset()


def test_array_dataset_and_prefetcher_preserve_order_and_values():
    rng = np.random.RandomState(1)
    n = 10
    arrays = dict(timeseries=rng.randn(n, 6).astype(np.float32), vision=rng.randint(0, 256, (n, 2, 3, 4, 4)).astype(np.uint8),
                  language=rng.randint(0, 50, (n, 7)).astype(np.float32), padvals=rng.randint(0, 5, (n, 3)).astype(np.int64),
                  vis_weights=rng.rand(n, 2), lang_weights=rng.rand(n, 8))
    ds = ArrayVLBDataset(arrays)
    assert len(ds) == n and ds[3]["vision"].dtype == torch.uint8              # raw frames stay uint8 end to end
    batches = [ds.batch(range(i, min(i + 3, n))) for i in range(0, n, 3)]
    same = [collate_contract([ds[j] for j in range(i, min(i + 3, n))]) for i in range(0, n, 3)]
    for a, b in zip(batches, same):
        for k in a:
            assert torch.equal(a[k], b[k]) and a[k].dtype == b[k].dtype
    got = [{k: v.clone() for k, v in b.items()} for b in DevicePrefetcher(batches, "cpu")]
    assert len(got) == len(batches)
    for a, b in zip(got, batches):
        for k in b:
            assert torch.equal(a[k], b[k])
