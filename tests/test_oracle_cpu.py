"""The oracle against (i) golden vectors produced by executing the reference's own files (oracle/make_goldens.py),
(ii) the known-answer indexing table of SURVEY.md 8(a-13), (iii) numpy/scipy for Pearson."""
import os
import types

import numpy as np
import torch

from oracle import indexing
from oracle.ref_model import RefConfig, RefVLB, STCConnector, pearson_from_stats, pearson_r, pearson_stats

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_head_layers_match_reference_utils():
    z = np.load(os.path.join(G, "head_utils.npz"))
    emb, w = torch.from_numpy(z["emb"]), torch.from_numpy(z["w"])
    pooled = torch.einsum("bse,bs->be", emb, w)
    assert np.allclose(pooled.numpy(), z["pooled"], atol=1e-6)
    W, b = torch.from_numpy(z["ridge_w"]), torch.from_numpy(z["ridge_b"])
    out = pooled @ W.t() + b
    assert np.allclose(out.numpy(), z["out"], atol=1e-5)
    assert abs(0.01 * float(torch.norm(W, p=2) ** 2) - float(z["l2"])) < 1e-6


def test_weight_mask_matches_reference_bit_exact():
    z = np.load(os.path.join(G, "weight_mask.npz"))
    cfg = RefConfig(num_frames=8)
    me = types.SimpleNamespace(cfg=cfg)
    wm = RefVLB.make_weight_mask(me, torch.from_numpy(z["padvals"]), torch.from_numpy(z["vis_w"]),
                                 torch.from_numpy(z["lang_w"]), 1204, 2048, dtype=torch.bfloat16).float().numpy()
    assert np.array_equal(wm, z["mask"])
    for i in range(3):       # float64 layout restatement, rounded the same way
        pad, inst, dialog = (int(v) for v in z["padvals"][i])
        row = indexing.weight_mask_row(pad, inst, dialog, z["vis_w"][i], z["lang_w"][i], 169, 2048)
        assert np.array_equal(torch.from_numpy(row).to(torch.bfloat16).float().numpy(), z["mask"][i])


def test_frame_sample_matches_reference():
    z = np.load(os.path.join(G, "frame_sample.npz"))
    for d in z["durations"]:
        for n in (4, 8, 12):
            assert np.array_equal(indexing.frame_sample_uniform(int(d), n), z["d%d_n%d" % (d, n)])


def test_indexing_known_answers():
    # fps=29.97, tr=1.49, window=3, frames_per_tr=4
    fps, N = 29.97, 40000
    exp = {1: (0, 43, 4, [5, 16, 27, 38], 8), 2: (0, 88, 8, [6, 17, 28, 39, 50, 61, 72, 83], 4),
           3: (0, 132, 12, [6, 17, 28, 39, 50, 61, 72, 83, 94, 105, 116, 127], 0),
           10: (311, 445, 12, [317, 328, 339, 350, 361, 372, 384, 395, 406, 417, 428, 439], 0)}
    for k, (fs, fe, nf, idx, pad) in exp.items():
        got = indexing.video_chunk_indices(k * 1.49, 3, fps, N)
        assert got[0] == fs and got[1] == fe and got[2] == nf and list(got[3]) == idx and got[4] == pad
    assert [indexing.text_budget(t) for t in (8, 12, 16)] == [1204, 866, 528]
    assert np.allclose(indexing.visual_token_lags(8), [8.195, 7.0775, 5.96, 4.8425, 3.725])
    assert np.allclose(indexing.visual_token_lags(12), [8.195, 7.45, 6.705, 5.96, 5.215, 4.47, 3.725])
    assert len(indexing.visual_token_lags(16)) == 9
    assert indexing.bold_alignment(100, 100) == (2, 5, 95)
    assert indexing.tr_end_times(10.0)[:3] == [1.49, 2.98, 4.47]


def test_stc_connector_matches_reference_glue():
    z = np.load(os.path.join(G, "stc_connector.npz"))
    cfg = RefConfig(vit_hidden=32, hidden=64, stc_depth=2)
    stc = STCConnector(cfg)
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd.")}
    stc.load_state_dict(sd)
    y = stc(torch.from_numpy(z["x"]))
    assert y.shape == tuple(z["y"].shape)
    assert np.allclose(y.detach().numpy(), z["y"], atol=1e-5)


def test_pearson_matches_numpy_and_merges():
    rng = np.random.RandomState(0)
    x, y = rng.randn(500, 37), rng.randn(500, 37)
    y[:, :10] += 0.7 * x[:, :10]
    x[3, 5] = np.nan
    r = pearson_r(x, y)
    xc = np.nan_to_num(x)
    for c in (0, 5, 20):
        assert abs(r[c] - np.corrcoef(xc[:, c], y[:, c])[0, 1]) < 1e-12
    s = pearson_stats(x[:200], y[:200]) + pearson_stats(x[200:], y[200:])
    assert np.allclose(pearson_from_stats(s, 500), r, atol=1e-12)


def test_backbone_matches_reference_model_forward():
    """backbone_ref.npz was produced by running the reference's own Videollama2MistralForCausalLM.forward (the call of
    litmodule.py:231-236), encode_images_or_videos (videollama2_arch.py:114-134), prepare_inputs_labels_for_multimodal
    (:161-263) and CLIPVisionTower.forward (encoder.py:40-53) on seeded weights; the oracle must reproduce every stage."""
    from oracle.make_goldens import BACKBONE_GOLDEN
    from oracle.ref_model import canonical_params, seed_weights_
    z = np.load(os.path.join(G, "backbone_ref.npz"))
    cfg = RefConfig(**BACKBONE_GOLDEN)
    ref = RefVLB(cfg, seed=0, lora_b_std=0.0).eval()
    named = canonical_params(ref)
    assert len(named) == int(z["n_keys"]) and all(str(k) in named for k in z["key_sample"])    # same state_dict key set
    seed_weights_(named, seed=1234)
    ids, mask, pixels = torch.from_numpy(z["ids"]), torch.from_numpy(z["mask"]), torch.from_numpy(z["pixels"])
    B, T = pixels.shape[:2]
    with torch.no_grad():
        frames = ref.encode_frames(pixels.reshape(B * T, *pixels.shape[2:]))
        mm = ref.encode_videos(pixels)
        embeds, mask_ext = ref.splice(ids, mask, mm)
        hidden, _, _ = ref.backbone(pixels, ids, mask)
    assert np.allclose(frames.numpy(), z["frame_feats"], atol=1e-5)
    assert np.allclose(mm.numpy(), z["mm"], atol=1e-5)
    assert np.array_equal(mask_ext.numpy(), z["mask_ext"])                       # left-extended attention mask
    assert np.allclose(embeds.numpy(), z["embeds"], atol=1e-5)
    assert hidden.shape == tuple(z["hidden_last"].shape)
    err = np.abs(hidden.numpy() - z["hidden_last"]).max()
    assert err < 2e-5, err


def test_video_chunk_indexing_matches_reference_execution():
    """video_chunks.npz: load_video / extract_video_chunk / get_max_token of videollama2_vlb_extractfeatures.py executed on a
    fake reader (frame k holds k+1, black pads hold 0)."""
    z = np.load(os.path.join(G, "video_chunks.npz"))
    ci = 0
    while "c%d_params" % ci in z.files:
        fps, nfr, win, fpt, dur = z["c%d_params" % ci]
        nfr, win, fpt = int(nfr), int(win), int(fpt)
        trs = indexing.tr_end_times(dur)
        assert np.array_equal(np.array(trs), z["c%d_tr_list" % ci])
        for k, frames in zip(z["c%d_k" % ci], z["c%d_frames" % ci]):
            _, _, nf, sampled, pads = indexing.video_chunk_indices(trs[int(k)], win, fps, nfr, frames_per_tr=fpt)
            exp = list(np.array(sampled) + 1) + [0] * pads
            assert list(frames) == exp, (ci, k)
        ci += 1
    assert ci == 5
    for w, f, mt in z["max_token"]:
        assert indexing.text_budget(int(w) * int(f)) == int(mt)


def test_lazyload_alignment_matches_reference_execution():
    """lazyload_align.npz: make_lazy_loading_dsets (videollama2_vlb_lazyloading.py:51-166) executed over an in-memory h5py
    double. The oracle's alignment formulas must give the same rows / lags."""
    from oracle import glover
    z = np.load(os.path.join(G, "lazyload_align.npz"))
    window, delay = int(z["window"]), int(z["delay"])
    eps = sorted({k.split(".")[1] for k in z.files if k.startswith("in.")})
    chunk_idx = np.floor(np.arange(len(eps)) / (len(eps) / len(z["files"]))).astype(int)      # lazyloading.py:89-91
    for fi in range(len(z["files"])):
        rows = {k: [] for k in ("timeseries", "vision", "language", "padvals", "lang_weights")}
        for ep in np.array(eps)[chunk_idx == fi]:
            feat, bold = z["in.%s.video_features" % ep], z["in.%s.bold" % ep]
            f0, b0, n = indexing.bold_alignment(feat.shape[0], bold.shape[0], delay, window)
            rows["timeseries"].append(bold[b0:b0 + n])
            rows["vision"].append(feat[f0:f0 + n])
            rows["language"].append(z["in.%s.transcript_features" % ep][f0:f0 + n])
            mp = z["in.%s.masking_params" % ep][f0:f0 + n]
            rows["padvals"].append(mp)
            on = z["in.%s.transcript_onsets" % ep][f0:f0 + n].copy()
            for i in range(n):
                d = int(mp[i, 2])
                t_mid = indexing.tr_mid_time(i, delay, window)
                on[i, :d] = [glover.get_hrf_weight(t) for t in t_mid - on[i, :d]]
            rows["lang_weights"].append(on)
        assert int(z["f%d.len" % fi]) == sum(len(r) for r in rows["timeseries"])
        for k, v in rows.items():
            assert np.allclose(np.concatenate(v), z["f%d.%s" % (fi, k)], atol=1e-12), (fi, k)
        T = z["f%d.vision" % fi].shape[1]
        lags = indexing.visual_token_lags(T, delay, window)
        vw = np.array([glover.get_hrf_weight(t) for t in lags])
        assert np.allclose(z["f%d.vis_weights" % fi], vw[None, :], atol=1e-12)
