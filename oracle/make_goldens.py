"""Generates tests/golden/*.npz by EXECUTING the reference's own files from /root/reference (build container only;
the GPU box has no /root/reference).  Third-party imports that are absent here (lightning, nilearn, torchmetrics, timm,
peft, h5py, decord ...) are stubbed; timm's RegStage/LayerNorm2d are provided by oracle/stc_timm.py, so the
STCConnector golden pins the reference's *glue* (einops rearranges, sampler, readout), not timm's arithmetic.

    python -m oracle.make_goldens            # rewrites tests/golden/
"""
import importlib.machinery
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def ref_utils():
    _stub("lightning"); _stub("lightning.pytorch"); _stub("lightning.pytorch.callbacks", Callback=object)
    _stub("nilearn"); _stub("nilearn.glm"); _stub("nilearn.glm.first_level", compute_regressor=None)
    _stub("torchmetrics", PearsonCorrCoef=None)
    return _load(os.path.join(REF, "src/utils.py"), "ref_src_utils")


def ref_projector():
    from oracle import stc_timm
    _stub("timm"); _stub("timm.models"); _stub("timm.models.regnet", RegStage=stc_timm.RegStage)
    _stub("timm.models.layers", LayerNorm=torch.nn.LayerNorm, LayerNorm2d=stc_timm.LayerNorm2d)
    import transformers
    if not hasattr(transformers, "TRANSFORMERS_CACHE"):      # removed in transformers 5 (reference pins 4.53.3)
        transformers.TRANSFORMERS_CACHE = "/tmp"
    return _load(os.path.join(REF, "VideoLLaMA2/videollama2/model/projector.py"), "ref_projector")


def ref_mm_utils_frame_sample():
    src = open(os.path.join(REF, "VideoLLaMA2/videollama2/mm_utils.py")).read()
    start = src.index("def frame_sample(")
    end = src.index("\ndef ", start + 10)
    ns = {"np": np, "NUM_FRAMES_PER_SECOND": 1}
    exec(src[start:end], ns)
    return ns["frame_sample"]


def ref_make_weight_mask():
    """Extract VLBLitModule.make_weight_mask (litmodule.py:178-203) and bind it to a minimal object."""
    src = open(os.path.join(REF, "src/litmodule/videollama2_vlb_litmodule.py")).read()
    start = src.index("    def make_weight_mask(")
    end = src.index("    def configure_model(", start)
    import textwrap
    ns = {"torch": torch}
    exec(textwrap.dedent(src[start:end]), ns)

    class Self:
        device = torch.device("cpu")

        class config:
            dtype = torch.bfloat16
    return lambda *a: ns["make_weight_mask"](Self(), *a)


def ref_videollama2():
    """Import the reference's Videollama2MistralForCausalLM (videollama2_mistral.py:47-108 + videollama2_arch.py +
    encoder.py + projector.py) WITHOUT running videollama2/__init__.py (which pulls decord etc.): the two packages are
    registered as path-only stubs, so the relative imports resolve to the reference's own files."""
    ref_projector()                                   # timm + TRANSFORMERS_CACHE stubs
    import importlib
    import transformers
    import transformers.generation.utils  # noqa: F401  (importing these can swap sys.modules["transformers"])
    sys.modules["transformers"].TRANSFORMERS_CACHE = "/tmp"
    transformers.TRANSFORMERS_CACHE = "/tmp"
    root = os.path.join(REF, "VideoLLaMA2/videollama2")
    _stub("videollama2").__path__ = [root]
    _stub("videollama2.model").__path__ = [os.path.join(root, "model")]
    mistral = importlib.import_module("videollama2.model.videollama2_mistral")
    enc = importlib.import_module("videollama2.model.encoder")
    clip = enc.CLIPVisionModel

    class Clip453(clip):
        """transformers 5.5 records `hidden_states` through forward hooks, and a CLIPVisionModel nested inside another
        PreTrainedModel gets them registered twice: the tuple comes back with every layer output duplicated (2N+1 entries), so
        the reference's `hidden_states[-2]` (encoder.py:32) would pick the LAST layer instead of the penultimate one it selects
        under its pinned transformers==4.53.3 (N+1 entries). Consecutive duplicates are dropped to restore the pinned layout."""

        def forward(self, *a, **kw):
            out = super().forward(*a, **kw)
            hs = out.hidden_states
            if hs is not None:
                out.hidden_states = (hs[0],) + tuple(h for p, h in zip(hs, hs[1:]) if h.data_ptr() != p.data_ptr())
                assert len(out.hidden_states) == self.config.num_hidden_layers + 1
            return out

    def eager_clip(config=None, **kw):
        config._attn_implementation = "eager"       # encoder.py:24 asks for flash_attention_2 (CUDA only): same math
        return Clip453(config=config, **kw)
    enc.CLIPVisionModel = eager_clip
    return mistral


# widths of the backbone golden (CPU-only fixture: no kernel alignment rules apply, so it can be very small)
BACKBONE_GOLDEN = dict(vit_hidden=32, vit_layers=3, vit_heads=2, vit_mlp=64, image_size=28, patch_size=14, hidden=64, layers=2,
                       heads=2, kv_heads=1, mlp=128, vocab=128, stc_depth=4, num_frames=4, num_target=8, max_len=40, lang_w_len=8)


def golden_backbone():
    """Runs the reference model's forward exactly as litmodule.forward does (litmodule.py:231-236):
    nnmodule(input_ids, attention_mask, output_hidden_states=True, images=[(clip, "video")]*B) -> hidden_states[-1],
    plus encode_images_or_videos (videollama2_arch.py:114-134) and prepare_inputs_labels_for_multimodal (:161-263)."""
    import tempfile
    from transformers import CLIPImageProcessor, CLIPVisionConfig
    from oracle.ref_model import seed_weights_
    mistral = ref_videollama2()
    c = BACKBONE_GOLDEN
    d = tempfile.mkdtemp(prefix="clip-vit-")          # build_vision_tower dispatches on "clip" in the path (encoder.py:156)
    CLIPVisionConfig(hidden_size=c["vit_hidden"], intermediate_size=c["vit_mlp"], num_hidden_layers=c["vit_layers"],
                     num_attention_heads=c["vit_heads"], image_size=c["image_size"], patch_size=c["patch_size"],
                     hidden_act="quick_gelu", layer_norm_eps=1e-5, projection_dim=c["vit_hidden"]).save_pretrained(d)
    CLIPImageProcessor(size={"shortest_edge": c["image_size"]}, crop_size={"height": c["image_size"], "width": c["image_size"]}).save_pretrained(d)
    cfg = mistral.Videollama2MistralConfig(
        vocab_size=c["vocab"], hidden_size=c["hidden"], intermediate_size=c["mlp"], num_hidden_layers=c["layers"],
        num_attention_heads=c["heads"], num_key_value_heads=c["kv_heads"], head_dim=c["hidden"] // c["heads"], rms_norm_eps=1e-5,
        rope_theta=1e6, sliding_window=None, max_position_embeddings=4096, use_cache=False)
    cfg.mm_vision_tower, cfg.mm_vision_select_layer, cfg.mm_vision_select_feature = d, -2, "patch"
    cfg.mm_projector_type, cfg.mm_hidden_size, cfg.num_frames = "stc_connector", c["vit_hidden"], c["num_frames"]
    cfg._attn_implementation = "eager"
    torch.manual_seed(0)
    model = mistral.Videollama2MistralForCausalLM(cfg).eval()
    sd = {k: v for k, v in model.state_dict().items() if k != "lm_head.weight"}
    seed_weights_(sd, seed=1234)
    B, T = 3, c["num_frames"]
    grid_out = (c["image_size"] // c["patch_size"]) // 2 + 1
    n_vis = (T // 2 + 1) * grid_out * grid_out
    L = c["max_len"] - n_vis + 1
    g = torch.Generator().manual_seed(77)
    ids = torch.randint(3, c["vocab"], (B, L), generator=g)
    for b, (pos, pad) in enumerate([(2, 0), (5, 4), (1, 9)]):
        ids[b, pos] = -201
        if pad:
            ids[b, L - pad:] = 0
    mask = ids.ne(0).long()
    pixels = torch.randn(B, T, 3, c["image_size"], c["image_size"], generator=g)
    images = [(pixels[i], "video") for i in range(B)]
    with torch.no_grad():
        out = model(input_ids=ids, attention_mask=mask, output_hidden_states=True, images=images)
        mm = model.encode_images_or_videos(images)
        _, mask2, _, embeds, _ = model.prepare_inputs_labels_for_multimodal(ids, mask, None, None, images)
        frames = model.get_model().get_vision_tower()(pixels.reshape(B * T, 3, c["image_size"], c["image_size"]))
    assert len(out.hidden_states) == c["layers"] + 1
    np.savez_compressed(os.path.join(OUT, "backbone_ref.npz"), ids=ids.numpy(), mask=mask.numpy(), pixels=pixels.numpy().astype(np.float32),
                        frame_feats=frames.numpy(), mm=mm.numpy(), embeds=embeds.numpy(), mask_ext=mask2.numpy(),
                        hidden_last=out.hidden_states[-1].numpy(), n_keys=len(sd),
                        key_sample=np.array(sorted(sd)[::17]))


def _func_source(path, name, until=None):
    src = open(path).read()
    start = src.index("def %s(" % name)
    end = src.index("\ndef ", start + 10) if until is None else src.index(until, start)
    return src[start:end]


def golden_video_chunks():
    """Executes load_video's TR list, extract_video_chunk's frame window / sampling / black-frame padding and get_max_token
    (src/preprocessing/videollama2_vlb_extractfeatures.py:198-212, 303-346) with a fake decord reader whose frame k is
    filled with the value k+1 (so the output tells which frames were sampled; 0 = black pad)."""
    import math
    path = os.path.join(REF, "src/preprocessing/videollama2_vlb_extractfeatures.py")

    class Img:
        def __init__(self, a):
            self.a = a
            self.size = (a.shape[1], a.shape[0])

    class Image:
        fromarray = staticmethod(lambda a: Img(a))

    class Batch:
        def __init__(self, a):
            self.a = a

        def asnumpy(self):
            return self.a

    class Reader:
        def __init__(self, n, fps):
            self.n, self.fps = n, fps

        def __len__(self):
            return self.n

        def get_avg_fps(self):
            return self.fps

        def get_batch(self, idx):
            return Batch(np.stack([np.full((2, 2, 3), i + 1, dtype=np.int64) for i in idx]))

    class Proc:
        image_mean = [0.5, 0.5, 0.5]

        @staticmethod
        def preprocess(images):
            return {"pixel_values": [im.a for im in images]}

    ns = {"np": np, "math": math, "Image": Image, "expand2square": lambda im, c: im, "frame_sample": ref_mm_utils_frame_sample(),
          "VideoReader": lambda p, ctx=None, num_threads=1: p, "cpu": lambda i: None}
    exec(_func_source(path, "load_video"), ns)
    exec(_func_source(path, "extract_video_chunk"), ns)
    exec(_func_source(path, "get_max_token"), ns)
    out = {}
    cases = [(29.97, 14000, 3, 4), (29.97, 14000, 2, 4), (29.97, 14000, 4, 4), (25.0, 1000, 3, 4), (23.976, 333, 3, 2)]
    for ci, (fps, nfr, win, fpt) in enumerate(cases):
        rd = Reader(nfr, fps)
        _, fps_, n_, dur, tr_list = ns["load_video"](rd, tr=1.49)
        out["c%d_params" % ci] = np.array([fps, nfr, win, fpt, dur])
        out["c%d_tr_list" % ci] = np.array(tr_list)
        ks = sorted(set([0, 1, 2, 3, 9, len(tr_list) // 2, len(tr_list) - 2, len(tr_list) - 1]))
        ks = [k for k in ks if 0 <= k < len(tr_list)]
        vids = [ns["extract_video_chunk"](rd, Proc, tr_list[k], win, fps_, n_, 1.49, fpt)[:, 0, 0, 0] for k in ks]
        out["c%d_k" % ci] = np.array(ks)
        out["c%d_frames" % ci] = np.stack(vids)             # value = sampled frame index + 1, 0 = black pad
    import types as _t
    out["max_token"] = np.array([[w, f, ns["get_max_token"](_t.SimpleNamespace(window_duration=w, frames_per_tr=f, model_max_length=2048))]
                                 for w, f in [(2, 4), (3, 4), (4, 4), (3, 2), (1, 8)]])
    np.savez_compressed(os.path.join(OUT, "video_chunks.npz"), **out)


def golden_lazyload():
    """Executes make_lazy_loading_dsets (src/preprocessing/videollama2_vlb_lazyloading.py:51-166) on a synthetic feature file
    and BOLD file through an in-memory h5py double, with get_hrf_weight = the restated nilearn Glover regressor
    (oracle/glover.py; nilearn is absent). Pins the TR alignment: which feature / BOLD rows pair up, visual token lags,
    language token lags, n_rows, per-sample dataset names, chunking of episodes into n_split files."""
    from oracle import fake_h5py, glover
    fake_h5py.reset()
    _stub("h5py", File=fake_h5py.File)
    _stub("src", get_hrf_weight=glover.get_hrf_weight)
    mod = _load(os.path.join(REF, "src/preprocessing/videollama2_vlb_lazyloading.py"), "ref_lazyloading")
    rng = np.random.RandomState(5)
    window, delay, T, L, P = 3, 3, 12, 10, 6
    eps = {"s01e01a": 14, "s01e01b": 11, "s01e02a": 9}              # TRs of extracted features per episode half
    bold_len = {"s01e01a": 20, "s01e01b": 13, "s01e02a": 30}        # TRs of BOLD per run (longer / shorter than the features)
    with fake_h5py.File("/feat.h5", "w") as f:
        for ep, n in eps.items():
            gp = f.create_group(ep)
            gp.create_dataset("video_features", data=(np.arange(n)[:, None, None] * 100 + np.arange(T)[None, :, None]).astype(np.float32))
            gp.create_dataset("transcript_features", data=rng.randint(3, 100, (n, L)).astype(np.int64))
            mp = np.stack([rng.randint(0, 3, n), np.full(n, 2), rng.randint(1, 5, n)], 1)
            on = np.zeros((n, 8))
            for i in range(n):
                on[i, :mp[i, 2]] = (i + 1) * 1.49 - rng.uniform(0.0, window * 1.49, mp[i, 2])    # words spoken inside the window ending at TR i
            gp.create_dataset("transcript_onsets", data=on)
            gp.create_dataset("masking_params", data=mp)
    with fake_h5py.File("/bold.h5", "w") as f:
        ses = f.create_group("ses-001")
        for ep, n in bold_len.items():
            ses.create_dataset("ses-001_task-%s_timeseries" % ep, data=rng.randn(n, P).astype(np.float32))
    import types as _t
    cfg = _t.SimpleNamespace(features_path="/feat.h5", timeseries_path="/bold.h5", lazyload_path="/ll", subject="sub-01",
                             season="s1", n_split=2, delay=delay, window=window)
    mod.make_lazy_loading_dsets(cfg)
    out = {"window": window, "delay": delay}
    for ep in eps:
        g = fake_h5py.File("/feat.h5")[ep]
        for k in ("video_features", "transcript_features", "transcript_onsets", "masking_params"):
            out["in.%s.%s" % (ep, k)] = np.array(g[k])
        out["in.%s.bold" % ep] = np.array(fake_h5py.File("/bold.h5")["ses-001"]["ses-001_task-%s_timeseries" % ep])
    files = [p for p in fake_h5py.paths() if p.startswith("/ll/")]
    out["files"] = np.array(files)
    for fi, p in enumerate(files):
        f = fake_h5py.File(p)
        n = int(np.array(f["dset_len"])[0])
        out["f%d.len" % fi] = n
        for mod_name in ("timeseries", "vision", "vis_weights", "language", "lang_weights", "padvals"):
            out["f%d.%s" % (fi, mod_name)] = np.stack([np.array(f["%d" % i]["%d_%s" % (i, mod_name)]) for i in range(n)])
    np.savez_compressed(os.path.join(OUT, "lazyload_align.npz"), **out)
    fake_h5py.reset()


def golden_yaml():
    """The datamodule / litmodule / trainer sections of the reference's two experiment YAMLs, as parsed (values only)."""
    import yaml
    for name in ("VLB_vllama2_friends_lora", "VLB_vllama2_friends_baseline"):
        d = yaml.safe_load(open(os.path.join(REF, "config/experiment/%s.yaml" % name)))
        with open(os.path.join(OUT, "%s.yaml" % name), "w") as f:
            f.write("# datamodule / litmodule / trainer sections of the reference's config/experiment/%s.yaml as parsed by yaml.safe_load\n"
                    "# (written by oracle/make_goldens.py; interpolations like ${subject} kept as strings)\n" % name)
            yaml.safe_dump({k: d[k] for k in ("datamodule", "litmodule", "trainer")}, f, sort_keys=False)

def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(1234)
    # ---- src/utils.py head layers
    U = ref_utils()
    emb, w = torch.randn(3, 40, 32), torch.rand(3, 40)
    pooled = U.HRFConvolveLayer()(emb, w)
    ridge = U.RidgeRegressionLayer(32, 24, l2_lambda=0.01)
    out, l2 = ridge(pooled)
    np.savez(os.path.join(OUT, "head_utils.npz"), emb=emb.numpy(), w=w.numpy(), pooled=pooled.detach().numpy(),
             ridge_w=ridge.linear.weight.detach().numpy(), ridge_b=ridge.linear.bias.detach().numpy(),
             out=out.detach().numpy(), l2=float(l2))
    # ---- make_weight_mask
    mk = ref_make_weight_mask()
    g = torch.Generator().manual_seed(7)
    padvals = torch.tensor([[5, 9, 20], [0, 9, 58], [300, 9, 2]])
    visw, langw = torch.rand(3, 5, generator=g, dtype=torch.float64), torch.rand(3, 64, generator=g, dtype=torch.float64)
    wm = mk(padvals, visw, langw, 1204, 2048)
    np.savez(os.path.join(OUT, "weight_mask.npz"), padvals=padvals.numpy(), vis_w=visw.numpy(), lang_w=langw.numpy(),
             mask=wm.float().numpy())
    # ---- frame_sample / chunk indices known answers
    fs = ref_mm_utils_frame_sample()
    durations = [44, 89, 133, 135, 7, 12, 500]
    np.savez(os.path.join(OUT, "frame_sample.npz"), durations=np.array(durations),
             **{"d%d_n%d" % (d, n): fs(d, "uniform", n) for d in durations for n in (4, 8, 12)})
    # ---- STCConnector glue on tiny widths
    P = ref_projector()

    class Cfg:
        mm_hidden_size, hidden_size = 32, 64
    torch.manual_seed(99)
    stc = P.STCConnector(Cfg(), depth=2)
    x = torch.randn(2, 4, 16, 32)
    y = stc(x)
    sd = {k: v.detach().numpy() for k, v in stc.state_dict().items()}
    np.savez(os.path.join(OUT, "stc_connector.npz"), x=x.numpy(), y=y.detach().numpy(), **{"sd." + k: v for k, v in sd.items()})
    golden_yaml()
    golden_video_chunks()
    golden_lazyload()
    golden_backbone()
    print("goldens written to", OUT, sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
