"""Frozen-weight loading from checkpoints that use the Hugging Face / VideoLLaMA2 key names.

Reference: load_pretrained_vllama2 (src/litmodule/videollama2_vlb_litmodule.py:58-122) calls
`Videollama2MistralForCausalLM.from_pretrained(config.model_path, torch_dtype=bfloat16, local_files_only=True)`; the module
tree behind it (VideoLLaMA2/videollama2/model/videollama2_mistral.py:47-61, videollama2_arch.py:31-37, encoder.py:12-29,
projector.py:133-187) fixes the key names:

  model.embed_tokens.weight, model.norm.weight
  model.layers.N.{input_layernorm,post_attention_layernorm}.weight
  model.layers.N.self_attn.{q,k,v,o}_proj.weight, model.layers.N.mlp.{gate,up,down}_proj.weight
  model.vision_tower.vision_tower.vision_model.{embeddings.*, pre_layrnorm.*, encoder.layers.N.*}
  model.mm_projector.{s1,s2}.bK.{conv1,conv2,conv3,downsample}.{conv,bn}.*, .se.{fc1,fc2}.*, sampler.0.*, readout.{0,2}.*
  (lm_head.weight and ViT layer 24 exist in the checkpoint and are never read: the loss does not use them)

peft prefixes (`base_model.model.`) and wrappers (`.base_layer`) are stripped, so a merged or an adapter-wrapped state dict
loads alike. `open_checkpoint(dir)` reads `config.json` + `*.safetensors` (sharded or not) or `pytorch_model*.bin` lazily,
one tensor at a time, straight into the kernel layouts of model.FrozenWeights (bf16, K-contiguous, fused qkv / gate-up).
"""
import glob
import json
import os

import torch

from .model import FrozenWeights, ModelConfig, _bf, rope_tables

VT = "model.vision_tower.vision_tower.vision_model."
PJ = "model.mm_projector."


def canonical_key(k):
    """Strip peft / wrapper decorations from a checkpoint key."""
    while k.startswith("base_model.model."):
        k = k[len("base_model.model."):]
    if k.startswith("nnmodule."):                      # a Lightning checkpoint of the reference litmodule
        k = canonical_key(k[len("nnmodule."):])
    return k.replace(".base_layer.", ".")


class TensorSource:
    """name -> tensor, lazily, over a dict or over the files of a checkpoint directory."""

    def __init__(self, mapping=None):
        self._map = {}
        self._files = {}
        if mapping is not None:
            for k, v in mapping.items():
                self._map[canonical_key(k)] = v

    def add_safetensors(self, path):
        from safetensors import safe_open
        with safe_open(path, framework="pt", device="cpu") as f:
            for k in f.keys():
                self._files[canonical_key(k)] = (path, k)

    def add_torch_file(self, path):
        sd = torch.load(path, map_location="cpu", weights_only=True)
        sd = sd.get("state_dict", sd)
        for k, v in sd.items():
            self._map[canonical_key(k)] = v

    def __contains__(self, k):
        return k in self._map or k in self._files

    def keys(self):
        return list(self._map) + list(self._files)

    def __getitem__(self, k):
        if k in self._map:
            return self._map[k]
        if k in self._files:
            from safetensors import safe_open
            path, raw = self._files[k]
            with safe_open(path, framework="pt", device="cpu") as f:
                return f.get_tensor(raw)
        raise KeyError("checkpoint has no tensor %r" % k)


def open_checkpoint(path):
    """-> (TensorSource, config dict) for a local directory holding config.json and the weight files."""
    if not os.path.isdir(path):
        raise FileNotFoundError("model_path %r is not a local directory (no network here: hub ids cannot be resolved; "
                                "the reference itself loads with local_files_only=True, litmodule.py:81)" % (path,))
    cfg_path = os.path.join(path, "config.json")
    cfg = json.load(open(cfg_path)) if os.path.exists(cfg_path) else {}
    src = TensorSource()
    st = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    for p in st:
        src.add_safetensors(p)
    if not st:
        bins = sorted(glob.glob(os.path.join(path, "pytorch_model*.bin"))) + sorted(glob.glob(os.path.join(path, "*.pt")))
        if not bins:
            raise FileNotFoundError("no *.safetensors / pytorch_model*.bin under %r" % (path,))
        for p in bins:
            src.add_torch_file(p)
    # the CLIP tower may live in its own directory (config.mm_vision_tower, encoder.py:21-29) when the VideoLLaMA2
    # checkpoint was saved without it
    vt = cfg.get("mm_vision_tower")
    if VT + "pre_layrnorm.weight" not in src and vt and os.path.isdir(str(vt)):
        sub = TensorSource()
        for p in sorted(glob.glob(os.path.join(vt, "*.safetensors"))):
            sub.add_safetensors(p)
        for k in sub.keys():
            if k.startswith("vision_model."):
                src._files["model.vision_tower.vision_tower." + k] = sub._files[k]
        vcfg_path = os.path.join(vt, "config.json")
        if os.path.exists(vcfg_path):
            vc = json.load(open(vcfg_path))
            cfg.setdefault("_vision_config", vc.get("vision_config", vc))
    return src, cfg


def model_config_from_hf(hf, **overrides):
    """config.json of a VideoLLaMA2 checkpoint -> ModelConfig (rms_norm_eps, rope_theta, mm_vision_select_layer,
    tokenizer_model_max_length come from the checkpoint, not from constants)."""
    vc = hf.get("_vision_config", {})
    kw = dict(
        hidden=hf.get("hidden_size", 4096), layers=hf.get("num_hidden_layers", 32), heads=hf.get("num_attention_heads", 32),
        kv_heads=hf.get("num_key_value_heads", hf.get("num_attention_heads", 32)), mlp=hf.get("intermediate_size", 14336),
        vocab=hf.get("vocab_size", 32000), rms_eps=hf.get("rms_norm_eps", 1e-5), rope_theta=hf.get("rope_theta", 1e6),
        max_len=hf.get("tokenizer_model_max_length", 2048), select_layer=hf.get("mm_vision_select_layer", -2),
        vit_hidden=vc.get("hidden_size", hf.get("mm_hidden_size", 1024)), vit_layers=vc.get("num_hidden_layers", 24),
        vit_heads=vc.get("num_attention_heads", 16), vit_mlp=vc.get("intermediate_size", 4096),
        image_size=vc.get("image_size", 336), patch_size=vc.get("patch_size", 14), vit_ln_eps=vc.get("layer_norm_eps", 1e-5))
    pt = hf.get("mm_projector_type", "stc_connector")
    if pt != "stc_connector":
        raise NotImplementedError("mm_projector_type=%r: only the STC connector of VideoLLaMA2-7B is built (projector.py:95-122)" % pt)
    if hf.get("head_dim") not in (None, kw["hidden"] // kw["heads"]):
        raise NotImplementedError("head_dim != hidden_size / num_attention_heads")
    kw.update(overrides)
    return ModelConfig(**kw)


def _stc_block(src, p, dev):
    C = src[p + "conv1.conv.weight"].shape[0]
    se1 = src[p + "se.fc1.weight"]
    d = dict(
        conv1_w=_bf(src[p + "conv1.conv.weight"].reshape(C, -1), dev), n1_w=_bf(src[p + "conv1.bn.weight"], dev),
        n1_b=_bf(src[p + "conv1.bn.bias"], dev),
        dw_w=_bf(src[p + "conv2.conv.weight"].reshape(C, 9).t(), dev), n2_w=_bf(src[p + "conv2.bn.weight"], dev),
        n2_b=_bf(src[p + "conv2.bn.bias"], dev),
        se1_w=_bf(se1.reshape(se1.shape[0], C), dev), se1_b=_bf(src[p + "se.fc1.bias"], dev),
        se2_w=_bf(src[p + "se.fc2.weight"].reshape(C, -1), dev), se2_b=_bf(src[p + "se.fc2.bias"], dev),
        conv3_w=_bf(src[p + "conv3.conv.weight"].reshape(C, C), dev), n3_w=_bf(src[p + "conv3.bn.weight"], dev),
        n3_b=_bf(src[p + "conv3.bn.bias"], dev), ds_w=None)
    if p + "downsample.conv.weight" in src:
        d.update(ds_w=_bf(src[p + "downsample.conv.weight"].reshape(C, -1), dev), dsn_w=_bf(src[p + "downsample.bn.weight"], dev),
                 dsn_b=_bf(src[p + "downsample.bn.bias"], dev))
    return d


def load_frozen(src, cfg: ModelConfig, dev, parts=("vit", "stc", "llm")):
    """Checkpoint tensors (HF / VideoLLaMA2 names) -> FrozenWeights in the kernels' layouts (model.py docstring)."""
    if not isinstance(src, TensorSource):
        src = TensorSource(src)
    W = FrozenWeights()
    cat = lambda names: torch.cat([src[n] for n in names], 0)
    if "vit" in parts:
        e = VT + "embeddings."
        pw = src[e + "patch_embedding.weight"].detach().reshape(cfg.vit_hidden, -1).float()
        pwp = torch.zeros(cfg.vit_hidden, cfg.kpad)
        pwp[:, :pw.shape[1]] = pw
        W.vit = dict(patch_w=_bf(pwp, dev), cls=_bf(src[e + "class_embedding"], dev), pos=_bf(src[e + "position_embedding.weight"], dev),
                     pre_w=_bf(src[VT + "pre_layrnorm.weight"], dev), pre_b=_bf(src[VT + "pre_layrnorm.bias"], dev))
        for i in range(cfg.vit_used_layers):
            p = VT + "encoder.layers.%d." % i
            a = p + "self_attn."
            W.vit_layers.append(dict(
                ln1_w=_bf(src[p + "layer_norm1.weight"], dev), ln1_b=_bf(src[p + "layer_norm1.bias"], dev),
                wqkv=_bf(cat([a + "q_proj.weight", a + "k_proj.weight", a + "v_proj.weight"]), dev),
                bqkv=_bf(cat([a + "q_proj.bias", a + "k_proj.bias", a + "v_proj.bias"]), dev),
                wo=_bf(src[a + "out_proj.weight"], dev), bo=_bf(src[a + "out_proj.bias"], dev),
                ln2_w=_bf(src[p + "layer_norm2.weight"], dev), ln2_b=_bf(src[p + "layer_norm2.bias"], dev),
                fc1_w=_bf(src[p + "mlp.fc1.weight"], dev), fc1_b=_bf(src[p + "mlp.fc1.bias"], dev),
                fc2_w=_bf(src[p + "mlp.fc2.weight"], dev), fc2_b=_bf(src[p + "mlp.fc2.bias"], dev)))
    if "stc" in parts:
        C = cfg.hidden
        W.stc = dict(
            s1=[_stc_block(src, PJ + "s1.b%d." % (i + 1), dev) for i in range(cfg.stc_depth)],
            s2=[_stc_block(src, PJ + "s2.b%d." % (i + 1), dev) for i in range(cfg.stc_depth)],
            sampler_w=_bf(src[PJ + "sampler.0.weight"].permute(0, 2, 3, 4, 1).reshape(C, 8 * C), dev),
            sampler_b=_bf(src[PJ + "sampler.0.bias"], dev),
            ro1_w=_bf(src[PJ + "readout.0.weight"], dev), ro1_b=_bf(src[PJ + "readout.0.bias"], dev),
            ro2_w=_bf(src[PJ + "readout.2.weight"], dev), ro2_b=_bf(src[PJ + "readout.2.bias"], dev))
    if "llm" in parts:
        W.llm = dict(embed=_bf(src["model.embed_tokens.weight"], dev), norm_w=_bf(src["model.norm.weight"], dev))
        W.llm["cos"], W.llm["sin"] = rope_tables(cfg, dev)
        for i in range(cfg.layers):
            p = "model.layers.%d." % i
            a, m = p + "self_attn.", p + "mlp."
            W.llm_layers.append(dict(
                ln1_w=_bf(src[p + "input_layernorm.weight"], dev),
                wqkv=_bf(cat([a + "q_proj.weight", a + "k_proj.weight", a + "v_proj.weight"]), dev),
                wo=_bf(src[a + "o_proj.weight"], dev),
                ln2_w=_bf(src[p + "post_attention_layernorm.weight"], dev),
                wgu=_bf(cat([m + "gate_proj.weight", m + "up_proj.weight"]), dev),
                wd=_bf(src[m + "down_proj.weight"], dev)))
        shapes = W.llm_layers[0]
        D = cfg.head_dim
        assert shapes["wqkv"].shape == ((cfg.heads + 2 * cfg.kv_heads) * D, cfg.hidden), "checkpoint / config.json mismatch"
        assert shapes["wgu"].shape == (2 * cfg.mlp, cfg.hidden)
    return W


def load_pretrained(path, dev, **overrides):
    """model_path=<local dir> -> (ModelConfig, FrozenWeights)."""
    src, hf = open_checkpoint(path)
    depth = len({k.split(".")[3] for k in src.keys() if k.startswith(PJ + "s1.b")})      # RegStage depth: blocks b1..bN
    if depth:
        overrides.setdefault("stc_depth", depth)
    cfg = model_config_from_hf(hf, **overrides)
    return cfg, load_frozen(src, cfg, dev)
