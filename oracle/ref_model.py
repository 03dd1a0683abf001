"""ORACLE (test infrastructure only -- only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference arm
may import this; the product path never does).

CPU fp32 restatement of the phantom_vlb training-step hot path, assembled from
  * the installed Hugging Face `transformers` CLIPVisionModel / MistralModel (the L0 arithmetic the reference itself
    reaches: VideoLLaMA2/videollama2/model/encoder.py:27-29,50; videollama2_mistral.py:23-24,47,94-104; the reference
    pins transformers==4.53.3, this image has 5.5 -- same math for these two models),
  * restated timm RegStage pieces (oracle/stc_timm.py), restated peft lora.Linear (peft==0.16.0 semantics:
    y = W x + (alpha/r) * B(A(dropout(x))), src/litmodule/videollama2_vlb_litmodule.py:113-120),
  * the reference's own logic followed function by function:
      STCConnector.forward                    VideoLLaMA2/videollama2/model/projector.py:189-215
      CLIPVisionTower.feature_select/forward  VideoLLaMA2/videollama2/model/encoder.py:31-53
      encode_images_or_videos                 VideoLLaMA2/videollama2/model/videollama2_arch.py:114-134
      prepare_inputs_labels_for_multimodal    VideoLLaMA2/videollama2/model/videollama2_arch.py:161-263
      HRFConvolveLayer / RidgeRegressionLayer src/utils.py:40-73
      make_weight_mask / forward / training_step   src/litmodule/videollama2_vlb_litmodule.py:178-203, 229-256, 259-306
      Pearson r (torchmetrics PearsonCorrCoef after nan_to_num)  src/utils.py:101-110
PARITY PINNING: the reference ships no golden vectors or tests for this path, so the oracle is pinned
against the reference EXECUTED in the build container by oracle/make_goldens.py (fixtures frozen under tests/golden/):
  backbone_ref.npz   Videollama2MistralForCausalLM.forward(input_ids, attention_mask, output_hidden_states=True, images=...)
                     exactly as litmodule.py:231-236 calls it, + encode_images_or_videos, prepare_inputs_labels_for_multimodal
                     (splice + left-extended attention mask) and CLIPVisionTower.forward, on seeded weights
                     -> RefVLB.encode_frames / encode_videos / splice / backbone          (tests/test_oracle_cpu.py)
  stc_connector.npz  STCConnector (projector.py:133-215) with the restated timm stage injected
  head_utils.npz     HRFConvolveLayer / RidgeRegressionLayer (src/utils.py:40-73)
  weight_mask.npz    VLBLitModule.make_weight_mask (litmodule.py:178-203)
  frame_sample.npz, video_chunks.npz, lazyload_align.npz   the integer indexing contract (oracle/indexing.py)
Still "parity unpinned" (third-party packages absent here, restated from their published behaviour): timm RegStage internals
(oracle/stc_timm.py), peft lora.Linear, torchmetrics PearsonCorrCoef (checked against numpy), nilearn's Glover regressor
(oracle/glover.py).
"""
import math
from dataclasses import dataclass, field

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .stc_timm import RegStage, LayerNorm2d

VIDEO_TOKEN_ID = -201          # MODAL_INDEX_MAP["<video>"], VideoLLaMA2/videollama2/constants.py:18,28-32


@dataclass
class RefConfig:
    # vision tower (CLIP ViT-L/14-336 at full size)
    vit_hidden: int = 1024
    vit_layers: int = 24
    vit_heads: int = 16
    vit_mlp: int = 4096
    image_size: int = 336
    patch_size: int = 14
    select_layer: int = -2
    # decoder (Mistral-7B at full size)
    hidden: int = 4096
    layers: int = 32
    heads: int = 32
    kv_heads: int = 8
    mlp: int = 14336
    vocab: int = 32000
    rms_eps: float = 1e-5
    rope_theta: float = 1e6
    # STC connector
    stc_depth: int = 4
    # task
    num_frames: int = 8
    num_target: int = 1000
    max_len: int = 2048
    lora_r: int = 16
    lora_alpha: int = 32
    l2_lambda: float = 0.001
    lang_w_len: int = 64
    attn_impl: str = "eager"      # "sdpa" only for bench.py's same-GPU timing of this module (parity runs stay eager, fp32)

    @property
    def grid(self):
        return self.image_size // self.patch_size

    @property
    def grid_out(self):
        return self.grid // 2 + 1

    @property
    def t_out(self):
        return self.num_frames // 2 + 1

    @property
    def tok_per_frame(self):
        return self.grid_out * self.grid_out

    @property
    def n_vis(self):
        return self.t_out * self.tok_per_frame

    @property
    def l_text(self):
        return self.max_len - self.n_vis + 1


def tiny_config(**kw):
    """Small widths that still satisfy the kernels' alignment rules (head_dim 64 ViT / 128 decoder)."""
    c = RefConfig(vit_hidden=128, vit_layers=3, vit_heads=2, vit_mlp=256, image_size=56, patch_size=14,
                  hidden=256, layers=2, heads=2, kv_heads=1, mlp=512, vocab=512, stc_depth=2,
                  num_frames=4, num_target=40, max_len=128, lang_w_len=16)
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def config1(**kw):
    """BASELINE.json configs[0]: reduced depth (2 ViT + 2 LLM layers), full widths, 8 frames @336^2, 1000 parcels."""
    c = RefConfig(vit_layers=2, layers=2)
    for k, v in kw.items():
        setattr(c, k, v)
    return c


# ------------------------------------------------------------------------------------------------ LoRA (peft restated)
class LoraLinear(nn.Module):
    def __init__(self, base: nn.Linear, r: int, alpha: int, dropout: float = 0.0):
        super().__init__()
        self.base_layer = base
        self.r, self.scaling = r, alpha / r
        self.lora_A = nn.Linear(base.in_features, r, bias=False)
        self.lora_B = nn.Linear(r, base.out_features, bias=False)
        nn.init.kaiming_uniform_(self.lora_A.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_B.weight)
        self.lora_dropout = nn.Dropout(dropout) if dropout > 0 else nn.Identity()
        for p in self.base_layer.parameters():
            p.requires_grad_(False)

    def forward(self, x):
        return self.base_layer(x) + self.lora_B(self.lora_A(self.lora_dropout(x))) * self.scaling


LORA_TARGETS = ("q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj")


# ------------------------------------------------------------------------------------------------ STC connector
class STCConnector(nn.Module):
    """projector.py:133-215 restated (padding=1 variant, 'stc_connector')."""

    def __init__(self, cfg: RefConfig):
        super().__init__()
        self.s1 = RegStage(depth=cfg.stc_depth, in_chs=cfg.vit_hidden, out_chs=cfg.hidden, stride=1, dilation=1,
                           act_layer=nn.SiLU, norm_layer=LayerNorm2d)
        self.sampler = nn.Sequential(nn.Conv3d(cfg.hidden, cfg.hidden, kernel_size=2, stride=2, padding=1, bias=True),
                                     nn.SiLU())
        self.s2 = RegStage(depth=cfg.stc_depth, in_chs=cfg.hidden, out_chs=cfg.hidden, stride=1, dilation=1,
                           act_layer=nn.SiLU, norm_layer=LayerNorm2d)
        self.readout = nn.Sequential(nn.Linear(cfg.hidden, cfg.hidden), nn.GELU(), nn.Linear(cfg.hidden, cfg.hidden))

    def forward(self, x):                       # x [b, t, l, d]
        b, t, l, d = x.shape
        hw = int(l ** 0.5)
        x = x.view(b, t, hw, hw, d).permute(0, 1, 4, 2, 3).reshape(b * t, d, hw, hw)     # (b t) d h w
        x = self.s1(x)
        x = x.view(b, t, -1, hw, hw).permute(0, 2, 1, 3, 4)                               # b d t h w
        x = self.sampler(x)
        new_t, nh, nw = x.shape[2], x.shape[3], x.shape[4]
        x = x.permute(0, 2, 1, 3, 4).reshape(b * new_t, -1, nh, nw)
        x = self.s2(x)
        x = x.view(b, new_t, -1, nh * nw).permute(0, 1, 3, 2).reshape(b, new_t * nh * nw, -1)   # b (t h w) d
        return self.readout(x)


# ------------------------------------------------------------------------------------------------ the model
class RefVLB(nn.Module):
    """Backbone + brain head of VLBLitModule (LoRA mode), fp32 on CPU."""

    def __init__(self, cfg: RefConfig, seed: int = 1234, lora_b_std: float = 0.02, lora_dropout: float = 0.0,
                 head_dropout: float = 0.0):
        super().__init__()
        from transformers import CLIPVisionConfig, CLIPVisionModel, MistralConfig, MistralModel
        self.cfg = cfg
        torch.manual_seed(seed)
        vc = CLIPVisionConfig(hidden_size=cfg.vit_hidden, intermediate_size=cfg.vit_mlp, num_hidden_layers=cfg.vit_layers,
                              num_attention_heads=cfg.vit_heads, image_size=cfg.image_size, patch_size=cfg.patch_size,
                              hidden_act="quick_gelu", layer_norm_eps=1e-5, projection_dim=cfg.vit_hidden)
        vc._attn_implementation = cfg.attn_impl
        self.vision_tower = CLIPVisionModel(vc)
        self.mm_projector = STCConnector(cfg)
        mc = MistralConfig(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.mlp,
                           num_hidden_layers=cfg.layers, num_attention_heads=cfg.heads, num_key_value_heads=cfg.kv_heads,
                           head_dim=cfg.hidden // cfg.heads, rms_norm_eps=cfg.rms_eps, rope_theta=cfg.rope_theta,
                           sliding_window=None, max_position_embeddings=max(cfg.max_len, 4096), use_cache=False)
        mc._attn_implementation = cfg.attn_impl
        self.llm = MistralModel(mc)
        for p in self.parameters():
            p.requires_grad_(False)
        # get_peft_model(...LoraConfig(target_modules=find_all_linear_names)) on the decoder (litmodule.py:101-120)
        for layer in self.llm.layers:
            for name in LORA_TARGETS:
                parent = layer.self_attn if name in ("q_proj", "k_proj", "v_proj", "o_proj") else layer.mlp
                setattr(parent, name, LoraLinear(getattr(parent, name), cfg.lora_r, cfg.lora_alpha, lora_dropout))
        if lora_b_std > 0:      # parity runs exercise the adapter path with a non-zero B (SURVEY 8c golden generation)
            g = torch.Generator().manual_seed(seed + 1)
            for m in self.modules():
                if isinstance(m, LoraLinear):
                    m.lora_B.weight.data.copy_(torch.randn(m.lora_B.weight.shape, generator=g) * lora_b_std)
        # head (litmodule.py:217-226)
        self.layer_norm1 = nn.LayerNorm(cfg.hidden)
        self.layer_norm2 = nn.LayerNorm(cfg.hidden)
        self.ridge = nn.Linear(cfg.hidden, cfg.num_target, bias=True)
        self.dropout = nn.Dropout(head_dropout) if head_dropout > 0 else nn.Identity()      # litmodule.py:226

    # ---- encoder.py:31-53
    def encode_frames(self, frames):
        out = self.vision_tower(frames, output_hidden_states=True)
        return out.hidden_states[self.cfg.select_layer][:, 1:]

    # ---- videollama2_arch.py:114-134
    def encode_videos(self, vision):               # [B, T, 3, H, W]
        B, T = vision.shape[:2]
        feats = self.encode_frames(vision.reshape(B * T, *vision.shape[2:]))
        feats = feats.view(B, T, feats.shape[1], feats.shape[2])
        self.last_frame_feats = feats.detach()          # per-stage parity checks (tests only)
        return self.mm_projector(feats)

    # ---- videollama2_arch.py:161-263 (one <video> per sample, equal lengths -> no re-padding branch)
    def splice(self, input_ids, attention_mask, mm_features):
        embeds = []
        emb = self.llm.embed_tokens
        for b in range(input_ids.shape[0]):
            ids = input_ids[b]
            idx = torch.where(ids == VIDEO_TOKEN_ID)[0]
            assert idx.numel() == 1, "exactly one <video> token per sample"
            s = int(idx[0])
            embeds.append(torch.cat([emb(ids[:s]), mm_features[b], emb(ids[s + 1:])], dim=0))
        embeds = torch.stack(embeds, dim=0)
        pad_left = torch.ones((attention_mask.shape[0], embeds.shape[1] - input_ids.shape[1]), dtype=attention_mask.dtype)
        attention_mask = torch.cat([pad_left, attention_mask], dim=1)
        assert attention_mask.shape == embeds.shape[:2]
        return embeds, attention_mask

    # ---- litmodule.py:178-203
    def make_weight_mask(self, pad_vals, vis_weights, lang_weights, lang_len, max_len, dtype=torch.float32):
        tok = self.cfg.tok_per_frame
        feature_len = vis_weights.shape[1] * tok + lang_len - 1
        assert feature_len == max_len
        rows = []
        for i in range(pad_vals.shape[0]):
            pad_len, inst_len, dialog_len = (int(v) for v in pad_vals[i])
            trial = torch.cat([
                vis_weights[i].repeat_interleave(tok).to(dtype),
                torch.zeros(2 + inst_len).to(dtype),
                lang_weights[i][:dialog_len].to(dtype),
                torch.zeros(4 + pad_len).to(dtype),
            ], dim=0)
            pad_left = feature_len - trial.shape[0]
            rows.append(torch.cat([torch.zeros(pad_left).to(dtype), trial], dim=0).unsqueeze(0))
        return torch.cat(rows, dim=0)

    def backbone(self, vision, input_ids, attention_mask):
        mm = self.encode_videos(vision)
        embeds, mask = self.splice(input_ids, attention_mask, mm)
        out = self.llm(inputs_embeds=embeds, attention_mask=mask, use_cache=False)
        return out.last_hidden_state, mm, embeds            # hidden_states[-1] (post final norm)

    # ---- litmodule.py:229-256
    def head(self, hidden, weight_mask):
        h = self.layer_norm1(hidden)
        pooled = torch.einsum("bse,bs->be", h, weight_mask)        # src/utils.py:56
        hrf = self.dropout(self.layer_norm2(pooled))
        pred = self.ridge(hrf)
        l2 = self.cfg.l2_lambda * torch.norm(self.ridge.weight, p=2) ** 2     # src/utils.py:70
        return pred, l2, pooled, hrf

    # ---- litmodule.py:259-306 (dropout off: parity runs)
    def step(self, batch, weight_dtype=torch.bfloat16, round_inputs=True):
        """Returns dict(loss, pred, hidden, ...).  round_inputs mimics the reference's .to(bf16) casts of pixels,
        BOLD targets and HRF weights (values only; arithmetic stays fp32)."""
        vision = batch["vision"]
        y = batch["timeseries"]
        if round_inputs:
            vision = vision.bfloat16().float()
            y = y.bfloat16().float()
        ids = batch["language"].long()
        mask = ids.ne(0).long()
        wm = self.make_weight_mask(batch["padvals"], batch["vis_weights"], batch["lang_weights"], ids.shape[1],
                                   self.cfg.max_len, dtype=weight_dtype if round_inputs else torch.float32).float()
        hidden, mm, embeds = self.backbone(vision, ids, mask)
        pred, l2, pooled, hrf = self.head(hidden, wm)
        loss = F.mse_loss(pred, y) + l2
        return dict(loss=loss, pred=pred, l2=l2, hidden=hidden, mm=mm, embeds=embeds, weight_mask=wm, pooled=pooled,
                    hrf=hrf, frame_feats=self.last_frame_feats, input_ids=ids)

    def trainable(self):
        return [(n, p) for n, p in self.named_parameters() if p.requires_grad]


def canonical_params(model: RefVLB):
    """{VideoLLaMA2 checkpoint key: tensor} for the backbone of a RefVLB -- the names Videollama2MistralForCausalLM's
    state_dict uses (videollama2_mistral.py:47-61, videollama2_arch.py:31-37, encoder.py:12-29)."""
    out = {}
    for k, v in model.vision_tower.state_dict().items():
        out["model.vision_tower.vision_tower." + k] = v
    for k, v in model.mm_projector.state_dict().items():
        out["model.mm_projector." + k] = v
    for k, v in model.llm.state_dict().items():
        if "lora_" not in k:
            out["model." + k.replace(".base_layer.", ".")] = v
    return out


def seed_weights_(named, seed=1234):
    """Deterministic values for a {key: tensor} map, a function of (seed, key, shape) only -- so the reference model (built in
    the container by oracle/make_goldens.py) and a RefVLB (built anywhere) hold identical weights without shipping them."""
    import zlib
    with torch.no_grad():
        for k in sorted(named):
            v = named[k]
            g = torch.Generator().manual_seed(seed * 1000003 + zlib.crc32(k.encode()))
            r = torch.randn(v.shape, generator=g)
            if v.dim() >= 2 or k.endswith("class_embedding"):
                v.copy_(r * 0.08)
            elif k.endswith("weight"):
                v.copy_(1.0 + 0.1 * r)          # norm scales
            else:
                v.copy_(0.05 * r)               # biases
    return named


def round_weights_to_bf16(model: RefVLB):
    """Make every parameter bf16-representable (values only; arithmetic stays fp32). The reference loads its backbone with
    torch_dtype=bfloat16 (litmodule.py:74-82) and builds the head with dtype=bfloat16 (:217-226), so "identical weights"
    for a parity run means weights both sides can hold exactly."""
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(p.bfloat16().float())
    return model


def enable_training(model: RefVLB):
    """LoRA A/B + head trainable, everything else frozen (load_pretrained_vllama2, litmodule.py:58-122)."""
    for n, p in model.named_parameters():
        p.requires_grad_(("lora_A" in n) or ("lora_B" in n) or n.startswith(("layer_norm1", "layer_norm2", "ridge")))
    return model


# ------------------------------------------------------------------------------------------------ Pearson (torchmetrics restated)
def pearson_r(preds, target):
    """Per-column Pearson r in fp64 on nan_to_num'd inputs, clamped to [-1, 1] (src/utils.py:101-106)."""
    x = np.nan_to_num(np.asarray(preds, dtype=np.float64))
    y = np.nan_to_num(np.asarray(target, dtype=np.float64))
    xm, ym = x - x.mean(0), y - y.mean(0)
    num = (xm * ym).sum(0)
    den = np.sqrt((xm * xm).sum(0) * (ym * ym).sum(0))
    with np.errstate(divide="ignore", invalid="ignore"):
        r = num / den
    return np.clip(r, -1.0, 1.0)


# ------------------------------------------------------------------------------------------------ synthetic batch
def synthetic_batch(cfg: RefConfig, B: int, seed: int = 1234):
    """Friends-shaped synthetic batch obeying the datamodule contract (src/datamodule/videollama2_vlb_datamodule.py:98-109)."""
    g = torch.Generator().manual_seed(seed)
    vision = torch.randn(B, cfg.num_frames, 3, cfg.image_size, cfg.image_size, generator=g)
    L = cfg.l_text
    inst_len = min(9, max(1, L // 16))
    lang = torch.zeros(B, L)
    padvals = torch.zeros(B, 3, dtype=torch.int64)
    for b in range(B):
        max_dialog = min(cfg.lang_w_len - 6, L // 4)
        dialog_len = int(torch.randint(2, max(3, max_dialog), (1,), generator=g))
        pad_len = int(torch.randint(0, max(1, L // 2 - dialog_len - inst_len - 8), (1,), generator=g))
        n_sys = L - (1 + 2 + inst_len + dialog_len + 4 + pad_len) - 1     # BOS + sys + <video> + ...
        assert n_sys >= 0
        toks = torch.randint(3, cfg.vocab, (L,), generator=g).float()
        toks[0] = 1
        toks[1 + n_sys] = VIDEO_TOKEN_ID
        if pad_len > 0:
            toks[L - pad_len:] = 0
        lang[b] = toks
        padvals[b] = torch.tensor([pad_len, inst_len, dialog_len])
    tp = cfg.t_out
    lag_v = 1.49 * (5.5 - np.arange(tp) * (3.0 / max(tp - 1, 1)))
    vis_w = torch.tensor(np.stack([glover_like(lag_v) for _ in range(B)]), dtype=torch.float64)
    lang_w = torch.zeros(B, cfg.lang_w_len, dtype=torch.float64)
    for b in range(B):
        d = int(padvals[b, 2])
        lags = 1.5 + 7.5 * torch.rand(d, generator=g).double().numpy()
        lang_w[b, :d] = torch.tensor(glover_like(lags))
    ts = torch.randn(B, cfg.num_target, generator=g)
    return dict(timeseries=ts, vision=vision, language=lang, padvals=padvals, vis_weights=vis_w, lang_weights=lang_w)


def glover_like(t):
    """Closed-form double-gamma HRF (Glover 1999 parameters: peaks 6 s / 16 s, ratio 1/6), normalised to peak 1.
    Only used to give the synthetic HRF weights a realistic shape; they are *inputs* of the step."""
    t = np.asarray(t, dtype=np.float64)
    a1, a2, b = 6.0, 16.0, 1.0
    h = (t ** (a1 - 1) * np.exp(-t / b) / math.gamma(a1)) - (t ** (a2 - 1) * np.exp(-t / b) / math.gamma(a2)) / 6.0
    return h / 0.1758


def pearson_stats(preds, target):
    """Raw moments [5, P] (sum x, sum y, sum x^2, sum y^2, sum xy) in fp64 -- additive across row shards."""
    x = np.nan_to_num(np.asarray(preds, dtype=np.float64))
    y = np.nan_to_num(np.asarray(target, dtype=np.float64))
    return np.stack([x.sum(0), y.sum(0), (x * x).sum(0), (y * y).sum(0), (x * y).sum(0)])


def pearson_from_stats(stats, n):
    sx, sy, sxx, syy, sxy = stats
    with np.errstate(divide="ignore", invalid="ignore"):
        r = (sxy - sx * sy / n) / np.sqrt((sxx - sx * sx / n) * (syy - sy * sy / n))
    return np.clip(r, -1.0, 1.0)
