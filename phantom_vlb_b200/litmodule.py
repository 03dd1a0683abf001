"""Host-side mirror of src/litmodule/videollama2_vlb_litmodule.py: same config keys, class / hook / attribute names
and argument meaning, with the arithmetic replaced by the libvlb_b200 engine.

  VLBLitModuleConfig            litmodule.py:126-157  (16 hydra keys + dtype / device_map from __post_init__)
  VLBLitModule.configure_model  :206-226   build the backbone + head (here: Engine over frozen bf16 weights)
              .make_weight_mask :178-203   -> vlb_weight_mask (bit-exact layout)
              .forward          :229-256   (regression_output, l2_reg)
              .training_step    :259-306   loss tensor wired to autograd (Lightning calls .backward() on it)
              .validation_step  :309-342   {'loss','brain_preds','brain_vals'}
              .configure_optimizers :345-379  AdamW + lr_scheduler (CosineAnnealingLR), here a fused clip+AdamW kernel
Trainable state is ONE nn.Parameter (`flat`, fp32) aliasing the engine's flat master buffer, so Lightning's DDP,
gradient clipping and checkpointing all operate on a single contiguous tensor; `named_trainables()` exposes the
PEFT-style views for checkpoint interchange.
`model_path` accepts "random:7b" | "random:config1" | "random:tiny" (random-init backbones for parity runs and
benchmarks, there is no network for checkpoints here) or a prepared FrozenWeights object via `frozen=`.
"""
from dataclasses import dataclass

import torch
from torch.optim import lr_scheduler

try:
    from lightning import LightningModule
except Exception:                                   # Lightning is not installed in this image
    class LightningModule(torch.nn.Module):
        def log(self, *a, **k):
            self._logged = getattr(self, "_logged", {})
            if a:
                self._logged[a[0]] = a[1]

        def log_dict(self, d, **k):
            for key, v in d.items():
                self.log(key, v)

from . import ops
from .engine import Engine
from .model import ModelConfig, Trainables, frozen_random
from .utils import HRFConvolveLayer, RidgeRegressionLayer


@dataclass
class VLBLitModuleConfig:
    model_path: str
    freeze_backbone: bool
    use_lora: bool
    lora_r: int | None
    lora_alpha: int | None
    lora_dropout: float | None
    dropout_rate: float
    num_target: int
    l2_lambda: float
    lr: float
    betas: list[float]
    eps: float
    weight_decay: float
    lr_scheduler_name: str
    last_epoch: int
    t_max: int

    def __post_init__(self):
        self.dtype = torch.bfloat16
        self.device_map = "auto"


def model_config_for(config: VLBLitModuleConfig, num_frames=8) -> ModelConfig:
    kind = str(config.model_path).split(":")[-1].lower()
    common = dict(num_target=config.num_target, l2_lambda=config.l2_lambda, num_frames=num_frames,
                  lora_r=config.lora_r or 16, lora_alpha=config.lora_alpha or 32, use_lora=bool(config.use_lora))
    if kind in ("7b", "videollama2-7b"):
        return ModelConfig(**common)
    if kind == "config1":
        return ModelConfig(vit_layers=2, layers=2, **common)
    if kind == "tiny":
        common["num_frames"] = 4
        return ModelConfig(vit_hidden=128, vit_layers=3, vit_heads=2, vit_mlp=256, image_size=56, hidden=256, layers=2,
                           heads=2, kv_heads=1, mlp=512, vocab=512, stc_depth=2, max_len=128, lang_w_len=16, **common)
    raise NotImplementedError("model_path=%r: loading a Hugging Face checkpoint is not wired yet; use random:7b / "
                              "random:config1 / random:tiny or pass frozen= to configure_model" % (config.model_path,))


class _StepFn(torch.autograd.Function):
    """Forward AND backward run eagerly inside the engine; autograd only has to hand the flat gradient to `flat`."""

    @staticmethod
    def forward(ctx, flat, module, batch):
        out, pred = module.engine.loss_and_grads(batch)
        ctx.module = module
        module._last = (out, pred)
        return out[0].clone()

    @staticmethod
    def backward(ctx, g):
        return ctx.module.engine.T.grad * g, None, None


class FusedClipAdamW(torch.optim.Optimizer):
    """torch.optim-compatible front of vlb_adamw_step: global-norm clip + AdamW + bf16 working-copy refresh in one pass."""

    def __init__(self, params, engine, lr, betas, eps, weight_decay, max_norm=1.0, world_size=1):
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.engine, self.max_norm, self.world_size = engine, max_norm, world_size
        self.allreduce = None       # set to a callable(tensor) to sum gradients across ranks outside Lightning's DDP

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        grp = self.param_groups[0]
        p = grp["params"][0]
        T = self.engine.T
        if p.grad is not None and p.grad.data_ptr() != T.grad.data_ptr():
            T.grad.copy_(p.grad)
        self.engine.optimizer_step(grp["lr"], betas=grp["betas"], eps=grp["eps"], weight_decay=grp["weight_decay"],
                                   max_norm=self.max_norm, world_size=self.world_size if self.allreduce else 1,
                                   allreduce=self.allreduce)                 # under DDP p.grad is already averaged
        return loss


class VLBLitModule(LightningModule):
    def __init__(self, config: VLBLitModuleConfig) -> None:
        super().__init__()
        self.config: VLBLitModuleConfig = config
        self.engine = None

    # ------------------------------------------------------------------ model construction (litmodule.py:206-226)
    def configure_model(self, frozen=None, model_config: ModelConfig = None, device="cuda", seed=1234, lora_b_std=0.0,
                        target_slice=None):
        if self.engine is not None:
            return
        # the two experiments the reference ships (config/experiment/*.yaml): LoRA (use_lora, backbone otherwise frozen)
        # and the frozen-backbone baseline (freeze_backbone=True, use_lora=False: head-only training)
        if self.config.use_lora == self.config.freeze_backbone:
            raise NotImplementedError("full fine-tuning (use_lora=False, freeze_backbone=False) is not built, and "
                                      "use_lora=True with freeze_backbone=True is contradictory (litmodule.py:86-120)")
        mc = model_config or model_config_for(self.config)
        dev = torch.device(device)
        W = frozen if frozen is not None else frozen_random(mc, dev, seed)
        T = Trainables(mc, dev, target_slice=target_slice)
        T.init_default(seed, lora_b_std=lora_b_std)
        self.engine = Engine(mc, W, T, device=dev, dropout_rate=self.config.dropout_rate,
                             lora_dropout=self.config.lora_dropout or 0.0)
        self.model_config = mc
        self.flat = torch.nn.Parameter(T.param)           # aliases the fp32 master buffer
        self.hrf_layer = HRFConvolveLayer()
        self.ridge_layer = RidgeRegressionLayer(mc.hidden, T.num_target_local, self.config.l2_lambda)
        self._params_dirty = False

    def named_trainables(self):
        """PEFT-style (name, fp32 view) pairs of the flat buffer, B matrices still transposed ([r, N])."""
        T = self.engine.T
        return [(n, T.p(n)) for n, _ in T.spec]

    # ------------------------------------------------------------------ adapter-only checkpoints
    # train.py:58-60 saves the full ~16 GB state dict and leaves "saving only LoRA adapters" as a TODO; here the checkpoint
    # holds the trainable tensors only (LoRA A/B + head, ~184 MB fp32) under the keys peft / the reference module tree use.
    _PEFT = "nnmodule.base_model.model.model.layers.%d.%s.%s.lora_%s.default.weight"
    _PROJ = (("self_attn", "q_proj", "Acat_qkv", 0, "Bt_q"), ("self_attn", "k_proj", "Acat_qkv", 1, "Bt_k"),
             ("self_attn", "v_proj", "Acat_qkv", 2, "Bt_v"), ("self_attn", "o_proj", "A_o", 0, "Bt_o"),
             ("mlp", "gate_proj", "Acat_gu", 0, "Bt_g"), ("mlp", "up_proj", "Acat_gu", 1, "Bt_u"),
             ("mlp", "down_proj", "A_d", 0, "Bt_d"))
    _HEAD = (("ridge_layer.linear.weight", "ridge_w"), ("ridge_layer.linear.bias", "ridge_b"),
             ("layer_norm1.weight", "ln1_w"), ("layer_norm1.bias", "ln1_b"),
             ("layer_norm2.weight", "ln2_w"), ("layer_norm2.bias", "ln2_b"))

    def _adapter_views(self):
        """(checkpoint key, fp32 view of the flat master, transposed?) for every trainable tensor."""
        T, r = self.engine.T, self.model_config.lora_r
        out = []
        for l in range(T.lora_layers):
            for blk, proj, a_name, a_idx, b_name in self._PROJ:
                a = T.p("layers.%d.%s" % (l, a_name))[a_idx * r:(a_idx + 1) * r]
                out.append((self._PEFT % (l, blk, proj, "A"), a, False))
                out.append((self._PEFT % (l, blk, proj, "B"), T.p("layers.%d.%s" % (l, b_name)), True))   # stored [r, N]
        out += [(k, T.p(n), False) for k, n in self._HEAD]
        return out

    def adapter_state_dict(self):
        return {k: (v.t() if tr else v).detach().clone().contiguous().cpu() for k, v, tr in self._adapter_views()}

    def load_adapter_state_dict(self, sd, strict=True):
        missing = []
        for k, v, tr in self._adapter_views():
            if k not in sd:
                missing.append(k)
                continue
            src = sd[k].to(device=v.device, dtype=torch.float32)
            v.copy_(src.t() if tr else src)
        if strict and missing:
            raise KeyError("adapter checkpoint is missing %d keys, e.g. %s" % (len(missing), missing[0]))
        self.engine.T.sync_working()
        return missing

    def save_adapters(self, path):
        torch.save({"state_dict": self.adapter_state_dict(), "step_count": self.engine.T.step_count,
                    "adam_m": self.engine.T.m.cpu(), "adam_v": self.engine.T.v.cpu()}, path)

    def load_adapters(self, path, load_optimizer=True):
        ck = torch.load(path, map_location="cpu")
        self.load_adapter_state_dict(ck["state_dict"])
        if load_optimizer and "adam_m" in ck:
            T = self.engine.T
            T.m.copy_(ck["adam_m"]); T.v.copy_(ck["adam_v"]); T.step_count = int(ck["step_count"])

    def on_save_checkpoint(self, checkpoint):           # Lightning hook: trainer.save_checkpoint(...) (train.py:58)
        checkpoint["state_dict"] = self.adapter_state_dict()

    # ------------------------------------------------------------------ helpers
    def make_weight_mask(self, pad_vals, vis_weights, lang_weights, lang_len, max_len):
        mc = self.model_config
        feature_len = vis_weights.shape[1] * mc.tok_per_frame + lang_len - 1
        assert feature_len == max_len
        dev = self.engine.dev
        return ops.weight_mask(pad_vals.to(dev).contiguous(), vis_weights.to(dev).contiguous(),
                               lang_weights.to(dev).contiguous(), mc.tok_per_frame, max_len)

    def _to_device(self, batch):
        dev = self.engine.dev
        return {k: (v if v.device == dev else v.to(dev, non_blocking=True)) for k, v in batch.items()}

    def _refresh(self):
        # an external optimiser (plain torch AdamW) may have written `flat`; FusedClipAdamW keeps copies in sync itself
        if self._params_dirty:
            self.engine.T.sync_working()
            self._params_dirty = False

    # ------------------------------------------------------------------ forward / steps
    def forward(self, x_video, x_lang, weight_mask=None, attention_mask=None, batch=None):
        """Reference signature (litmodule.py:229). The weight mask and attention mask are recomputed on the device
        from `batch`; returns (regression_output, l2_reg)."""
        assert batch is not None, "pass the datamodule batch dict (batch=...)"
        out, pred = self.engine.eval_step(self._to_device(batch))
        return pred, out[2]

    def training_step(self, batch):
        self._refresh()
        loss = _StepFn.apply(self.flat, self, self._to_device(batch))
        self.log("train/brain_loss", loss)
        return loss

    def validation_step(self, batch):
        self._refresh()
        b = self._to_device(batch)
        out, pred = self.engine.eval_step(b)
        y = b["timeseries"].to(self.config.dtype)
        self.log("val/brain_loss", out[0])
        return {"loss": out[0], "brain_preds": pred.detach().clone(), "brain_vals": y.detach()}

    def configure_optimizers(self):
        self.optimizer = FusedClipAdamW([self.flat], self.engine, lr=self.config.lr, betas=self.config.betas,
                                        eps=self.config.eps, weight_decay=self.config.weight_decay)
        self.lr_scheduler_args = {"last_epoch": self.config.last_epoch, "T_max": self.config.t_max}
        self.scheduler = getattr(lr_scheduler, self.config.lr_scheduler_name)(self.optimizer, **self.lr_scheduler_args)
        return [self.optimizer], [{"scheduler": self.scheduler, "interval": "step", "frequency": 1}]
