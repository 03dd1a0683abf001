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
`model_path` is a LOCAL checkpoint directory (config.json + *.safetensors / pytorch_model*.bin with the Hugging Face /
VideoLLaMA2 key names -- weights.py; the reference itself loads with local_files_only=True, litmodule.py:81), or
"random:7b" | "random:config1" | "random:tiny" (random-init backbones of the exact architecture for parity runs and
benchmarks), or a prepared FrozenWeights object passed as configure_model(frozen=...).
"""
import os
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
    raise FileNotFoundError("model_path=%r is neither a local checkpoint directory nor random:7b / random:config1 / "
                            "random:tiny (hub ids cannot be resolved: the reference loads with local_files_only=True, "
                            "litmodule.py:81)" % (config.model_path,))


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
        eng = ctx.module.engine
        wait = getattr(eng.grad_ready, "wait", None)
        if wait is not None:                       # overlapped data-parallel reduction (dp.OverlappedGradReducer): let it land
            wait()
        return eng.T.grad * g, None, None


class FusedClipAdamW(torch.optim.Optimizer):
    """torch.optim-compatible front of vlb_adamw_step: global-norm clip + AdamW + bf16 working-copy refresh in one pass."""

    def __init__(self, params, engine, lr, betas, eps, weight_decay, max_norm=1.0, world_size=1):
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.engine, self.max_norm, self.world_size = engine, max_norm, world_size
        self.allreduce = None       # set to a callable(tensor) to sum gradients across ranks outside Lightning's DDP
        self.prereduced = False     # True: gradients were already SUMMED across world_size ranks (engine.grad_ready reducer)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():                   # torch.optim convention: the closure re-evaluates the model
                loss = closure()
        grp = self.param_groups[0]
        p = grp["params"][0]
        T = self.engine.T
        if torch.distributed.is_available() and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1 \
                and self.engine.fused_head_opt and self.engine.shard_reduce is None and self.allreduce is None \
                and not self.prereduced:
            raise RuntimeError("wide replicated readout head under DDP: its rank-B gradient is not part of flat.grad; shard the "
                               "head by voxel (target_slice + engine.shard_reduce) or set optimizer.allreduce + "
                               "engine.row_gather")
        if p.grad is not None and p.grad.data_ptr() != T.grad.data_ptr():
            T.grad.copy_(p.grad)
        self.engine.optimizer_step(grp["lr"], betas=grp["betas"], eps=grp["eps"], weight_decay=grp["weight_decay"],
                                   max_norm=self.max_norm,
                                   world_size=self.world_size if (self.allreduce or self.prereduced) else 1,
                                   allreduce=self.allreduce)                 # under DDP p.grad is already averaged
        return loss

    # Adam moments and the step count live in the engine's flat buffers: expose them the torch.optim way, so Lightning's
    # checkpoint["optimizer_states"] carries them and trainer.fit(ckpt_path=...) resumes bias correction and moments
    def state_dict(self):
        T = self.engine.T
        sd = super().state_dict()
        sd["vlb"] = {"adam_m": T.m.detach().cpu(), "adam_v": T.v.detach().cpu(), "step_count": int(T.step_count)}
        return sd

    def load_state_dict(self, state_dict):
        vlb = state_dict.get("vlb")
        super().load_state_dict({k: v for k, v in state_dict.items() if k != "vlb"})
        if vlb is not None:
            T = self.engine.T
            T.m.copy_(vlb["adam_m"]); T.v.copy_(vlb["adam_v"]); T.step_count = int(vlb["step_count"])


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
        dev = torch.device(device)
        path = str(self.config.model_path)
        if frozen is None and model_config is None and os.path.isdir(path):
            # load_pretrained_vllama2 (litmodule.py:58-122): config.json + weight files of a local VideoLLaMA2 checkpoint
            from .weights import load_pretrained
            mc, W = load_pretrained(path, dev, num_target=self.config.num_target, l2_lambda=self.config.l2_lambda,
                                    lora_r=self.config.lora_r or 16, lora_alpha=self.config.lora_alpha or 32,
                                    use_lora=bool(self.config.use_lora))
        else:
            mc = model_config or model_config_for(self.config)
            W = frozen if frozen is not None else frozen_random(mc, dev, seed)
        T = Trainables(mc, dev, target_slice=target_slice)
        T.init_default(seed, lora_b_std=lora_b_std)
        self.engine = Engine(mc, W, T, device=dev, dropout_rate=self.config.dropout_rate,
                             lora_dropout=self.config.lora_dropout or 0.0)
        self.model_config = mc
        self.flat = torch.nn.Parameter(T.param)           # aliases the fp32 master buffer
        self.hrf_layer = HRFConvolveLayer()
        self.ridge_layer = RidgeRegressionLayer(mc.hidden, T.num_target_local, self.config.l2_lambda, own_parameters=False)
        self.ridge_layer.bind(lambda: T.w("ridge_w"), lambda: T.w("ridge_b"))    # the readout slice of the flat buffer
        self._seen_version = self.flat._version
        rank = torch.distributed.get_rank() if torch.distributed.is_available() and torch.distributed.is_initialized() else 0
        self.engine.set_seed(seed, rank)

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

    # nn.Module protocol: the module's state IS the adapter + head tensors under the reference's key names, so Lightning's
    # save_checkpoint / load_from_checkpoint / fit(ckpt_path=...) round-trip without knowing about the flat buffer
    def state_dict(self, *args, destination=None, prefix="", keep_vars=False):
        if self.engine is None:
            return super().state_dict(*args, destination=destination, prefix=prefix, keep_vars=keep_vars)
        out = destination if destination is not None else {}
        for k, v in self.adapter_state_dict().items():
            out[prefix + k] = v
        return out

    def load_state_dict(self, state_dict, strict=True, assign=False):
        if self.engine is None:
            raise RuntimeError("call configure_model() before load_state_dict (Lightning does: configure_model hook)")
        if "flat" in state_dict:                        # a raw dump of the flat master
            self.engine.T.param.copy_(state_dict["flat"].to(self.engine.T.param.device))
            self.engine.T.sync_working()
            missing = []
        else:
            missing = self.load_adapter_state_dict(state_dict, strict=strict)
        known = {k for k, _, _ in self._adapter_views()} | {"flat"}
        unexpected = [k for k in state_dict if k not in known]
        if strict and unexpected:
            raise KeyError("unexpected keys in checkpoint, e.g. %s" % unexpected[0])
        self._seen_version = self.flat._version
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)

    def on_save_checkpoint(self, checkpoint):           # Lightning hook: trainer.save_checkpoint(...) (train.py:58)
        checkpoint["state_dict"] = self.adapter_state_dict()
        T = self.engine.T                               # also kept here for checkpoints written without optimizer_states
        checkpoint["vlb_optimizer"] = {"adam_m": T.m.detach().cpu(), "adam_v": T.v.detach().cpu(), "step_count": int(T.step_count)}

    def on_load_checkpoint(self, checkpoint):           # Lightning hook, runs before load_state_dict
        self.configure_model()
        vo = checkpoint.get("vlb_optimizer")
        if vo is not None:
            T = self.engine.T
            T.m.copy_(vo["adam_m"]); T.v.copy_(vo["adam_v"]); T.step_count = int(vo["step_count"])

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
        # anything other than FusedClipAdamW that wrote `flat` in place (a plain torch optimiser, a manual copy_) bumps the
        # tensor's version counter: refresh the bf16 working copy and the staged LoRA B matrices the kernels read
        if self.flat._version != self._seen_version:
            self.engine.T.sync_working()
            self._seen_version = self.flat._version

    # ------------------------------------------------------------------ forward / steps
    def forward(self, x_video, x_lang, weight_mask, attention_mask=None):
        """Reference signature and argument meaning (litmodule.py:229-256): x_video = [(Tensor[T,3,H,W], "video")] * B (or a
        stacked [B,T,3,H,W] tensor), x_lang = token ids [B, L_text] incl. the <video> id, weight_mask = HRF weight row
        [B, max_len] (make_weight_mask), attention_mask = x_lang.ne(0). Returns (regression_output [B,P], l2_reg).
        The attention mask carries no extra information here: padding is on the right, so under the causal mask no valid
        position can see a padded one, and padded positions have HRF weight 0. Inference only (dropout follows
        self.training); gradients flow through training_step."""
        self._refresh()
        dev = self.engine.dev
        if isinstance(x_video, (list, tuple)):
            x_video = torch.stack([v[0] if isinstance(v, (list, tuple)) else v for v in x_video], 0)
        vision = x_video.to(dev)
        if vision.dtype != torch.uint8:
            vision = vision.float()
        batch = {"vision": vision.contiguous(), "language": x_lang.to(dev).long(), "weight_mask": weight_mask.to(dev)}
        B = vision.shape[0]
        with torch.no_grad():
            hidden, wm = self.engine.backbone_forward(batch, save=False)
            y0 = self.engine.buf("fw_zero_targets", (B, self.engine.T.num_target_local), torch.float32)
            out, pred = self.engine.head_forward(hidden, wm, y0, B, training=self.training)
        return pred.clone(), out[2].clone()

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
