"""Model description, frozen-weight containers (bf16, HBM-resident, laid out for the kernels) and the flat buffer of
trainable tensors (LoRA A/B of the 7 decoder projections x L layers + brain head).

Layouts (row-major, K contiguous unless noted):
  ViT    : patch_w [Hv, kpad] (conv weight flattened (c,ky,kx), zero padded to a multiple of 8), fused wqkv [3Hv, Hv]
  STC    : 1x1 convs as [Cout, Cin]; depthwise 3x3 as [9, C]; Conv3d sampler as [C, 8C] with k = ((kt*2+kh)*2+kw)*C + c
  decoder: fused wqkv [(Hq+2Hkv)D, H], wo [H, H], fused wgu [2I, H] (gate rows then up rows), wd [H, I]
  LoRA   : A stacked per fused projection (Acat_qkv [3r, H], Acat_gu [2r, H]); B kept TRANSPOSED ([r, N]) in the fp32
           master so its gradient GEMM writes contiguous rows; a bf16 working copy in [N, r*k] block-diagonal form
           (Bcat) feeds the fused k-block of the forward GEMM.
Reference: load_pretrained_vllama2 / LoraConfig (src/litmodule/videollama2_vlb_litmodule.py:58-122).
"""
import math
from dataclasses import dataclass

import torch

from . import ops

VIDEO_TOKEN_ID = -201


@dataclass
class ModelConfig:
    vit_hidden: int = 1024
    vit_layers: int = 24
    vit_heads: int = 16
    vit_mlp: int = 4096
    image_size: int = 336
    patch_size: int = 14
    select_layer: int = -2
    vit_ln_eps: float = 1e-5
    hidden: int = 4096
    layers: int = 32
    heads: int = 32
    kv_heads: int = 8
    mlp: int = 14336
    vocab: int = 32000
    rms_eps: float = 1e-5
    rope_theta: float = 1e6
    stc_depth: int = 4
    stc_ln_eps: float = 1e-5
    num_frames: int = 8
    num_target: int = 1000
    max_len: int = 2048
    lora_r: int = 16
    lora_alpha: int = 32
    l2_lambda: float = 0.001
    head_ln_eps: float = 1e-5
    lang_w_len: int = 64
    use_lora: bool = True          # False: frozen-backbone baseline (VLB_vllama2_friends_baseline.yaml:26-27), head-only training

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
        return self.grid_out ** 2

    @property
    def n_vis(self):
        return self.t_out * self.tok_per_frame

    @property
    def l_text(self):
        return self.max_len - self.n_vis + 1

    @property
    def head_dim(self):
        return self.hidden // self.heads

    @property
    def vit_used_layers(self):
        # hidden_states[select_layer] with select_layer=-2 is the output of layer L-2 (0-based): L-1 layers needed
        return self.vit_layers + 1 + self.select_layer if self.select_layer < 0 else self.select_layer

    @property
    def kpad(self):
        return (3 * self.patch_size ** 2 + 7) // 8 * 8

    @classmethod
    def from_any(cls, other):
        names = cls.__dataclass_fields__.keys()
        return cls(**{k: getattr(other, k) for k in names if hasattr(other, k)})


def _bf(t, dev):
    return t.detach().to(device=dev, dtype=torch.bfloat16).contiguous()


def rope_tables(cfg, dev):
    """cos/sin [S, D/2] in bf16, computed the way HF MistralRotaryEmbedding does (fp32 angles, cast to bf16)."""
    D = cfg.head_dim
    inv_freq = 1.0 / (cfg.rope_theta ** (torch.arange(0, D, 2, dtype=torch.int64).float() / D))
    pos = torch.arange(cfg.max_len, dtype=torch.float32)
    freqs = pos[:, None] * inv_freq[None, :]
    return _bf(freqs.cos(), dev), _bf(freqs.sin(), dev)


class FrozenWeights:
    """Plain attribute bags of bf16 CUDA tensors (no nn.Module: nothing here is trainable)."""

    def __init__(self):
        self.vit = {}
        self.vit_layers = []
        self.stc = {}
        self.llm = {}
        self.llm_layers = []

    def nbytes(self):
        tot = 0
        for d in [self.vit, self.stc, self.llm] + self.vit_layers + self.llm_layers:
            for v in _flatten(d):
                tot += v.numel() * v.element_size()
        return tot


def _flatten(d):
    for v in d.values():
        if isinstance(v, torch.Tensor):
            yield v
        elif isinstance(v, dict):
            yield from _flatten(v)
        elif isinstance(v, list):
            for x in v:
                yield from _flatten(x)


# ------------------------------------------------------------------------------------------------ from the oracle
def oracle_state_dict(ref):
    """state_dict of an oracle RefVLB under the key names a VideoLLaMA2 checkpoint uses (weights.py docstring)."""
    sd = {"model.vision_tower.vision_tower." + k: v for k, v in ref.vision_tower.state_dict().items()}
    sd.update({"model.mm_projector." + k: v for k, v in ref.mm_projector.state_dict().items()})
    sd.update({"model." + k: v for k, v in ref.llm.state_dict().items() if "lora_" not in k})
    return sd


def frozen_from_oracle(ref, cfg, dev):
    """Convert an oracle RefVLB (HF CLIP / Mistral modules + restated STC) into kernel layouts through the same key-name
    path a real checkpoint takes (weights.load_frozen). Test-side helper: the oracle object is passed in, this module
    never imports it."""
    from .weights import load_frozen
    return load_frozen(oracle_state_dict(ref), cfg, dev)


# ------------------------------------------------------------------------------------------------ random init (benchmarks)
def _rand(shape, dev, std, gen):
    return (torch.randn(shape, device=dev, dtype=torch.float32, generator=gen) * std).to(torch.bfloat16)


def frozen_random(cfg, dev, seed=1234):
    """Random-init frozen weights generated directly in HBM (HF init: N(0, 0.02) linears, LN weight 1 / bias 0)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    ones = lambda n: torch.ones(n, device=dev, dtype=torch.bfloat16)
    zeros = lambda n: torch.zeros(n, device=dev, dtype=torch.bfloat16)
    W = FrozenWeights()
    Hv, Mv = cfg.vit_hidden, cfg.vit_mlp
    pw = _rand((Hv, cfg.kpad), dev, 0.02, g)
    pw[:, 3 * cfg.patch_size ** 2:] = 0
    W.vit = dict(patch_w=pw, cls=_rand((Hv,), dev, 0.02, g), pos=_rand((cfg.grid ** 2 + 1, Hv), dev, 0.02, g),
                 pre_w=ones(Hv), pre_b=zeros(Hv))
    for _ in range(cfg.vit_used_layers):
        W.vit_layers.append(dict(ln1_w=ones(Hv), ln1_b=zeros(Hv), wqkv=_rand((3 * Hv, Hv), dev, 0.02, g), bqkv=zeros(3 * Hv),
                                 wo=_rand((Hv, Hv), dev, 0.02, g), bo=zeros(Hv), ln2_w=ones(Hv), ln2_b=zeros(Hv),
                                 fc1_w=_rand((Mv, Hv), dev, 0.02, g), fc1_b=zeros(Mv),
                                 fc2_w=_rand((Hv, Mv), dev, 0.02, g), fc2_b=zeros(Hv)))
    C = cfg.hidden

    def block(cin):
        rd = int(round(cin * 0.25))
        d = dict(conv1_w=_rand((C, cin), dev, cin ** -0.5, g), n1_w=ones(C), n1_b=zeros(C),
                 dw_w=_rand((9, C), dev, 1.0 / 3, g), n2_w=ones(C), n2_b=zeros(C),
                 se1_w=_rand((rd, C), dev, C ** -0.5, g), se1_b=zeros(rd), se2_w=_rand((C, rd), dev, rd ** -0.5, g), se2_b=zeros(C),
                 conv3_w=_rand((C, C), dev, C ** -0.5, g), n3_w=ones(C), n3_b=zeros(C), ds_w=None)
        if cin != C:
            d.update(ds_w=_rand((C, cin), dev, cin ** -0.5, g), dsn_w=ones(C), dsn_b=zeros(C))
        return d

    W.stc = dict(s1=[block(cfg.vit_hidden if i == 0 else C) for i in range(cfg.stc_depth)],
                 s2=[block(C) for _ in range(cfg.stc_depth)],
                 sampler_w=_rand((C, 8 * C), dev, (8 * C) ** -0.5, g), sampler_b=zeros(C),
                 ro1_w=_rand((C, C), dev, C ** -0.5, g), ro1_b=zeros(C), ro2_w=_rand((C, C), dev, C ** -0.5, g), ro2_b=zeros(C))
    W.llm = dict(embed=_rand((cfg.vocab, C), dev, 0.02, g), norm_w=ones(C))
    W.llm["cos"], W.llm["sin"] = rope_tables(cfg, dev)
    D, I = cfg.head_dim, cfg.mlp
    for _ in range(cfg.layers):
        W.llm_layers.append(dict(ln1_w=ones(C), wqkv=_rand(((cfg.heads + 2 * cfg.kv_heads) * D, C), dev, 0.02, g),
                                 wo=_rand((C, C), dev, 0.02, g), ln2_w=ones(C),
                                 wgu=_rand((2 * I, C), dev, 0.02, g), wd=_rand((C, I), dev, 0.02, g)))
    return W


# ------------------------------------------------------------------------------------------------ trainables
class Trainables:
    """One flat fp32 master buffer (+ grad, Adam m/v, bf16 working copy) holding every trainable tensor, so that the
    data-parallel all-reduce, the global-norm clip and AdamW are each ONE pass over one contiguous buffer."""

    def __init__(self, cfg, dev, target_slice=None):
        """target_slice=(lo, hi): this rank owns readout rows [lo, hi) of the P voxels (head sharded by voxel, SURVEY 8e
        config 4); None = all of them."""
        self.cfg, self.dev = cfg, dev
        r, H, I, D = cfg.lora_r, cfg.hidden, cfg.mlp, cfg.head_dim
        nq, nkv = cfg.heads * D, cfg.kv_heads * D
        self.target_slice = (0, cfg.num_target) if target_slice is None else (int(target_slice[0]), int(target_slice[1]))
        self.lora_layers = cfg.layers if cfg.use_lora else 0
        self.spec = []
        for l in range(self.lora_layers):
            p = "layers.%d." % l
            self.spec += [(p + "Acat_qkv", (3 * r, H)), (p + "Bt_q", (r, nq)), (p + "Bt_k", (r, nkv)), (p + "Bt_v", (r, nkv)),
                          (p + "A_o", (r, nq)), (p + "Bt_o", (r, H)),
                          (p + "Acat_gu", (2 * r, H)), (p + "Bt_g", (r, I)), (p + "Bt_u", (r, I)),
                          (p + "A_d", (r, I)), (p + "Bt_d", (r, H))]
        P = self.num_target_local = self.target_slice[1] - self.target_slice[0]
        self.spec += [("ridge_w", (P, H)), ("ridge_b", (P,)), ("ln1_w", (H,)), ("ln1_b", (H,)), ("ln2_w", (H,)), ("ln2_b", (H,))]
        self.offsets, off = {}, 0
        for name, shape in self.spec:
            n = math.prod(shape)
            self.offsets[name] = (off, shape)
            off += (n + 7) // 8 * 8
        self.numel = off
        z = lambda dt: torch.zeros(off, device=dev, dtype=dt)
        self.param, self.grad, self.m, self.v = z(torch.float32), z(torch.float32), z(torch.float32), z(torch.float32)
        self.param_bf16 = z(torch.bfloat16)
        self.sumsq = torch.zeros(1, device=dev, dtype=torch.float32)
        self.step_count = 0
        # bf16 block-diagonal / transposed staging of the LoRA B matrices for the forward fused k-block
        self.bcat = []
        for l in range(self.lora_layers):
            self.bcat.append(dict(qkv=torch.zeros((nq + 2 * nkv, 3 * r), device=dev, dtype=torch.bfloat16),
                                  o=torch.zeros((H, r), device=dev, dtype=torch.bfloat16),
                                  gu=torch.zeros((2 * I, 2 * r), device=dev, dtype=torch.bfloat16),
                                  d=torch.zeros((H, r), device=dev, dtype=torch.bfloat16)))

    def view(self, buf, name):
        off, shape = self.offsets[name]
        return buf[off:off + math.prod(shape)].view(shape)

    def p(self, name):       # fp32 master
        return self.view(self.param, name)

    def w(self, name):       # bf16 working copy
        return self.view(self.param_bf16, name)

    def g(self, name):       # fp32 gradient
        return self.view(self.grad, name)

    def init_default(self, seed=1234, lora_b_std=0.0):
        """peft init (A kaiming-uniform(a=sqrt(5)), B = 0 unless lora_b_std) and torch defaults for the head."""
        g = torch.Generator(device=self.dev).manual_seed(seed)
        for name, shape in self.spec:
            t = self.p(name)
            leaf = name.split(".")[-1]
            if leaf.startswith("A"):
                bound = 1.0 / math.sqrt(shape[1])
                t.copy_((torch.rand(shape, device=self.dev, generator=g) * 2 - 1) * bound)
            elif leaf.startswith("Bt"):
                if lora_b_std > 0:
                    t.copy_(torch.randn(shape, device=self.dev, generator=g) * lora_b_std)
            elif leaf == "ridge_w" or leaf == "ridge_b":
                bound = 1.0 / math.sqrt(self.cfg.hidden)
                t.copy_((torch.rand(shape, device=self.dev, generator=g) * 2 - 1) * bound)
            elif leaf in ("ln1_w", "ln2_w"):
                t.fill_(1.0)
        self.sync_working()

    def load_from_oracle(self, ref):
        """Copy LoRA + head parameters of an oracle RefVLB into the flat master (test-side helper)."""
        for l, layer in enumerate(ref.llm.layers):
            if l >= self.lora_layers:
                break
            a, m = layer.self_attn, layer.mlp
            p = "layers.%d." % l
            self.p(p + "Acat_qkv").copy_(torch.cat([a.q_proj.lora_A.weight, a.k_proj.lora_A.weight, a.v_proj.lora_A.weight], 0))
            self.p(p + "Bt_q").copy_(a.q_proj.lora_B.weight.t())
            self.p(p + "Bt_k").copy_(a.k_proj.lora_B.weight.t())
            self.p(p + "Bt_v").copy_(a.v_proj.lora_B.weight.t())
            self.p(p + "A_o").copy_(a.o_proj.lora_A.weight)
            self.p(p + "Bt_o").copy_(a.o_proj.lora_B.weight.t())
            self.p(p + "Acat_gu").copy_(torch.cat([m.gate_proj.lora_A.weight, m.up_proj.lora_A.weight], 0))
            self.p(p + "Bt_g").copy_(m.gate_proj.lora_B.weight.t())
            self.p(p + "Bt_u").copy_(m.up_proj.lora_B.weight.t())
            self.p(p + "A_d").copy_(m.down_proj.lora_A.weight)
            self.p(p + "Bt_d").copy_(m.down_proj.lora_B.weight.t())
        lo, hi = self.target_slice
        self.p("ridge_w").copy_(ref.ridge.weight[lo:hi])
        self.p("ridge_b").copy_(ref.ridge.bias[lo:hi])
        self.p("ln1_w").copy_(ref.layer_norm1.weight); self.p("ln1_b").copy_(ref.layer_norm1.bias)
        self.p("ln2_w").copy_(ref.layer_norm2.weight); self.p("ln2_b").copy_(ref.layer_norm2.bias)
        self.sync_working()

    def sync_working(self):
        """fp32 master -> bf16 working copy -> staged Bcat matrices."""
        from . import lib as _l
        _l.call("vlb_f32_to_bf16", _l.ptr(self.param), _l.ptr(self.param_bf16), self.numel, _l.cur_stream())
        self.stage_b()

    def stage_b(self):
        """LoRA B^T [r, N] (bf16 working copy) -> the [N, r] column blocks of the staged Bcat matrices: one batched launch over
        a descriptor table built once (source and destination buffers never move)."""
        from . import lib as _l
        if not self.lora_layers:
            return
        if getattr(self, "_stage_desc", None) is None:
            cfg = self.cfg
            r, D = cfg.lora_r, cfg.head_dim
            nq, nkv, I = cfg.heads * D, cfg.kv_heads * D, cfg.mlp
            rows = []
            for l in range(self.lora_layers):
                p = "layers.%d." % l
                bc = self.bcat[l]
                for src, dst, n in ((self.w(p + "Bt_q"), bc["qkv"][0:nq, 0:r], nq), (self.w(p + "Bt_k"), bc["qkv"][nq:nq + nkv, r:2 * r], nkv),
                                    (self.w(p + "Bt_v"), bc["qkv"][nq + nkv:, 2 * r:3 * r], nkv), (self.w(p + "Bt_o"), bc["o"], cfg.hidden),
                                    (self.w(p + "Bt_g"), bc["gu"][0:I, 0:r], I), (self.w(p + "Bt_u"), bc["gu"][I:, r:2 * r], I),
                                    (self.w(p + "Bt_d"), bc["d"], cfg.hidden)):
                    assert src.shape == (r, n) and dst.shape == (n, r)
                    rows.append([src.data_ptr(), dst.data_ptr(), src.stride(0), dst.stride(0), r, n])
            self._stage_desc = torch.tensor(rows, dtype=torch.int64, device=self.param_bf16.device)
            self._stage_blocks = max(1, min(64, (r * max(I, nq) + 2047) // 2048))
        _l.call("vlb_transpose2d_batch", _l.ptr(self._stage_desc), self._stage_desc.shape[0], self._stage_blocks, _l.cur_stream())
