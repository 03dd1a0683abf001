"""Training-step engine: runs the phantom_vlb hot path (VLBLitModule.training_step + backward + optimiser step,
src/litmodule/videollama2_vlb_litmodule.py:259-306, 345-379) as an explicit sequence of libvlb_b200 kernel launches
on the current CUDA stream. PyTorch only owns the device buffers; every arithmetic op below is a C-ABI call.

Forward  : ViT frames (no grad) -> STC connector (no grad) -> <video> splice -> L decoder layers with LoRA fused
           into the base GEMMs -> final RMSNorm + LN1 + HRF pooling + LN2 -> readout GEMM -> MSE + ridge penalty.
Backward : head -> decoder layers in reverse (dX through the frozen weights, rank-r dA/dB written straight into the
           flat fp32 gradient buffer) -> stops at the spliced embeddings (tower, connector, embeddings are frozen).
"""
import math

import os

import torch

from . import lib as _l
from . import ops
from .model import ModelConfig, Trainables
from .ops import ACT_GELU, ACT_NONE, ACT_QUICK_GELU, ACT_SIGMOID, ACT_SILU


class Engine:
    def __init__(self, cfg: ModelConfig, frozen, trainables: Trainables = None, device="cuda", dropout_rate=0.0,
                 lora_dropout=0.0):
        _l.call("vlb_check_device")
        self.cfg, self.W, self.dev = cfg, frozen, torch.device(device)
        self.T = trainables if trainables is not None else Trainables(cfg, self.dev)
        self.scale = cfg.lora_alpha / cfg.lora_r
        self.dropout_rate = float(dropout_rate)
        # peft applies dropout to the adapter's INPUT: B(A(dropout(x))), one independent mask per adapter. With r = 16 the masks
        # are packed bits applied inside the kernels that consume the masked values (csrc/lora_dropout.cu: no masked activation
        # is ever written); other ranks fall back to materialised dropout(x) + a rank-r GEMM + masked add per projection input.
        self.lora_dropout = float(lora_dropout or 0.0)
        self.masked_lora = cfg.lora_r == 16 and os.environ.get("VLB_LORA_MASKED", "1") != "0"
        self.drop_bits = None       # test hook: dict filled with the packed keep masks keyed by (layer, site)
        self.drop_step = 0
        self.seed_mix = 0           # set_seed(seed, rank): every data-parallel rank draws its own dropout masks
        self.drop_masks = None      # test hook: dict filled with uint8 keep masks keyed by (layer, site)
        self.saved = None
        # wide readouts (config 3, P ~ 250k): never materialise the rank-B ridge gradient; clip norm from Gram matrices and
        # AdamW in the epilogue of the gradient GEMM (vlb_head_gradnorm / vlb_head_adamw_fused)
        self.fused_head_opt = self.T.num_target_local * cfg.hidden >= (1 << 26)
        self.shard_reduce = None    # callable(tensor): in-place sum across voxel shards of the readout head (config 4)
        self.row_gather = None      # callable(tensor) -> rows of all ranks (dp.gather_rows): replicated wide head under DP
        self.bcast0 = None          # callable(tensor): broadcast from rank 0 (same case: one clip norm for all replicas)
        # data parallel overlap: callable(flat gradient slice) invoked as soon as a slice of T.grad is final -- the head's
        # right after the head backward, then groups of `grad_chunk_layers` decoder layers as the backward leaves them
        # (dp.OverlappedGradReducer all-reduces each on a side stream, so only the last group is exposed)
        self.grad_ready = None
        self.grad_chunk_layers = 4
        self.launches = 0
        self.profile = None
        self.fuse_swiglu = os.environ.get("VLB_FUSE_SWIGLU", "1") != "0"
        self._bufs = {}

    # ------------------------------------------------------------------ buffers
    def buf(self, name, shape, dtype=torch.bfloat16):
        """Persistent workspace keyed by name (re-used across steps so TMA descriptors stay cached)."""
        key = (name, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, device=self.dev, dtype=dtype)
            self._bufs[key] = t
        return t

    def gemm(self, a, b, out, **kw):
        self.launches += 1
        if self.profile is None:
            return ops.gemm(a, b, out, **kw)
        # instrumented pass (bench.py roofline): CUDA events around this launch + its algorithmic FLOPs
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = ops.gemm(a, b, out, **kw)
        e1.record()
        M, N = out.shape
        K = a.shape[0] if kw.get("a_mn") else a.shape[1]
        a2 = kw.get("a2")
        K2 = 0 if a2 is None else (a2.shape[0] if kw.get("a_mn") else a2.shape[1])
        self.profile.append((e0, e1, 2.0 * M * N * (K + K2), (M, N, K, K2)))
        return r

    def _timed(self, flops, shape, fn):
        """Run one launch; in the instrumented pass (bench.py roofline) bracket it with CUDA events and record its algorithmic FLOPs."""
        self.launches += 1
        if self.profile is None:
            return fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn()
        e1.record()
        self.profile.append((e0, e1, flops, shape))
        return r

    def gateup(self, x, wgu, gu, act, a2, b2):
        """Gate/up projection + SwiGLU: one pair-GEMM launch with the activation in its epilogue when the intermediate size
        tiles by 128 (Mistral: 14336), else the GEMM followed by the row kernel."""
        if self.fuse_swiglu and act.shape[1] % 128 == 0 and x.shape[0] >= 256:
            self.launches += 1
            if self.profile is None:
                return ops.gemm_gateup_swiglu(x, wgu, gu, act, a2=a2, b2=b2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.gemm_gateup_swiglu(x, wgu, gu, act, a2=a2, b2=b2)
            e1.record()
            K2 = 0 if a2 is None else a2.shape[1]
            self.profile.append((e0, e1, 2.0 * gu.shape[0] * gu.shape[1] * (x.shape[1] + K2), (gu.shape[0], gu.shape[1], x.shape[1], K2)))
            return
        if a2 is not None:
            self.gemm(x, wgu, gu, a2=a2, b2=b2)
        else:
            self.gemm(x, wgu, gu)
        ops.swiglu_fwd(gu, out=act)
        self.launches += 1

    def down_dx_swiglu(self, dy, wd, gu, dgu, t, a_lora):
        self.launches += 1
        if self.profile is None:
            return ops.gemm_down_dx_swiglu_bwd(dy, wd, gu, dgu, a2=t, b2=a_lora)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemm_down_dx_swiglu_bwd(dy, wd, gu, dgu, a2=t, b2=a_lora)
        e1.record()
        M, H = dy.shape
        I, K2 = wd.shape[1], t.shape[1]
        self.profile.append((e0, e1, 2.0 * M * I * (H + K2), (M, I, H, K2)))

    # ------------------------------------------------------------------ vision tower (frozen, forward only)
    def vit_forward(self, pixels):
        """pixels fp32 (normalised) or uint8 (raw) [N, 3, H, W] -> patch features bf16 [N*G*G, Hv] (hidden_states[-2][:, 1:], encoder.py:31-53)."""
        cfg, W = self.cfg, self.W
        N = pixels.shape[0]
        G2, Hv, T = cfg.grid ** 2, cfg.vit_hidden, cfg.grid ** 2 + 1
        cols = ops.patch_im2col(pixels, cfg.patch_size, cfg.kpad, self.buf("vit_cols", (N * G2, cfg.kpad)))
        patch = self.gemm(cols, W.vit["patch_w"], self.buf("vit_patch", (N * G2, Hv)))
        h = ops.clip_embed(patch, W.vit["cls"], W.vit["pos"], W.vit["pre_w"], W.vit["pre_b"], N, T, cfg.vit_ln_eps,
                           self.buf("vit_h", (N * T, Hv)))
        y = self.buf("vit_y", (N * T, Hv))
        qkv = self.buf("vit_qkv", (N * T, 3 * Hv))
        o = self.buf("vit_o", (N * T, Hv))
        f = self.buf("vit_f", (N * T, cfg.vit_mlp))
        h2 = self.buf("vit_h2", (N * T, Hv))
        hd = Hv // cfg.vit_heads
        for L in W.vit_layers:
            ops.layernorm_fwd(h, L["ln1_w"], L["ln1_b"], cfg.vit_ln_eps, out=y)
            self.gemm(y, L["wqkv"], qkv, bias=L["bqkv"])
            ops.attn_fwd(qkv, N, T, cfg.vit_heads, cfg.vit_heads, hd, False, out=o, need_lse=False)
            self.gemm(o, L["wo"], h2, bias=L["bo"], residual=h)
            ops.layernorm_fwd(h2, L["ln2_w"], L["ln2_b"], cfg.vit_ln_eps, out=y)
            self.gemm(y, L["fc1_w"], f, bias=L["fc1_b"], act=ACT_QUICK_GELU)
            self.gemm(f, L["fc2_w"], h, bias=L["fc2_b"], residual=h2)
            self.launches += 3
        self.launches += 4
        return ops.drop_cls(h, N, T, self.buf("vit_feats", (N * G2, Hv)))

    # ------------------------------------------------------------------ STC connector (frozen, forward only)
    def _stc_block(self, x, blk, F, Hh, Ww, tag, idx):
        cfg = self.cfg
        C, HW, eps = cfg.hidden, Hh * Ww, cfg.stc_ln_eps
        M = F * HW
        a = self.gemm(x, blk["conv1_w"], self.buf(tag + "a", (M, C)))
        ops.layernorm_fwd(a, blk["n1_w"], blk["n1_b"], eps, act=ACT_SILU, out=a)
        b = ops.dwconv3x3(a, blk["dw_w"], F, Hh, Ww, self.buf(tag + "b", (M, C)))
        ops.layernorm_fwd(b, blk["n2_w"], blk["n2_b"], eps, act=ACT_SILU, out=b)
        pool = ops.se_pool(b, F, HW, self.buf(tag + "pool", (F, C)))
        rd = blk["se1_w"].shape[0]
        g1 = self.gemm(pool, blk["se1_w"], self.buf(tag + "g1_%d" % rd, (F, rd)), bias=blk["se1_b"], act=ACT_SILU)
        gate = self.gemm(g1, blk["se2_w"], self.buf(tag + "gate", (F, C)), bias=blk["se2_b"], act=ACT_SIGMOID)
        ops.se_scale(b, gate, F, HW, out=b)
        c = self.gemm(b, blk["conv3_w"], self.buf(tag + "c", (M, C)))
        if blk["ds_w"] is not None:
            sc = self.gemm(x, blk["ds_w"], self.buf(tag + "sc", (M, C)))
            ops.layernorm_fwd(sc, blk["dsn_w"], blk["dsn_b"], eps, out=sc)
        else:
            sc = x
        out = self.buf(tag + "out%d" % (idx & 1), (M, C))       # ping-pong: never aliases the block input
        ops.layernorm_fwd(c, blk["n3_w"], blk["n3_b"], eps, residual=sc, act=ACT_SILU, out=out)
        self.launches += 7
        return out

    def stc_forward(self, feats, B):
        """feats [B*T*G*G, Hv] -> visual tokens [B*Nv, H] (STCConnector.forward, projector.py:189-215)."""
        cfg, W = self.cfg, self.W
        T, G, C = cfg.num_frames, cfg.grid, cfg.hidden
        x = feats
        for i, blk in enumerate(W.stc["s1"]):
            x = self._stc_block(x, blk, B * T, G, G, "s1_", i)
        cols = ops.conv3d_gather(x, B, T, G, G, self.buf("stc_cols", (B * cfg.n_vis, 8 * C)))
        x = self.gemm(cols, W.stc["sampler_w"], self.buf("stc_samp", (B * cfg.n_vis, C)), bias=W.stc["sampler_b"], act=ACT_SILU)
        for i, blk in enumerate(W.stc["s2"]):
            x = self._stc_block(x, blk, B * cfg.t_out, cfg.grid_out, cfg.grid_out, "s2_", i)
        r1 = self.gemm(x, W.stc["ro1_w"], self.buf("stc_r1", (B * cfg.n_vis, C)), bias=W.stc["ro1_b"], act=ACT_GELU)
        self.launches += 1
        return self.gemm(r1, W.stc["ro2_w"], self.buf("stc_vis", (B * cfg.n_vis, C)), bias=W.stc["ro2_b"])

    # ------------------------------------------------------------------ decoder (LoRA fused), forward
    def _layer_bufs(self, l, M):
        cfg = self.cfg
        H, I, D, r = cfg.hidden, cfg.mlp, cfg.head_dim, cfg.lora_r
        nqkv = (cfg.heads + 2 * cfg.kv_heads) * D
        t = "L%d_" % l
        B = M // cfg.max_len
        return dict(
            x=self.buf(t + "x", (M, H)), rstd1=self.buf(t + "rstd1", (M,), torch.float32), xn1=self.buf(t + "xn1", (M, H)),
            xa1=self.buf(t + "xa1", (M, 3 * r)), qkv=self.buf(t + "qkv", (M, nqkv)), o=self.buf(t + "o", (M, cfg.heads * D)),
            lse=self.buf(t + "lse", (B, cfg.heads, cfg.max_len), torch.float32), xa2=self.buf(t + "xa2", (M, r)),
            h1=self.buf(t + "h1", (M, H)), rstd2=self.buf(t + "rstd2", (M,), torch.float32), xn2=self.buf(t + "xn2", (M, H)),
            xa3=self.buf(t + "xa3", (M, 2 * r)), gu=self.buf(t + "gu", (M, 2 * I)), act=self.buf(t + "act", (M, I)),
            xa4=self.buf(t + "xa4", (M, r)))

    def set_seed(self, seed=0, rank=0):
        self.seed_mix = ((int(seed) & 0xFFFFFFFF) * 0x9E3779B1 + (int(rank) + 1) * 0x85EBCA77) & 0xFFFFFFFFFFFF

    def _drop_seed(self, l, site, adapter=0):
        return 0xD0C0FFEE + self.seed_mix + (self.drop_step << 16) + (l * 4 + site) * 4 + adapter

    def _masked(self, training=True):
        return training and self.lora_dropout > 0 and self.masked_lora

    def _site_bits(self, l, site, n_ad, M, K, generate):
        """Packed keep masks of the n_ad adapters reading projection input `site` of layer l: int64 [n_ad, K/64, M] (kept until
        the backward of the same step, 1 bit per element)."""
        bits = self.buf("L%d_dropbits%d" % (l, site), (n_ad, K // 64, M), torch.int64)
        if generate:
            for a in range(n_ad):
                ops.dropout_bits(M, K, self.lora_dropout, self._drop_seed(l, site, a), bits[a])
            self.launches += n_ad
            if self.drop_bits is not None:
                self.drop_bits[(l, site)] = bits
        return [bits[a] for a in range(n_ad)]

    def _lora_down(self, x, A, xa, l, site, training):
        """xa = s * dropout(x) A^T for the adapters of one projection input (forward)."""
        n_ad = A.shape[0] // self.cfg.lora_r
        if self._masked(training):
            bits = self._site_bits(l, site, n_ad, x.shape[0], x.shape[1], True)
            ops.lora_down_masked(x, A, xa, bits, self.scale / (1.0 - self.lora_dropout))
            self.launches += 1
            return
        self.gemm(self._lora_in(x, l, site, training), A, xa, alpha=self.scale)

    def _lora_grad_a(self, t, x, gA, l, site):
        """dA += t^T dropout(x) (backward)."""
        if self._masked():
            bits = self._site_bits(l, site, gA.shape[0] // self.cfg.lora_r, x.shape[0], x.shape[1], False)
            ops.lora_grad_a_masked(x, t, gA, bits, 1.0 / (1.0 - self.lora_dropout))
            self.launches += 1
            return
        self.gemm(t, self._lora_in_bwd(x, l, site), gA, a_mn=True, b_mn=True, out_f32=True, accumulate=True)

    def _lora_in(self, x, l, site, training):
        """Input of the LoRA A projection: x itself, or dropout(x) into a scratch buffer (regenerated in the backward)."""
        if not (training and self.lora_dropout > 0):
            return x
        xd = self.buf("lora_xd_%d" % x.shape[1], tuple(x.shape))
        mask = None
        if self.drop_masks is not None:
            mask = torch.empty(tuple(x.shape), device=x.device, dtype=torch.uint8)
            self.drop_masks[(l, site)] = mask
        ops.dropout(x, self.lora_dropout, self._drop_seed(l, site), out=xd, mask=mask)
        self.launches += 1
        return xd

    def _lora_in_bwd(self, x, l, site):
        if self.lora_dropout <= 0:
            return x
        xd = self.buf("lora_xd_%d" % x.shape[1], tuple(x.shape))
        ops.dropout(x, self.lora_dropout, self._drop_seed(l, site), out=xd)
        self.launches += 1
        return xd

    def _dx_with_lora(self, dy, w_base, t, a_lora, out, l, site):
        """out = dy @ W (frozen) + adapter branch (t @ A): one GEMM with the fused k-block, or -- with LoRA dropout -- the
        base GEMM, a rank-r GEMM and a masked add (the adapter's input gradient passes through the dropout mask)."""
        if self.lora_dropout <= 0:
            return self.gemm(dy, w_base, out, a2=t, b2=a_lora, b_mn=True)
        if self._masked() and out.shape[1] % 64 == 0:
            bits = self._site_bits(l, site, a_lora.shape[0] // self.cfg.lora_r, out.shape[0], out.shape[1], False)
            M, N, K = out.shape[0], out.shape[1], dy.shape[1]
            return self._timed(2.0 * M * N * (K + t.shape[1]), (M, N, K, t.shape[1]),
                               lambda: ops.gemm_dx_lora_masked(dy, w_base, out, t, a_lora, bits, 1.0 / (1.0 - self.lora_dropout)))
        self.gemm(dy, w_base, out, b_mn=True)
        tmp = self.buf("lora_tmp_%d" % out.shape[1], tuple(out.shape))
        self.gemm(t, a_lora, tmp, b_mn=True)
        ops.dropout_add(tmp, out, self.lora_dropout, self._drop_seed(l, site))
        self.launches += 1
        return out

    def decoder_forward(self, x0, B, save=True):
        """x0 [B*S, H] (buffer L0_x) -> pre-final-norm hidden [B*S, H]."""
        cfg, W, T = self.cfg, self.W, self.T
        S, M, D, s = cfg.max_len, B * cfg.max_len, cfg.head_dim, self.scale
        lora = cfg.use_lora
        assert lora or not save, "a frozen backbone (use_lora=False) has no decoder backward: call with save=False"
        x = x0
        for l, L in enumerate(W.llm_layers):
            b = self._layer_bufs(l if save else 0, M)
            p = "layers.%d." % l
            bc = T.bcat[l] if lora else None
            nxt = self.buf("L%d_x" % (l + 1), (M, cfg.hidden)) if save else self.buf("Lx_%d" % (l & 1), (M, cfg.hidden))
            ops.rmsnorm_fwd(x, L["ln1_w"], cfg.rms_eps, out=b["xn1"], rstd=b["rstd1"])
            if lora:
                self._lora_down(b["xn1"], T.w(p + "Acat_qkv"), b["xa1"], l, 0, save)
            # q/k/v projection (+ fused LoRA k-block) with RoPE in the GEMM epilogue (head_dim 128; other widths: GEMM, then rope)
            xa1, bq = (b["xa1"], bc["qkv"]) if lora else (None, None)
            Nq = L["wqkv"].shape[0]
            self._timed(2.0 * M * Nq * (cfg.hidden + (xa1.shape[1] if lora else 0)), (M, Nq, cfg.hidden, xa1.shape[1] if lora else 0),
                        lambda: ops.gemm_qkv_rope(b["xn1"], L["wqkv"], b["qkv"], W.llm["cos"], W.llm["sin"], S,
                                                  cfg.heads + cfg.kv_heads, D, a2=xa1, b2=bq))
            ops.attn_fwd_into(b["qkv"], b["o"], b["lse"], B, S, cfg.heads, cfg.kv_heads, D, True)
            if lora:
                self._lora_down(b["o"], T.w(p + "A_o"), b["xa2"], l, 1, save)
                self.gemm(b["o"], L["wo"], b["h1"], a2=b["xa2"], b2=bc["o"], residual=x)
            else:
                self.gemm(b["o"], L["wo"], b["h1"], residual=x)
            ops.rmsnorm_fwd(b["h1"], L["ln2_w"], cfg.rms_eps, out=b["xn2"], rstd=b["rstd2"])
            if lora:
                self._lora_down(b["xn2"], T.w(p + "Acat_gu"), b["xa3"], l, 2, save)
            self.gateup(b["xn2"], L["wgu"], b["gu"], b["act"], b["xa3"] if lora else None, bc["gu"] if lora else None)
            if lora:
                self._lora_down(b["act"], T.w(p + "A_d"), b["xa4"], l, 3, save)
                self.gemm(b["act"], L["wd"], nxt, a2=b["xa4"], b2=bc["d"], residual=b["h1"])
            else:
                self.gemm(b["act"], L["wd"], nxt, residual=b["h1"])
            self.launches += 3 + (0 if D == 128 else 1)  # rmsnorm x2, attention (+ the stand-alone rope when it is not fused)
            x = nxt
        return x

    # ------------------------------------------------------------------ decoder backward
    def decoder_backward(self, dx, B):
        """dx: gradient wrt the pre-final-norm hidden [B*S, H] (bf16). Fills LoRA gradients in T.grad."""
        cfg, W, T = self.cfg, self.W, self.T
        S, M, D, s, r = cfg.max_len, B * cfg.max_len, cfg.head_dim, self.scale, cfg.lora_r
        H, I = cfg.hidden, cfg.mlp
        nq, nkv = cfg.heads * D, cfg.kv_heads * D
        dact = self.buf("bw_dact", (M, I))
        dgu = self.buf("bw_dgu", (M, 2 * I))
        dxn = self.buf("bw_dxn", (M, H))
        dh1 = self.buf("bw_dh1", (M, H))
        do = self.buf("bw_do", (M, nq))
        dqkv = self.buf("bw_dqkv", (M, nq + 2 * nkv))
        dq32 = self.buf("bw_dq32", (M, nq), torch.float32)
        delta = self.buf("bw_delta", (B, cfg.heads, S), torch.float32)
        t48 = self.buf("bw_t48", (M, 3 * r))
        t32 = self.buf("bw_t32", (M, 2 * r))
        t16 = self.buf("bw_t16", (M, r))
        dxa = self.buf("bw_dx_a", (M, H))
        dxb = self.buf("bw_dx_b", (M, H))
        pending_hi = cfg.layers                             # layers [l, pending_hi) have final gradients not yet handed to grad_ready
        for l in range(cfg.layers - 1, -1, -1):
            L, b, bc = W.llm_layers[l], self._layer_bufs(l, M), T.bcat[l]
            p = "layers.%d." % l
            x = self.buf("L%d_x" % l, (M, H))
            # ---- down_proj: x_next = h1 + act Wd^T + s (act A_d^T) B_d^T
            self.gemm(dx, bc["d"], t16, b_mn=True, alpha=s)
            masked = self._masked()
            fused_bwd = self.fuse_swiglu and (self.lora_dropout <= 0 or masked) and I % 256 == 0 and M >= 256
            if fused_bwd and masked:                        # the adapter's masked input gradient is added in the same epilogue
                bits3 = self._site_bits(l, 3, 1, M, I, False)[0]
                self._timed(2.0 * M * I * (H + r), (M, I, H, r),
                            lambda: ops.gemm_down_dx_swiglu_bwd_masked(dx, L["wd"], b["gu"], dgu, t16, T.w(p + "A_d"), bits3,
                                                                       1.0 / (1.0 - self.lora_dropout)))
            elif fused_bwd:                                 # d(act) stays in TMEM; dgu leaves the GEMM epilogue
                self.down_dx_swiglu(dx, L["wd"], b["gu"], dgu, t16, T.w(p + "A_d"))
            else:
                self._dx_with_lora(dx, L["wd"], t16, T.w(p + "A_d"), dact, l, 3)
            self._lora_grad_a(t16, b["act"], T.g(p + "A_d"), l, 3)
            self.gemm(b["xa4"], dx, T.g(p + "Bt_d"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            if not fused_bwd:
                ops.swiglu_bwd(b["gu"], dact, out=dgu)
            # ---- gate/up
            self.gemm(dgu, bc["gu"], t32, b_mn=True, alpha=s)
            self._dx_with_lora(dgu, L["wgu"], t32, T.w(p + "Acat_gu"), dxn, l, 2)
            self._lora_grad_a(t32, b["xn2"], T.g(p + "Acat_gu"), l, 2)
            self.gemm(b["xa3"][:, 0:r], dgu[:, 0:I], T.g(p + "Bt_g"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            self.gemm(b["xa3"][:, r:2 * r], dgu[:, I:2 * I], T.g(p + "Bt_u"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            ops.rmsnorm_bwd(dxn, b["h1"], L["ln2_w"], b["rstd2"], dres=dx, out=dh1)
            # ---- o_proj: h1 = x + o Wo^T + s (o A_o^T) B_o^T
            self.gemm(dh1, bc["o"], t16, b_mn=True, alpha=s)
            self._dx_with_lora(dh1, L["wo"], t16, T.w(p + "A_o"), do, l, 1)
            self._lora_grad_a(t16, b["o"], T.g(p + "A_o"), l, 1)
            self.gemm(b["xa2"], dh1, T.g(p + "Bt_o"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            # ---- attention + RoPE
            ops.attn_bwd(b["qkv"], b["o"], do, b["lse"], B, S, cfg.heads, cfg.kv_heads, D, True, dqkv=dqkv, dq_f32=dq32,
                         delta=delta)
            ops.rope_bwd_f32(dq32, dqkv, W.llm["cos"], W.llm["sin"], S, cfg.heads, D)
            ops.rope_(dqkv[:, nq:], W.llm["cos"], W.llm["sin"], S, cfg.kv_heads, D, inverse=True)
            # ---- q/k/v
            self.gemm(dqkv, bc["qkv"], t48, b_mn=True, alpha=s)
            self._lora_grad_a(t48, b["xn1"], T.g(p + "Acat_qkv"), l, 0)
            self.gemm(b["xa1"][:, 0:r], dqkv[:, 0:nq], T.g(p + "Bt_q"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            self.gemm(b["xa1"][:, r:2 * r], dqkv[:, nq:nq + nkv], T.g(p + "Bt_k"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            self.gemm(b["xa1"][:, 2 * r:3 * r], dqkv[:, nq + nkv:], T.g(p + "Bt_v"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            self.launches += 6 if fused_bwd else 7
            # chunks of grad_chunk_layers layers, except that the last chunk is split into layers [1, chunk) and layer 0 alone: the
            # reduce launched after layer 0 is the one nothing overlaps, so it should be as small as possible
            if self.grad_ready is not None and (l % self.grad_chunk_layers == 0 or l == 1):
                hi_l = pending_hi
                pending_hi = l
                lo = T.offsets["layers.%d.Acat_qkv" % l][0]
                hi = T.offsets["layers.%d.Acat_qkv" % hi_l][0] if hi_l < cfg.layers else T.offsets["ridge_w"][0]
                # every kernel that writes these layers' slice of T.grad has been enqueued above (the q/k/v input-gradient GEMM
                # and the RMSNorm backward below do not touch it)
                self.grad_ready(T.grad[lo:hi])
            if l > 0:        # the spliced embeddings are frozen: layer 0 needs no input gradient
                self._dx_with_lora(dqkv, L["wqkv"], t48, T.w(p + "Acat_qkv"), dxn, l, 0)
                nxt = dxa if dx.data_ptr() != dxa.data_ptr() else dxb
                ops.rmsnorm_bwd(dxn, x, L["ln1_w"], b["rstd1"], dres=dh1, out=nxt)
                self.launches += 1
                dx = nxt

    # ------------------------------------------------------------------ full forward (backbone)
    def backbone_forward(self, batch, save=True):
        """batch: device tensors of the datamodule contract (src/datamodule/videollama2_vlb_datamodule.py:98-109).
        Returns (pre-final-norm hidden [B*S, H], HRF weight mask fp32 [B, S])."""
        cfg = self.cfg
        vision = batch["vision"]
        B, Tn = vision.shape[0], vision.shape[1]
        assert vision.dtype in (torch.float32, torch.uint8)                               # uint8: raw frames (f-1)
        if Tn != cfg.num_frames:
            # the frame count is a property of the data (window * frames_per_tr, extractfeatures.py:198-212): the reference
            # passes whatever T the datamodule holds, so the engine follows the batch (buffers are keyed by shape)
            cfg.num_frames = Tn
        assert batch["language"].shape[1] == cfg.l_text, "text length must be max_len - n_vis + 1 (litmodule.py:180-181)"
        feats = self.vit_forward(vision.reshape(B * Tn, 3, cfg.image_size, cfg.image_size))
        vis = self.stc_forward(feats, B)
        lang = batch["language"].contiguous()
        ids = lang if lang.dtype == torch.int64 else ops.ids_to_long(lang)
        x0 = self.buf("L0_x", (B * cfg.max_len, cfg.hidden))
        _, self.video_pos = ops.splice_embed(ids, self.W.llm["embed"], vis, cfg.n_vis, out=x0)
        if batch.get("weight_mask") is not None:          # forward(x_video, x_lang, weight_mask): caller-built row
            wm = batch["weight_mask"].to(torch.float32).contiguous()
            assert wm.shape == (B, cfg.max_len)
        else:
            wm = ops.weight_mask(batch["padvals"].contiguous(), batch["vis_weights"].contiguous(),
                                 batch["lang_weights"].contiguous(), cfg.tok_per_frame, cfg.max_len)
        self.launches += 4
        self.vis_tokens, self.embeds, self.frame_feats, self.weight_mask = vis, x0, feats, wm
        hidden = self.decoder_forward(x0, B, save=save)
        return hidden, wm

    # ------------------------------------------------------------------ brain head
    def _nch(self):
        S = self.cfg.max_len
        return 64 if S >= 1024 else max(1, S // 16)

    def _pool_partial(self, hidden, wm, B, unit_affine=False):
        """final RMSNorm + LN1 + HRF pooling of [B*S, H] -> fp32 partial sums [B, nch, H] (one read of `hidden`)."""
        cfg, T = self.cfg, self.T
        H, S, nch = cfg.hidden, cfg.max_len, self._nch()
        partial = self.buf("hd_partial", (B, nch, H), torch.float32)
        if unit_affine:         # feature extraction: LN1 affine factored out of the pooling (see vlb_pool_reduce)
            g1 = self.buf("hd_ones", (H,)); b1 = self.buf("hd_zeros", (H,))
            g1.fill_(1.0); b1.zero_()
        else:
            g1, b1 = T.w("ln1_w"), T.w("ln1_b")
        _l.call("vlb_hrf_pool_fwd", _l.ptr(hidden), _l.ptr(self.W.llm["norm_w"]), _l.ptr(g1), _l.ptr(b1),
                _l.ptr(wm), _l.ptr(partial), B, S, H, nch, cfg.rms_eps, cfg.head_ln_eps, _l.cur_stream())
        self.launches += 1
        return partial, nch

    def head_forward(self, hidden, wm, y, B, training, partial=None, nch=None):
        """LN2 -> dropout -> readout -> MSE + ridge penalty. `y` is this rank's [B, P_local] target slice; with a
        voxel-sharded head (`self.shard_reduce` set) the loss partial sums are summed across ranks before finalize."""
        cfg, T = self.cfg, self.T
        H, P = cfg.hidden, T.num_target_local
        if partial is None:
            partial, nch = self._pool_partial(hidden, wm, B)
        pooled, hrf = self.buf("hd_pooled", (B, H)), self.buf("hd_hrf", (B, H))
        drop = self.dropout_rate if training else 0.0
        keep = self.buf("hd_keep", (B, H), torch.float32) if drop > 0 else None
        self.rng_step = getattr(self, "rng_step", 0) + 1
        _l.call("vlb_hrf_pool_finalize", _l.ptr(partial), _l.ptr(T.w("ln2_w")), _l.ptr(T.w("ln2_b")), _l.ptr(pooled),
                _l.ptr(hrf), _l.ptr(keep), B, H, nch, cfg.head_ln_eps, float(drop), 0x5EED0000 + self.seed_mix + self.rng_step, _l.cur_stream())
        pred = self.gemm(hrf, T.w("ridge_w"), self.buf("hd_pred", (B, P)), bias=T.w("ridge_b"))
        acc = self.buf("hd_acc", (2,), torch.float32)
        acc.zero_()
        dpred = self.buf("hd_dpred", (B, P))
        out = self.buf("hd_loss", (3,), torch.float32)
        n_total = B * cfg.num_target
        assert y.shape == (B, P) and y.dtype == torch.float32
        _l.call("vlb_head_loss", _l.ptr(pred), _l.ptr(y), _l.ptr(dpred), _l.ptr(acc), B * P, n_total, _l.cur_stream())
        _l.call("vlb_l2_prefill", _l.ptr(T.w("ridge_w")), _l.ptr(T.g("ridge_w")), _l.ptr(acc), P * H, cfg.l2_lambda,
                int(training and not self.fused_head_opt), _l.cur_stream())
        if self.fused_head_opt:
            self.buf("hd_wsq_local", (1,), torch.float32).copy_(acc[1:2])      # this shard's sum of W^2 (before the reduce)
        if self.shard_reduce is not None:
            self.shard_reduce(acc)
        _l.call("vlb_loss_finalize", _l.ptr(acc), _l.ptr(out), n_total, cfg.l2_lambda, _l.cur_stream())
        self.launches += 4
        self.head_saved = dict(pooled=pooled, hrf=hrf, keep=keep, dpred=dpred, pred=pred, nch=nch)
        return out, pred

    def _head_backward_to_pooled(self, B):
        """ridge W/b and LN2 gradients into T.grad; returns d(loss)/d(pooled) fp32 [B, H]."""
        cfg, T, hs = self.cfg, self.T, self.head_saved
        H, P = cfg.hidden, T.num_target_local
        if not self.fused_head_opt:
            self.gemm(hs["dpred"], hs["hrf"], T.g("ridge_w"), a_mn=True, b_mn=True, out_f32=True, accumulate=True)
        _l.call("vlb_colsum", _l.ptr(hs["dpred"]), _l.ptr(T.g("ridge_b")), B, P, _l.cur_stream())
        dhrf = self.buf("hd_dhrf", (B, H), torch.float32)
        dhrf.zero_()            # accumulate mode lets the GEMM split the P-long contraction over all SMs (red.add)
        self.gemm(hs["dpred"], T.w("ridge_w"), dhrf, b_mn=True, out_f32=True, accumulate=True)
        if self.shard_reduce is not None:
            self.shard_reduce(dhrf)         # each voxel shard contributes a partial d(hrf); LN grads end up replicated
        dpooled = self.buf("hd_dpooled", (B, H), torch.float32)
        _l.call("vlb_ln2_bwd", _l.ptr(dhrf), _l.ptr(hs["keep"]), _l.ptr(hs["pooled"]), _l.ptr(T.w("ln2_w")), _l.ptr(dpooled),
                _l.ptr(T.g("ln2_w")), _l.ptr(T.g("ln2_b")), B, H, cfg.head_ln_eps, _l.cur_stream())
        self.launches += 2
        return dpooled

    def head_backward(self, hidden, wm, B, need_dx=True):
        """Gradients of ridge W/b, LN1/LN2 into T.grad; returns d(loss)/d(pre-final-norm hidden) [B*S, H] bf16
        (None when need_dx is False: frozen backbone, only the LN1 gradients are accumulated)."""
        cfg, T, hs = self.cfg, self.T, self.head_saved
        H, S = cfg.hidden, cfg.max_len
        dpooled = self._head_backward_to_pooled(B)
        dx = self.buf("bw_dx_a", (B * S, H)) if need_dx else None
        _l.call("vlb_hrf_pool_bwd", _l.ptr(hidden), _l.ptr(self.W.llm["norm_w"]), _l.ptr(T.w("ln1_w")), _l.ptr(wm),
                _l.ptr(dpooled), _l.ptr(dx), _l.ptr(T.g("ln1_w")), _l.ptr(T.g("ln1_b")), B, S, H, hs["nch"], cfg.rms_eps,
                cfg.head_ln_eps, _l.cur_stream())
        self.launches += 1
        return dx

    def _zero_grads(self):
        """Every gradient kernel ACCUMULATES (split-K red.add GEMMs, LN atomics). With the fused wide-head optimiser the
        ridge gradient is never written, so its (multi-GB) slice of the flat buffer is not touched at all."""
        T = self.T
        if not self.fused_head_opt:
            T.grad.zero_()
            return
        off_w, shape_w = T.offsets["ridge_w"]
        end_w = off_w + (math.prod(shape_w) + 7) // 8 * 8
        T.grad[:off_w].zero_()
        T.grad[end_w:].zero_()

    # ------------------------------------------------------------------ steps
    def loss_and_grads(self, batch):
        """One forward + backward of training_step (litmodule.py:259-306). Returns (loss[3] fp32 device tensor, pred).
        LoRA mode back-propagates through the decoder; the frozen-backbone baseline (use_lora=False, reference
        freeze_backbone=True litmodule.py:86-90) keeps no activations and stops at the head."""
        B = batch["vision"].shape[0]
        lora = self.cfg.use_lora
        self.drop_step += 1          # fresh LoRA dropout masks every training step (same masks in its backward)
        self._zero_grads()
        hidden, wm = self.backbone_forward(batch, save=lora)
        self.hidden_pre_norm = hidden
        y = batch["timeseries"].contiguous()
        out, pred = self.head_forward(hidden, wm, y, B, training=True)
        dx = self.head_backward(hidden, wm, B, need_dx=lora)
        if self.grad_ready is not None and not self.fused_head_opt:
            self.grad_ready(self.T.grad[self.T.offsets["ridge_w"][0]:])     # ridge W/b, LN1, LN2: final before the decoder backward
        if lora:
            self.decoder_backward(dx, B)
        return out, pred

    def eval_step(self, batch):
        """validation_step (litmodule.py:309-342): loss[3], brain_preds bf16 [B,P]."""
        B = batch["vision"].shape[0]
        hidden, wm = self.backbone_forward(batch, save=False)
        return self.head_forward(hidden, wm, batch["timeseries"].contiguous(), B, training=False)

    # ------------------------------------------------------------------ frozen-backbone feature cache
    def extract_features(self, batch, out_u=None, out_sw=None):
        """Backbone forward only -> (u fp32 [B, H], sw fp32 [B]) with u = sum_s w_s LN1hat(RMSNorm(h_s)), sw = sum_s w_s.
        While the backbone is frozen these never change, so the head can train from them without re-running the
        34-39 TF/clip forward every epoch (the reference recomputes it: VLB_vllama2_friends_baseline.yaml:26-27)."""
        cfg = self.cfg
        B, H, S = batch["vision"].shape[0], cfg.hidden, cfg.max_len
        hidden, wm = self.backbone_forward(batch, save=False)
        partial, nch = self._pool_partial(hidden, wm, B, unit_affine=True)
        u = out_u if out_u is not None else torch.empty((B, H), device=self.dev, dtype=torch.float32)
        sw = out_sw if out_sw is not None else torch.empty((B,), device=self.dev, dtype=torch.float32)
        _l.call("vlb_pool_reduce", _l.ptr(partial), _l.ptr(wm), _l.ptr(u), _l.ptr(sw), B, S, H, nch, _l.cur_stream())
        self.launches += 1
        return u, sw

    def cached_loss_and_grads(self, u, sw, y, training=True):
        """Head-only train step from cached features: pooled = g1*u + b1*sw -> LN2 -> dropout -> readout -> loss, and the
        ridge / LN1 / LN2 gradients. u [B, H] fp32, sw [B] fp32, y [B, P_local] fp32 (all B samples, this rank's voxels)."""
        cfg, T = self.cfg, self.T
        B, H = u.shape
        if training:
            self._zero_grads()
        pre = self.buf("hd_partial1", (B, 1, H), torch.float32)
        _l.call("vlb_ln1_affine_fwd", _l.ptr(u), _l.ptr(sw), _l.ptr(T.w("ln1_w")), _l.ptr(T.w("ln1_b")), _l.ptr(pre), B, H,
                _l.cur_stream())
        out, pred = self.head_forward(None, None, y.contiguous(), B, training=training, partial=pre, nch=1)
        self.launches += 1
        if training:
            dpooled = self._head_backward_to_pooled(B)
            _l.call("vlb_ln1_affine_bwd", _l.ptr(dpooled), _l.ptr(u), _l.ptr(sw), _l.ptr(T.g("ln1_w")), _l.ptr(T.g("ln1_b")),
                    B, H, _l.cur_stream())
            self.launches += 1
        return out, pred

    def optimizer_step(self, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_norm=1.0, world_size=1,
                       allreduce=None):
        """clip_grad_norm_(max_norm) + AdamW on the flat buffer; `allreduce(tensor)` sums gradients across ranks first (pass
        None with world_size > 1 when the slices were already summed through `grad_ready`)."""
        T = self.T
        fused = self.fused_head_opt
        off_w, shape_w = T.offsets["ridge_w"]
        n_w = math.prod(shape_w)
        end_w = off_w + (n_w + 7) // 8 * 8
        dp_rows = None
        if allreduce is not None:
            if fused and self.shard_reduce is None:
                # replicated wide head under data parallel: its rank-B gradient is never materialised, so there is nothing to
                # all-reduce for it -- the ranks exchange the B rows of dpred / hrf / pred instead (sum over ranks of
                # dpred_r^T hrf_r == gathered^T gathered) and every rank forms the same global gradient tile by tile
                if self.row_gather is None:
                    raise RuntimeError("fused wide-head optimiser with a replicated head under data parallel needs "
                                       "engine.row_gather (dp.gather_rows); refusing to step with rank-local gradients")
                hs = self.head_saved
                dp_rows = tuple(self.row_gather(hs[k]) for k in ("dpred", "hrf", "pred"))
                allreduce(T.grad[:off_w])
                allreduce(T.grad[end_w:])
            else:
                allreduce(T.grad)
        T.step_count += 1
        hyper = (float(lr), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), T.step_count, float(max_norm),
                 1.0 / world_size)
        st = _l.cur_stream()
        T.sumsq.zero_()
        gp = T.grad.data_ptr()
        if fused:
            hs = self.head_saved
            h_dpred, h_hrf, h_pred = dp_rows if dp_rows is not None else (hs["dpred"], hs["hrf"], hs["pred"])
            # gradients are kept as SUMS over ranks and scaled by 1/world inside the kernels; the ridge term 2*lambda*W each
            # rank would have contributed is therefore counted world times
            lam = self.cfg.l2_lambda * (world_size if dp_rows is not None else 1)

        def sumsq(lo, hi):                      # squared norm of the materialised gradients in [lo, hi), ridge_w skipped if fused
            for a, b in (((lo, min(hi, off_w)), (max(lo, end_w), hi)) if fused else ((lo, hi),)):
                if b > a:
                    _l.call("vlb_grad_sumsq", gp + 4 * a, b - a, _l.ptr(T.sumsq), st)
                    self.launches += 1

        split = T.offsets["ln1_w"][0] if self.shard_reduce is not None else T.numel
        sumsq(0, split)
        if fused:                               # ||dW_ridge||^2 from Gram matrices: the 4 GB gradient is never formed
            N = h_dpred.shape[0]
            N8 = (N + 7) // 8 * 8               # GEMM output columns come in multiples of 8: zero-padded sample rows
            dp, hr = h_dpred, h_hrf
            if N8 != N:
                dp = self.buf("hd_dpred_pad", (N8, dp.shape[1])); hr = self.buf("hd_hrf_pad", (N8, hr.shape[1]))
                dp.zero_(); hr.zero_()
                dp[:N].copy_(h_dpred); hr[:N].copy_(h_hrf)
            gd = self.buf("hd_gram_d", (N8, N8), torch.float32); gh = self.buf("hd_gram_h", (N8, N8), torch.float32)
            gd.zero_(); gh.zero_()
            self.gemm(dp, dp, gd, out_f32=True, accumulate=True)
            self.gemm(hr, hr, gh, out_f32=True, accumulate=True)
            _l.call("vlb_head_gradnorm", _l.ptr(h_dpred), _l.ptr(h_pred), _l.ptr(T.w("ridge_b")), _l.ptr(gd), _l.ptr(gh),
                    N8, _l.ptr(self.buf("hd_wsq_local", (1,), torch.float32)), _l.ptr(T.sumsq), N, T.num_target_local,
                    lam, st)
            self.launches += 1
        if dp_rows is not None and self.bcast0 is not None:
            # the Gram GEMMs above are split-K (red.add: summation order varies), so each replica's norm differs in its last bits;
            # rank 0's value is used everywhere to keep the replicas bit-identical
            self.bcast0(T.sumsq)
        if self.shard_reduce is not None:
            # voxel-sharded head: ridge W/b gradients are rank-private (sum their squares across ranks), the LN1/LN2
            # gradients that follow them in the flat buffer are replicated (count them once)
            self.shard_reduce(T.sumsq)
            sumsq(split, T.numel)

        def adamw(lo, hi):
            if hi > lo:
                _l.call("vlb_adamw_step", T.param.data_ptr() + 4 * lo, gp + 4 * lo, T.m.data_ptr() + 4 * lo,
                        T.v.data_ptr() + 4 * lo, T.param_bf16.data_ptr() + 2 * lo, hi - lo, _l.ptr(T.sumsq), *hyper, st)
                self.launches += 1

        if fused:
            adamw(0, off_w)
            _l.call("vlb_head_adamw_fused", _l.ptr(h_dpred), h_dpred.stride(0), _l.ptr(h_hrf), h_hrf.stride(0),
                    h_dpred.shape[0], T.num_target_local, self.cfg.hidden, T.param.data_ptr() + 4 * off_w,
                    T.m.data_ptr() + 4 * off_w, T.v.data_ptr() + 4 * off_w, T.param_bf16.data_ptr() + 2 * off_w,
                    _l.ptr(T.sumsq), *hyper, lam, st)
            self.launches += 1
            adamw(end_w, T.numel)
        else:
            adamw(0, T.numel)
        T.stage_b()
        self.launches += 1 if T.lora_layers else 0

    def train_step(self, batch, lr, **opt):
        out, _ = self.loss_and_grads(batch)
        self.optimizer_step(lr, **opt)
        return out


def cosine_lr(base_lr, step, t_max, eta_min=0.0):
    """torch CosineAnnealingLR closed form (litmodule.py:365-378), value used for optimiser step `step` (0-based)."""
    return eta_min + (base_lr - eta_min) * (1 + math.cos(math.pi * step / t_max)) / 2
