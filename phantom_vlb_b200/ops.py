"""Thin Python wrappers over the C ABI (tensor -> raw pointer + shape marshalling only; no arithmetic here)."""
import torch

from . import lib as _l

ACT_NONE, ACT_QUICK_GELU, ACT_GELU, ACT_SILU, ACT_SIGMOID = 0, 1, 2, 3, 4


def _chk(t, dtype=None):
    assert t.is_cuda and t.stride(-1) == 1, "tensor must be a CUDA tensor with unit inner stride"
    if dtype is not None:
        assert t.dtype == dtype, "expected %s got %s" % (dtype, t.dtype)
    return t


def gemm(a, b, out=None, *, a2=None, b2=None, alpha=1.0, bias=None, residual=None, act=ACT_NONE,
         a_mn=False, b_mn=False, out_f32=False, accumulate=False, block_n=0, max_ctas=0, split_k=0):
    """out[M,N] = epi(alpha * (a @ b.T + a2 @ b2.T)).

    a: [M,K] (a_mn: stored [K,M]); b: [N,K] (b_mn: stored [K,N]); a2/b2 likewise with their own K2 (LoRA block).
    2-D tensors with unit inner stride; leading dimensions are taken from stride(0).
    """
    _chk(a, torch.bfloat16), _chk(b, torch.bfloat16)
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, "contraction mismatch %d vs %d" % (K, Kb)
    K2 = 0
    if a2 is not None:
        _chk(a2, torch.bfloat16), _chk(b2, torch.bfloat16)
        K2 = a2.shape[0] if a_mn else a2.shape[1]
        assert (b2.shape[0] if b_mn else b2.shape[1]) == K2
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=torch.float32 if out_f32 else torch.bfloat16)
    _chk(out, torch.float32 if out_f32 else torch.bfloat16)
    assert out.shape[0] == M and out.shape[1] == N
    _l.call("vlb_gemm_bf16", _l.ptr(a), a.stride(0), _l.ptr(b), b.stride(0), _l.ptr(out), out.stride(0),
            M, N, K,
            _l.ptr(a2), a2.stride(0) if a2 is not None else 0, _l.ptr(b2), b2.stride(0) if b2 is not None else 0, K2,
            float(alpha), _l.ptr(bias), _l.ptr(residual), residual.stride(0) if residual is not None else 0, int(act),
            int(a_mn), int(b_mn), int(out_f32), int(accumulate), int(block_n), int(max_ctas), int(split_k), _l.cur_stream())
    return out


def gemm_qkv_rope(x, wqkv, qkv, cos, sin, S, n_rot_heads, D, *, a2=None, b2=None, alpha=1.0):
    """qkv[M, N] = x @ wqkv.T (+ a2 @ b2.T), then RoPE on the first n_rot_heads heads (position = row % S), in one launch when
    D == 128 (CTA-pair GEMM with the rotation in its epilogue); bit-identical to gemm() followed by rope_()."""
    _chk(x, torch.bfloat16), _chk(wqkv, torch.bfloat16), _chk(qkv, torch.bfloat16)
    M, K = x.shape
    N = wqkv.shape[0]
    assert wqkv.shape == (N, K) and qkv.shape == (M, N) and cos.dtype == torch.bfloat16 and cos.shape[1] == D // 2
    K2 = 0
    if a2 is not None:
        _chk(a2, torch.bfloat16), _chk(b2, torch.bfloat16)
        K2 = a2.shape[1]
        assert b2.shape == (N, K2) and a2.shape[0] == M
    _l.call("vlb_gemm_qkv_rope", _l.ptr(x), x.stride(0), _l.ptr(wqkv), wqkv.stride(0), _l.ptr(qkv), qkv.stride(0), M, N, K,
            _l.ptr(a2), a2.stride(0) if a2 is not None else 0, _l.ptr(b2), b2.stride(0) if b2 is not None else 0, K2,
            float(alpha), _l.ptr(cos), _l.ptr(sin), int(S), int(n_rot_heads), int(D), _l.cur_stream())
    return qkv


def gemm_gateup_swiglu(x, wgu, gu, act, *, a2=None, b2=None, alpha=1.0):
    """gu[M, 2I] = x @ wgu.T (+ a2 @ b2.T) and act[M, I] = silu(gu[:, :I]) * gu[:, I:] from one CTA-pair GEMM launch."""
    _chk(x, torch.bfloat16), _chk(wgu, torch.bfloat16), _chk(gu, torch.bfloat16), _chk(act, torch.bfloat16)
    M, K = x.shape
    I = wgu.shape[0] // 2
    assert wgu.shape == (2 * I, K) and gu.shape == (M, 2 * I) and act.shape == (M, I) and I % 128 == 0
    K2 = 0
    if a2 is not None:
        _chk(a2, torch.bfloat16), _chk(b2, torch.bfloat16)
        K2 = a2.shape[1]
        assert b2.shape == (2 * I, K2) and a2.shape[0] == M
    _l.call("vlb_gemm_gateup_swiglu", _l.ptr(x), x.stride(0), _l.ptr(wgu), wgu.stride(0), _l.ptr(gu), gu.stride(0),
            _l.ptr(act), act.stride(0), M, I, K,
            _l.ptr(a2), a2.stride(0) if a2 is not None else 0, _l.ptr(b2), b2.stride(0) if b2 is not None else 0, K2,
            float(alpha), _l.cur_stream())
    return gu, act


def gemm_down_dx_swiglu_bwd(dy, wd, gu, dgu, *, a2=None, b2=None, alpha=1.0):
    """dgu[M, 2I] = swiglu_bwd(gu, dy @ wd (+ a2 @ b2)) from one CTA-pair GEMM launch; wd is the forward weight [H, I],
    b2 the adapter's A [r, I]."""
    _chk(dy, torch.bfloat16), _chk(wd, torch.bfloat16), _chk(gu, torch.bfloat16), _chk(dgu, torch.bfloat16)
    M, H = dy.shape
    I = wd.shape[1]
    assert wd.shape[0] == H and gu.shape == (M, 2 * I) and dgu.shape == (M, 2 * I) and I % 256 == 0
    K2 = 0
    if a2 is not None:
        _chk(a2, torch.bfloat16), _chk(b2, torch.bfloat16)
        K2 = a2.shape[1]
        assert b2.shape == (K2, I) and a2.shape[0] == M
    _l.call("vlb_gemm_down_dx_swiglu_bwd", _l.ptr(dy), dy.stride(0), _l.ptr(wd), wd.stride(0), _l.ptr(gu), gu.stride(0),
            _l.ptr(dgu), dgu.stride(0), M, I, H,
            _l.ptr(a2), a2.stride(0) if a2 is not None else 0, _l.ptr(b2), b2.stride(0) if b2 is not None else 0, K2,
            float(alpha), _l.cur_stream())
    return dgu


def attn_fwd(qkv, B, S, Hq, Hkv, D, causal, scale=None, out=None, need_lse=True):
    """Flash attention over a fused projection buffer qkv [B*S, (Hq+2*Hkv)*D] -> (o [B*S, Hq*D], lse [B,Hq,S])."""
    _chk(qkv, torch.bfloat16)
    assert qkv.shape[0] == B * S and qkv.shape[1] == (Hq + 2 * Hkv) * D
    if out is None:
        out = torch.empty((B * S, Hq * D), device=qkv.device, dtype=torch.bfloat16)
    lse = torch.empty((B, Hq, S), device=qkv.device, dtype=torch.float32) if need_lse else None
    if scale is None:
        scale = D ** -0.5
    es = qkv.element_size()
    base = qkv.data_ptr()
    _l.call("vlb_attn_fwd", base, base + Hq * D * es, base + (Hq + Hkv) * D * es, qkv.stride(0), _l.ptr(out),
            out.stride(0), _l.ptr(lse), B, S, Hq, Hkv, D, int(causal), float(scale), _l.cur_stream())
    return out, lse


def attn_bwd(qkv, o, do, lse, B, S, Hq, Hkv, D, causal, scale=None, dqkv=None, dq_f32=None, delta=None):
    """Backward of attn_fwd. Returns (dq_f32 [B*S, Hq*D] fp32, dqkv bf16 with the k/v gradient columns filled)."""
    _chk(qkv, torch.bfloat16), _chk(o, torch.bfloat16), _chk(do, torch.bfloat16)
    dev = qkv.device
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    if dq_f32 is None:
        dq_f32 = torch.empty((B * S, Hq * D), device=dev, dtype=torch.float32)
    if delta is None:
        delta = torch.empty((B, Hq, S), device=dev, dtype=torch.float32)
    if scale is None:
        scale = D ** -0.5
    assert o.stride(0) == do.stride(0)
    base, dbase, es = qkv.data_ptr(), dqkv.data_ptr(), 2
    _l.call("vlb_attn_bwd", base, base + Hq * D * es, base + (Hq + Hkv) * D * es, qkv.stride(0), _l.ptr(o), _l.ptr(do),
            o.stride(0), _l.ptr(lse), _l.ptr(delta), _l.ptr(dq_f32), dbase + Hq * D * es, dbase + (Hq + Hkv) * D * es,
            dqkv.stride(0), B, S, Hq, Hkv, D, int(causal), float(scale), _l.cur_stream())
    return dq_f32, dqkv


# ------------------------------------------------------------------------------------------------ row kernels
def _st():
    return _l.cur_stream()


def rmsnorm_fwd(x, w, eps, out=None, rstd=None):
    M, H = x.shape
    out = torch.empty_like(x) if out is None else out
    _l.call("vlb_rmsnorm_fwd", _l.ptr(x), _l.ptr(w), _l.ptr(out), _l.ptr(rstd), M, H, float(eps), _st())
    return out


def rmsnorm_bwd(dy, x, w, rstd, dres=None, out=None):
    M, H = x.shape
    out = torch.empty_like(x) if out is None else out
    _l.call("vlb_rmsnorm_bwd", _l.ptr(dy), _l.ptr(x), _l.ptr(w), _l.ptr(rstd), _l.ptr(dres), _l.ptr(out), M, H, _st())
    return out


def layernorm_fwd(x, w, b, eps, residual=None, act=ACT_NONE, out=None):
    M, H = x.shape
    out = torch.empty_like(x) if out is None else out
    _l.call("vlb_layernorm_fwd", _l.ptr(x), _l.ptr(w), _l.ptr(b), _l.ptr(residual), _l.ptr(out), M, H, float(eps),
            int(act), _st())
    return out


def rope_(qkv, cos, sin, S, nheads, D, inverse=False):
    _l.call("vlb_rope", _l.ptr(qkv), qkv.stride(0), _l.ptr(cos), _l.ptr(sin), qkv.shape[0], S, nheads, D, int(inverse), _st())
    return qkv


def rope_bwd_f32(dq_f32, dqkv, cos, sin, S, nheads, D):
    _l.call("vlb_rope_bwd_f32", _l.ptr(dq_f32), _l.ptr(dqkv), dqkv.stride(0), _l.ptr(cos), _l.ptr(sin), dqkv.shape[0], S,
            nheads, D, _st())
    return dqkv


def swiglu_fwd(gu, out=None):
    M, I2 = gu.shape
    out = torch.empty((M, I2 // 2), device=gu.device, dtype=gu.dtype) if out is None else out
    _l.call("vlb_swiglu_fwd", _l.ptr(gu), _l.ptr(out), M, I2 // 2, _st())
    return out


def swiglu_bwd(gu, dact, out=None):
    M, I2 = gu.shape
    out = torch.empty_like(gu) if out is None else out
    _l.call("vlb_swiglu_bwd", _l.ptr(gu), _l.ptr(dact), _l.ptr(out), M, I2 // 2, _st())
    return out


# ------------------------------------------------------------------------------------------------ vision / splice
CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)      # transformers OPENAI_CLIP_MEAN / OPENAI_CLIP_STD (CLIPImageProcessor)
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)


def patch_im2col(pixels, patch, kpad, out=None, mean=CLIP_MEAN, std=CLIP_STD):
    """pixels [N,3,H,W]: fp32 CLIP-normalised values (the reference batch contract) or raw uint8 frames, in which case the
    rescale + normalise of CLIPImageProcessor is fused into the gather."""
    N, C, H, W = pixels.shape
    assert C == 3 and H == W and pixels.is_contiguous() and pixels.dtype in (torch.float32, torch.uint8)
    G = H // patch
    out = torch.empty((N * G * G, kpad), device=pixels.device, dtype=torch.bfloat16) if out is None else out
    if pixels.dtype == torch.uint8:
        import ctypes
        m3, s3 = (ctypes.c_float * 3)(*mean), (ctypes.c_float * 3)(*std)
        _l.call("vlb_patch_im2col_u8", _l.ptr(pixels), _l.ptr(out), N, H, patch, kpad, m3, s3, _st())
    else:
        _l.call("vlb_patch_im2col", _l.ptr(pixels), _l.ptr(out), N, H, patch, kpad, _st())
    return out


def clip_embed(patch, cls, pos, lnw, lnb, N, T, eps, out=None):
    H = patch.shape[1]
    out = torch.empty((N * T, H), device=patch.device, dtype=torch.bfloat16) if out is None else out
    _l.call("vlb_clip_embed", _l.ptr(patch), _l.ptr(cls), _l.ptr(pos), _l.ptr(lnw), _l.ptr(lnb), _l.ptr(out), N, T, H,
            float(eps), _st())
    return out


def drop_cls(x, N, T, out=None):
    H = x.shape[1]
    out = torch.empty((N * (T - 1), H), device=x.device, dtype=x.dtype) if out is None else out
    _l.call("vlb_drop_cls", _l.ptr(x), _l.ptr(out), N, T, H, _st())
    return out


def dwconv3x3(x, w9c, F, Hh, Ww, out=None):
    C = x.shape[1]
    out = torch.empty_like(x) if out is None else out
    _l.call("vlb_dwconv3x3", _l.ptr(x), _l.ptr(w9c), _l.ptr(out), F, Hh, Ww, C, _st())
    return out


def se_pool(x, F, HW, out=None):
    C = x.shape[1]
    out = torch.empty((F, C), device=x.device, dtype=x.dtype) if out is None else out
    _l.call("vlb_se_pool", _l.ptr(x), _l.ptr(out), F, HW, C, _st())
    return out


def se_scale(x, gate, F, HW, out=None):
    C = x.shape[1]
    out = torch.empty_like(x) if out is None else out
    _l.call("vlb_se_scale", _l.ptr(x), _l.ptr(gate), _l.ptr(out), F, HW, C, _st())
    return out


def conv3d_gather(x, B, T, Hh, Ww, out=None):
    C = x.shape[1]
    To, Ho, Wo = T // 2 + 1, Hh // 2 + 1, Ww // 2 + 1
    out = torch.empty((B * To * Ho * Wo, 8 * C), device=x.device, dtype=x.dtype) if out is None else out
    _l.call("vlb_conv3d_gather", _l.ptr(x), _l.ptr(out), B, T, Hh, Ww, C, _st())
    return out


def ids_to_long(ids_f32):
    out = torch.empty(ids_f32.shape, device=ids_f32.device, dtype=torch.int64)
    _l.call("vlb_ids_to_long", _l.ptr(ids_f32), _l.ptr(out), ids_f32.numel(), _st())
    return out


def splice_embed(ids, table, vis, Nv, video_id=-201, out=None):
    B, L = ids.shape
    V, H = table.shape
    S = L + Nv - 1
    out = torch.empty((B * S, H), device=table.device, dtype=table.dtype) if out is None else out
    pos = torch.empty((B,), device=table.device, dtype=torch.int32)
    _l.call("vlb_splice_embed", _l.ptr(ids), _l.ptr(table), _l.ptr(vis), _l.ptr(out), _l.ptr(pos), B, L, Nv, H, V,
            int(video_id), _st())
    return out, pos


def weight_mask(padvals, vis_w, lang_w, tok_per_frame, S):
    B, Tp = vis_w.shape
    assert padvals.dtype == torch.int64 and vis_w.dtype == torch.float64 and lang_w.dtype == torch.float64
    out = torch.empty((B, S), device=vis_w.device, dtype=torch.float32)
    _l.call("vlb_weight_mask", _l.ptr(padvals), _l.ptr(vis_w), _l.ptr(lang_w), _l.ptr(out), B, Tp, tok_per_frame,
            lang_w.shape[1], S, _st())
    return out


# ------------------------------------------------------------------------------------------------ optimiser / pearson
def copy2d(src, dst, rows, cols, transpose=False):
    _l.call("vlb_copy2d", _l.ptr(src), src.stride(0), _l.ptr(dst), dst.stride(0), rows, cols, int(transpose), _st())
    return dst


def pearson(preds, targets, stats=None, finalize=True):
    """Per-column Pearson r of [N,P] tensors (bf16 or fp32). Returns (r fp32 [P] or None, stats fp64 [5,P])."""
    N, P = preds.shape
    assert preds.dtype == targets.dtype and preds.is_contiguous() and targets.is_contiguous()
    if stats is None:
        stats = torch.zeros((5, P), device=preds.device, dtype=torch.float64)
    _l.call("vlb_pearson_update", _l.ptr(preds), _l.ptr(targets), _l.ptr(stats), N, P,
            int(preds.dtype == torch.bfloat16), _st())
    r = pearson_finalize(stats, N) if finalize else None
    return r, stats


def pearson_finalize(stats, n_rows):
    P = stats.shape[1]
    r = torch.empty((P,), device=stats.device, dtype=torch.float32)
    _l.call("vlb_pearson_finalize", _l.ptr(stats), int(n_rows), _l.ptr(r), P, _st())
    return r


def paint_labels(r, labels, square=True):
    vol = torch.empty(labels.shape, device=r.device, dtype=torch.float32)
    _l.call("vlb_paint_labels", _l.ptr(r), _l.ptr(labels), _l.ptr(vol), labels.numel(), r.numel(), int(square), _st())
    return vol


def attn_fwd_into(qkv, out, lse, B, S, Hq, Hkv, D, causal, scale=None):
    """attn_fwd writing into caller-owned o / lse buffers (engine path: no allocation per step)."""
    if scale is None:
        scale = D ** -0.5
    base = qkv.data_ptr()
    _l.call("vlb_attn_fwd", base, base + Hq * D * 2, base + (Hq + Hkv) * D * 2, qkv.stride(0), _l.ptr(out),
            out.stride(0), _l.ptr(lse), B, S, Hq, Hkv, D, int(causal), float(scale), _l.cur_stream())
    return out


def dropout(x, p, seed, out=None, mask=None):
    """out = x * keep / (1 - p) with a counter-based mask of (seed, element index); mask (uint8 [M, K]) is a test hook."""
    M, K = x.shape
    out = torch.empty((M, K), device=x.device, dtype=x.dtype) if out is None else out
    _l.call("vlb_dropout", _l.ptr(x), x.stride(0), _l.ptr(out), out.stride(0), _l.ptr(mask), M, K, float(p), int(seed), _st())
    return out


def dropout_add(t, y, p, seed):
    """y += t * keep / (1 - p) with the same mask dropout(x, p, seed) used."""
    M, K = t.shape
    _l.call("vlb_dropout_add", _l.ptr(t), t.stride(0), _l.ptr(y), y.stride(0), M, K, float(p), int(seed), _st())
    return y


# ------------------------------------------------------------------------------------------------ LoRA dropout on the fused path
def dropout_bits(n_tok, n_feat, p, seed, out):
    """Keep mask of one adapter over its input [n_tok, n_feat] as packed bits: out int64 [n_feat // 64, n_tok]."""
    assert out.dtype == torch.int64 and out.shape == (n_feat // 64, n_tok) and out.is_contiguous()
    _l.call("vlb_dropout_bits", _l.ptr(out), int(n_tok), int(n_feat), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, _st())
    return out


def unpack_bits(bits):
    """int64 [n_feat // 64, n_tok] -> uint8 keep mask [n_tok, n_feat] (tests: replay inside the oracle)."""
    nfb, n_tok = bits.shape
    sh = torch.arange(64, device=bits.device, dtype=torch.int64)
    m = (bits[:, :, None] >> sh[None, None, :]) & 1                 # [fb, tok, 64]
    return m.permute(1, 0, 2).reshape(n_tok, nfb * 64).to(torch.uint8)


def _bits3(bits):
    b = [_l.ptr(t) for t in bits] + [None, None]
    return b[0], b[1], b[2]


def lora_down_masked(x, A, xa, bits, alpha):
    """xa[M, 16 n] = alpha * (keep_a . x) A_a^T for the n adapters sharing x; A [16 n, K] stacked, bits = list of n bit arrays."""
    _chk(x, torch.bfloat16), _chk(A, torch.bfloat16), _chk(xa, torch.bfloat16)
    M, K = x.shape
    n = len(bits)
    assert A.shape == (16 * n, K) and xa.shape == (M, 16 * n)
    _l.call("vlb_lora_down_masked", _l.ptr(x), x.stride(0), _l.ptr(A), A.stride(0), _l.ptr(xa), xa.stride(0), M, K, n,
            *_bits3(bits), float(alpha), _st())
    return xa


def lora_grad_a_masked(x, t, dA, bits, alpha):
    """dA[16 n, K] += alpha * t_a^T (keep_a . x); x [M, K], t [M, 16 n], dA fp32."""
    _chk(x, torch.bfloat16), _chk(t, torch.bfloat16), _chk(dA, torch.float32)
    M, K = x.shape
    n = len(bits)
    assert t.shape == (M, 16 * n) and dA.shape == (16 * n, K)
    _l.call("vlb_lora_grad_a_masked", _l.ptr(x), x.stride(0), _l.ptr(t), t.stride(0), _l.ptr(dA), dA.stride(0), M, K, n,
            *_bits3(bits), float(alpha), _st())
    return dA


def gemm_dx_lora_masked(dy, W, out, t, A, bits, scale):
    """out[M, N] = dy[M, K] @ W[K, N] + scale * sum_a keep_a . (t_a @ A_a)  (W = forward weight [out = K, in = N])."""
    _chk(dy, torch.bfloat16), _chk(W, torch.bfloat16), _chk(out, torch.bfloat16), _chk(t, torch.bfloat16), _chk(A, torch.bfloat16)
    M, K = dy.shape
    N = W.shape[1]
    n = len(bits)
    assert W.shape[0] == K and out.shape == (M, N) and t.shape == (M, 16 * n) and A.shape == (16 * n, N)
    _l.call("vlb_gemm_dx_lora_masked", _l.ptr(dy), dy.stride(0), _l.ptr(W), W.stride(0), _l.ptr(out), out.stride(0), M, N, K,
            _l.ptr(t), t.stride(0), _l.ptr(A), A.stride(0), n, *_bits3(bits), float(scale), _st())
    return out


def gemm_down_dx_swiglu_bwd_masked(dy, wd, gu, dgu, t, A, bits, scale):
    """dgu = swiglu_bwd(gu, dy @ wd + scale * keep . (t @ A)) from one CTA-pair GEMM launch (wd [H, I], A [16, I])."""
    _chk(dy, torch.bfloat16), _chk(wd, torch.bfloat16), _chk(gu, torch.bfloat16), _chk(dgu, torch.bfloat16)
    M, H = dy.shape
    I = wd.shape[1]
    assert wd.shape[0] == H and gu.shape == (M, 2 * I) and dgu.shape == (M, 2 * I) and t.shape == (M, 16) and A.shape == (16, I)
    _l.call("vlb_gemm_down_dx_swiglu_bwd_masked", _l.ptr(dy), dy.stride(0), _l.ptr(wd), wd.stride(0), _l.ptr(gu), gu.stride(0),
            _l.ptr(dgu), dgu.stride(0), M, I, H, _l.ptr(t), t.stride(0), _l.ptr(A), A.stride(0), _l.ptr(bits), float(scale), _st())
    return dgu
