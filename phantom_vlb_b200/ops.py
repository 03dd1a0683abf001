"""Thin Python wrappers over the C ABI (tensor -> raw pointer + shape marshalling only; no arithmetic here)."""
import torch

from . import lib as _l

ACT_NONE, ACT_QUICK_GELU, ACT_GELU, ACT_SILU = 0, 1, 2, 3


def _chk(t, dtype=None):
    assert t.is_cuda and t.stride(-1) == 1, "tensor must be a CUDA tensor with unit inner stride"
    if dtype is not None:
        assert t.dtype == dtype, "expected %s got %s" % (dtype, t.dtype)
    return t


def gemm(a, b, out=None, *, a2=None, b2=None, alpha=1.0, bias=None, residual=None, act=ACT_NONE,
         a_mn=False, b_mn=False, out_f32=False, accumulate=False, block_n=0, max_ctas=0):
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
            int(a_mn), int(b_mn), int(out_f32), int(accumulate), int(block_n), int(max_ctas), _l.cur_stream())
    return out


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
