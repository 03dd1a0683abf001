"""tcgen05 GEMM vs torch fp32 matmul on bf16-rounded inputs (tolerance: bf16 output rounding, 2^-8 rel)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a2, b2, alpha, bias, residual, act, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    y = A @ B.t()
    if a2 is not None:
        A2 = a2.float().t() if a_mn else a2.float()
        B2 = b2.float().t() if b_mn else b2.float()
        y = y + A2 @ B2.t()
    y = y * alpha
    if bias is not None:
        y = y + bias.float()
    if act == 1:
        y = y * torch.sigmoid(1.702 * y)
    elif act == 2:
        y = torch.nn.functional.gelu(y)
    elif act == 3:
        y = torch.nn.functional.silu(y)
    if residual is not None:
        y = y.bfloat16().float() + residual.float()
    return y


def _close(out, ref, tol=1e-2):
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err / scale < tol, "max err %g scale %g" % (err, scale)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K,bn", [(256, 256, 128, 0), (300, 200, 136, 0), (1024, 512, 4096, 256),
                                      (128, 64, 64, 64), (77, 48, 1000, 0), (2048, 1024, 512, 128)])
def test_gemm_majors(M, N, K, bn, a_mn, b_mn):
    from phantom_vlb_b200 import ops
    if a_mn and M % 8:
        M = (M + 7) // 8 * 8
    torch.manual_seed(M * 7 + N * 3 + K)
    dev = "cuda"
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    out = ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, block_n=bn)
    torch.cuda.synchronize()
    _close(out, _ref(a, b, None, None, 1.0, None, None, 0, a_mn, b_mn))


@pytest.mark.parametrize("act", [0, 1, 2, 3])
def test_gemm_epilogue_lora(act):
    from phantom_vlb_b200 import ops
    torch.manual_seed(act)
    M, N, K, R = 640, 384, 1024, 16
    dev = "cuda"
    a = torch.randn(M, K, device=dev).bfloat16()
    b = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    a2 = torch.randn(M, R, device=dev).bfloat16()
    b2 = torch.randn(N, R, device=dev).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    res = torch.randn(M, N, device=dev).bfloat16()
    out = ops.gemm(a, b, a2=a2, b2=b2, alpha=0.5, bias=bias, residual=res, act=act)
    torch.cuda.synchronize()
    _close(out, _ref(a, b, a2, b2, 0.5, bias, res, act, False, False))


def test_gemm_f32_out_accumulate():
    from phantom_vlb_b200 import ops
    torch.manual_seed(0)
    M, N, K = 48, 4096, 2048          # dA-like: contraction over the token axis, both operands MN-major
    a = torch.randn(K, M, device="cuda").bfloat16()
    b = torch.randn(K, N, device="cuda").bfloat16()
    out = torch.ones(M, N, device="cuda", dtype=torch.float32)
    ops.gemm(a, b, out, a_mn=True, b_mn=True, out_f32=True, accumulate=True, alpha=2.0)
    torch.cuda.synchronize()
    ref = 1.0 + 2.0 * (a.float().t() @ b.float())
    _close(out, ref, 1e-4)


def test_gemm_persistent_many_tiles():
    from phantom_vlb_b200 import ops
    torch.manual_seed(1)
    M, N, K = 4096, 6144, 512
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    out = ops.gemm(a, b)
    torch.cuda.synchronize()
    _close(out, a.float() @ b.float().t())


@pytest.mark.parametrize("M,N,K,split", [(16, 4096, 8192, 0), (48, 4096, 8192, 0), (16, 14336, 4096 + 64, 7), (32, 1000, 3 * 64, 0),
                                         (16, 4096, 8192, 1)])
def test_gemm_split_k_red_add(M, N, K, split):
    """LoRA-gradient shape: tiny [r, N] fp32 output, contraction over all tokens split across CTAs with red.add."""
    from phantom_vlb_b200 import ops
    torch.manual_seed(2)
    a = (torch.randn(K, M, device="cuda") * 0.1).bfloat16()
    b = (torch.randn(K, N, device="cuda") * 0.1).bfloat16()
    base = torch.randn(M, N, device="cuda", dtype=torch.float32)
    out = base.clone()
    ops.gemm(a, b, out, a_mn=True, b_mn=True, out_f32=True, accumulate=True, alpha=0.5, split_k=split)
    torch.cuda.synchronize()
    ref = base + 0.5 * (a.float().t() @ b.float())
    _close(out, ref, 1e-4)


def test_gemm_split_k_needs_accumulate():
    from phantom_vlb_b200 import ops
    a = torch.zeros(256, 16, device="cuda").bfloat16()
    b = torch.zeros(256, 128, device="cuda").bfloat16()
    with pytest.raises(RuntimeError):
        ops.gemm(a, b, a_mn=True, b_mn=True, out_f32=True, accumulate=False, split_k=4)


@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("M,N,K,R", [(4096, 4096, 512, 0), (4096 + 200, 5120, 320, 48), (16384, 2304, 192, 16), (2560, 8192, 1088, 32),
                                     (2560 + 72, 4096, 8192 + 136, 16)])      # K >= 8192: serpentine tile / K order, three tiles per CTA
def test_gemm_cta_pair_kernel(M, N, K, R, b_mn):
    """Shapes with >= 74 tiles of 256 x 256 take the cta_group::2 kernel (gemm2_kernel.cuh): ragged M (the peer CTA's rows
    partly / fully out of range), fused LoRA k-block (K2 < 64: zero-filled by TMA), bias + activation + residual epilogue."""
    from phantom_vlb_b200 import ops
    torch.manual_seed(M + N + K + R)
    dev = "cuda"
    a = torch.randn(M, K, device=dev).bfloat16()
    b = (torch.randn((K, N) if b_mn else (N, K), device=dev) * 0.05).bfloat16()
    a2 = b2 = None
    if R:
        a2 = torch.randn(M, R, device=dev).bfloat16()
        b2 = (torch.randn((R, N) if b_mn else (N, R), device=dev) * 0.1).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    res = torch.randn(M, N, device=dev).bfloat16()
    out = ops.gemm(a, b, a2=a2, b2=b2, alpha=0.5, bias=bias, residual=res, act=3, b_mn=b_mn)
    torch.cuda.synchronize()
    _close(out, _ref(a, b, a2, b2, 0.5, bias, res, 3, False, b_mn))
    # and the plain product, repeated to cover the accumulator double-buffer / ring wrap-around across many tiles
    out2 = ops.gemm(a, b, b_mn=b_mn)
    torch.cuda.synchronize()
    _close(out2, _ref(a, b, None, None, 1.0, None, None, 0, False, b_mn))


@pytest.mark.parametrize("M,I,K,R", [(256, 128, 64, 0), (256, 512, 128, 16), (4096 + 72, 1792, 1024, 48), (16384, 14336 // 4, 512, 48)])
def test_gemm_gateup_swiglu_fused(M, I, K, R):
    from phantom_vlb_b200 import ops
    """The pair kernel with SwiGLU in its epilogue (leader stages gate rows, peer the matching up rows) == the plain GEMM
    followed by the row kernel, bit for bit: gu in the standard [M, 2I] layout and act [M, I]."""
    g = torch.Generator(device="cuda").manual_seed(M + I)
    x = (torch.randn((M, K), device="cuda", generator=g)).bfloat16()
    w = (torch.randn((2 * I, K), device="cuda", generator=g) * K ** -0.5).bfloat16()
    a2 = (torch.randn((M, R), device="cuda", generator=g)).bfloat16() if R else None
    b2 = (torch.randn((2 * I, R), device="cuda", generator=g) * 0.1).bfloat16() if R else None
    gu_ref = ops.gemm(x, w, a2=a2, b2=b2)
    act_ref = ops.swiglu_fwd(gu_ref)
    gu = torch.full((M, 2 * I), float("nan"), device="cuda", dtype=torch.bfloat16)
    act = torch.full((M, I), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_gateup_swiglu(x, w, gu, act, a2=a2, b2=b2)
    torch.cuda.synchronize()
    ref32 = x.float() @ w.float().t() + (a2.float() @ b2.float().t() if R else 0)
    _close(gu, ref32)
    assert torch.equal(gu, gu_ref), "gu differs from the unfused GEMM: max %g" % float((gu.float() - gu_ref.float()).abs().max())
    assert torch.equal(act, act_ref), "act differs: max %g" % float((act.float() - act_ref.float()).abs().max())


@pytest.mark.parametrize("M,I,H,R", [(256, 256, 64, 0), (256, 512, 128, 16), (4096 + 72, 1792, 1024, 16), (16384, 14336 // 4, 512, 16)])
def test_gemm_down_dx_swiglu_bwd_fused(M, I, H, R):
    """d(act) = dy Wd (+ t A) with the SwiGLU backward in the pair kernel's epilogue == GEMM then the row kernel, bit for bit."""
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + I + 1)
    dy = (torch.randn((M, H), device="cuda", generator=g)).bfloat16()
    wd = (torch.randn((H, I), device="cuda", generator=g) * H ** -0.5).bfloat16()
    gu = (torch.randn((M, 2 * I), device="cuda", generator=g) * 2).bfloat16()
    t = (torch.randn((M, R), device="cuda", generator=g)).bfloat16() if R else None
    a_l = (torch.randn((R, I), device="cuda", generator=g) * 0.1).bfloat16() if R else None
    dact = ops.gemm(dy, wd, a2=t, b2=a_l, b_mn=True)
    ref = ops.swiglu_bwd(gu, dact)
    dgu = torch.full((M, 2 * I), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_down_dx_swiglu_bwd(dy, wd, gu, dgu, a2=t, b2=a_l)
    torch.cuda.synchronize()
    assert torch.equal(dgu, ref), "dgu differs: max %g" % float((dgu.float() - ref.float()).abs().max())
    # and against fp32 torch on the same bf16 operands (bf16 rounding of d(act) and of the result)
    d32 = dy.float() @ wd.float() + (t.float() @ a_l.float() if R else 0)
    gf, uf = gu[:, :I].float(), gu[:, I:].float()
    sg = torch.sigmoid(gf)
    ref32 = torch.cat([d32 * uf * sg * (1 + gf * (1 - sg)), d32 * gf * sg], dim=1)
    _close(dgu, ref32, tol=2e-2)


@pytest.mark.parametrize("M,Hq,Hkv,D,K,R,S", [(16384, 32, 8, 128, 4096, 48, 2048), (4096 + 72, 8, 2, 128, 512, 0, 521),
                                              (256, 4, 2, 32, 128, 48, 64), (300, 2, 2, 128, 256, 16, 100)])
def test_gemm_qkv_rope_fused(M, Hq, Hkv, D, K, R, S):
    """q/k/v projection (+ LoRA k-block) with RoPE in the pair kernel's epilogue == GEMM then the rope kernel, bit for bit (the 7B
    shape and a ragged one run fused; head_dim 32 and a grid too small for the pair kernel take the two-kernel route inside the
    entry point), and against fp32 torch with HF's rotate_half formula."""
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + K + D)
    N = (Hq + 2 * Hkv) * D
    x = torch.randn((M, K), device="cuda", generator=g).bfloat16()
    w = (torch.randn((N, K), device="cuda", generator=g) * K ** -0.5).bfloat16()
    a2 = torch.randn((M, R), device="cuda", generator=g).bfloat16() if R else None
    b2 = (torch.randn((N, R), device="cuda", generator=g) * 0.1).bfloat16() if R else None
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device="cuda").float() / D))
    ang = torch.arange(S, device="cuda").float()[:, None] * inv[None, :]
    cos, sin = ang.cos().bfloat16(), ang.sin().bfloat16()
    ref = ops.gemm(x, w, a2=a2, b2=b2)
    plain = ref.clone()
    ops.rope_(ref, cos, sin, S, Hq + Hkv, D)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_qkv_rope(x, w, out, cos, sin, S, Hq + Hkv, D, a2=a2, b2=b2)
    torch.cuda.synchronize()
    assert torch.equal(out, ref), "fused qkv+rope differs from gemm -> rope: max %g" % float((out.float() - ref.float()).abs().max())
    assert torch.equal(out[:, (Hq + Hkv) * D:], plain[:, (Hq + Hkv) * D:])          # v passes through
    p32 = (x.float() @ w.float().t() + (a2.float() @ b2.float().t() if R else 0)).view(M, Hq + 2 * Hkv, D)
    pos = torch.arange(M, device="cuda") % S
    c = torch.cat([cos, cos], 1).float()[pos][:, None, :]
    s = torch.cat([sin, sin], 1).float()[pos][:, None, :]
    rh = torch.cat([-p32[..., D // 2:], p32[..., :D // 2]], -1)
    r32 = p32.clone()
    r32[:, :Hq + Hkv] = (p32 * c + rh * s)[:, :Hq + Hkv]
    _close(out, r32.view(M, N), tol=2e-2)
