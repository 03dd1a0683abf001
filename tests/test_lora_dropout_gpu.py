"""LoRA input dropout on the fused path (csrc/lora_dropout.cu, LORA_EPI in csrc/gemm2_kernel.cuh) against fp32 torch on the
same bf16 operands and the SAME masks (unpacked from the kernels' bit arrays). peft semantics: y = W x + s B(A(dropout(x))),
one independent mask per adapter, exact p (reference lora_dropout = 0.1, VLB_vllama2_friends_lora.yaml:30)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float(), b.detach().float()
    return float((a - b).norm() / (b.norm() + 1e-12))


def _bits(n_ad, M, K, p=0.1, seed=11):
    from phantom_vlb_b200 import ops
    bits = torch.empty((n_ad, K // 64, M), device="cuda", dtype=torch.int64)
    for a in range(n_ad):
        ops.dropout_bits(M, K, p, seed + a, bits[a])
    masks = [ops.unpack_bits(bits[a]).float() for a in range(n_ad)]
    return [bits[a] for a in range(n_ad)], masks


def test_dropout_bits_exact_probability_and_independence():
    from phantom_vlb_b200 import ops
    M, K = 4096, 4096
    for p in (0.1, 0.05, 0.5):
        (b0, b1), (m0, m1) = _bits(2, M, K, p=p)
        for m in (m0, m1):
            frac = float(m.mean())
            assert abs(frac - (1 - p)) < 6e-4, (p, frac)              # 16.7 M draws: sigma <= 1.2e-4
        agree = float((m0 == m1).float().mean())
        assert abs(agree - ((1 - p) ** 2 + p ** 2)) < 1e-3            # independent masks per adapter
        assert abs(float(m0.mean(0).std()) - (p * (1 - p) / M) ** 0.5) < 2e-3 and abs(float(m0.mean(1).std()) - (p * (1 - p) / K) ** 0.5) < 2e-3
    again = torch.empty_like(b0)
    ops.dropout_bits(M, K, 0.5, 11, again)
    assert torch.equal(again, b0)                                      # a function of (seed, index) only


@pytest.mark.parametrize("M,K,n_ad", [(16384, 4096, 3), (16384, 14336, 1), (1000, 4096, 2), (256, 256, 3), (300, 512, 1)])
def test_lora_down_masked_matches_torch(M, K, n_ad):
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    A = (torch.randn(16 * n_ad, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
    bits, masks = _bits(n_ad, M, K)
    xa = torch.full((M, 16 * n_ad), float("nan"), device="cuda", dtype=torch.bfloat16)
    alpha = 2.0 / 0.9
    ops.lora_down_masked(x, A, xa, bits, alpha)
    ref = torch.cat([alpha * (x.float() * masks[a]) @ A[16 * a:16 * a + 16].float().t() for a in range(n_ad)], 1)
    assert torch.isfinite(xa.float()).all()
    assert rel(xa, ref) < 4e-3, rel(xa, ref)


@pytest.mark.parametrize("M,K,n_ad", [(16384, 4096, 3), (16384, 14336, 1), (1000, 4096, 2), (256, 256, 3), (264, 512, 1)])
def test_lora_grad_a_masked_matches_torch(M, K, n_ad):
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    t = torch.randn(M, 16 * n_ad, device="cuda", generator=g).bfloat16()
    bits, masks = _bits(n_ad, M, K)
    base = torch.randn(16 * n_ad, K, device="cuda", generator=g)
    dA = base.clone()
    ops.lora_grad_a_masked(x, t, dA, bits, 1 / 0.9)
    ref = torch.cat([(1 / 0.9) * t[:, 16 * a:16 * a + 16].float().t() @ (x.float() * masks[a]) for a in range(n_ad)], 0)
    assert rel(dA - base, ref) < 2e-3, rel(dA - base, ref)


@pytest.mark.parametrize("M,N,K,n_ad", [(16384, 4096, 6144, 3), (16384, 4096, 4096, 1), (2048, 4096, 28672, 2), (256, 256, 512, 3), (300, 512, 256, 2)])
def test_gemm_dx_lora_masked_matches_torch(M, N, K, n_ad):
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(2)
    dy = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    W = (torch.randn(K, N, device="cuda", generator=g) * K ** -0.5).bfloat16()
    t = torch.randn(M, 16 * n_ad, device="cuda", generator=g).bfloat16()
    A = (torch.randn(16 * n_ad, N, device="cuda", generator=g) * 0.25).bfloat16()
    bits, masks = _bits(n_ad, M, N)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_dx_lora_masked(dy, W, out, t, A, bits, 1 / 0.9)
    ref = dy.float() @ W.float()
    for a in range(n_ad):
        ref += (1 / 0.9) * masks[a] * (t[:, 16 * a:16 * a + 16].float() @ A[16 * a:16 * a + 16].float())
    assert torch.isfinite(out.float()).all()
    assert rel(out, ref) < 4e-3, rel(out, ref)
    # the adapter term really is there and really is masked
    assert rel(out, dy.float() @ W.float()) > 0.1


@pytest.mark.parametrize("M,I,H", [(16384, 14336, 4096), (256, 512, 256), (300, 256, 512)])
def test_gemm_down_dx_swiglu_bwd_masked_matches_unfused(M, I, H):
    from phantom_vlb_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    dy = torch.randn(M, H, device="cuda", generator=g).bfloat16()
    wd = (torch.randn(H, I, device="cuda", generator=g) * H ** -0.5).bfloat16()
    gu = torch.randn(M, 2 * I, device="cuda", generator=g).bfloat16()
    t = torch.randn(M, 16, device="cuda", generator=g).bfloat16()
    A = (torch.randn(16, I, device="cuda", generator=g) * 0.25).bfloat16()
    bits, masks = _bits(1, M, I)
    dgu = torch.empty_like(gu)
    ops.gemm_down_dx_swiglu_bwd_masked(dy, wd, gu, dgu, t, A, bits[0], 1 / 0.9)
    dact = (dy.float() @ wd.float() + (1 / 0.9) * masks[0] * (t.float() @ A.float())).bfloat16().float()
    gt, up = gu[:, :I].float(), gu[:, I:].float()
    sg = torch.sigmoid(gt)
    ref = torch.cat([dact * up * (sg * (1 + gt * (1 - sg))), dact * (gt * sg)], 1)
    assert rel(dgu, ref) < 6e-3, rel(dgu, ref)
