// Row-wise normalisation, RoPE and SwiGLU kernels (HBM-bound; 16-byte vector accesses, warp-shuffle reductions).
//
// Rounding points follow the HF modules the reference reaches under bf16 weights (the oracle is the fp32 restatement):
//   MistralRMSNorm : fp32 variance, x*rsqrt -> bf16, then * weight (bf16)     [transformers modeling_mistral.py]
//   nn.LayerNorm   : fp32 statistics, affine in fp32, one rounding to bf16
//   apply_rotary   : q*cos + rotate_half(q)*sin with bf16 cos/sin tables
//   MistralMLP     : act_fn(gate) * up, act rounded to bf16 before the product
#include "common.cuh"

namespace vlb {

// --------------------------------------------------------------------------------------------- RMSNorm
// one CTA per row; H % 8 == 0; each thread walks the row in 8-element (16 B) chunks kept in registers (H <= 8*8*threads)
template <int THREADS, int MAXV>
__global__ void __launch_bounds__(THREADS) rmsnorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                              bf16* __restrict__ y, float* __restrict__ rstd_out,
                                                              int H, float eps) {
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const bf16* xr = x + row * H;
    float v[MAXV][8];
    float ss = 0.f;
    const int nchunk = H / 8;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            load8(xr + c * 8, v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
        }
    }
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / (float)H + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
    bf16* yr = y + row * H;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            float wv[8], o[8];
            load8(w + c * 8, wv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = wv[j] * bf2f(f2bf(v[i][j] * rstd));
            store8(yr + c * 8, o);
        }
    }
}

// dx = dres + rstd * (g - xhat * mean(g * xhat)),  g = dy * w,  xhat = x * rstd      (w frozen: no dw)
template <int THREADS, int MAXV>
__global__ void __launch_bounds__(THREADS, 768 / THREADS) rmsnorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                                              const bf16* __restrict__ w, const float* __restrict__ rstd_in,
                                                              const bf16* __restrict__ dres, bf16* __restrict__ dx, int H) {
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const float rstd = rstd_in[row];
    // the row stays in registers as packed bf16 (48 instead of 96 registers at H=4096: more rows in flight per SM); the
    // second pass redoes the same fp32 products, so the result is bit-identical to holding them in fp32
    uint4 gq[MAXV], xq[MAXV], rq[MAXV];
    float dot = 0.f;
    const int nchunk = H / 8;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            gq[i] = *reinterpret_cast<const uint4*>(dy + row * H + c * 8);
            xq[i] = *reinterpret_cast<const uint4*>(x + row * H + c * 8);
            if (dres) rq[i] = *reinterpret_cast<const uint4*>(dres + row * H + c * 8);   // issued before the row reduction
        }
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            float wv[8], g[8], xh[8];
            load8(w + c * 8, wv);
            unpack8(gq[i], g);
            unpack8(xq[i], xh);
#pragma unroll
            for (int j = 0; j < 8; ++j) { g[j] *= wv[j]; xh[j] *= rstd; dot += g[j] * xh[j]; }
        }
    }
    dot = block_sum(dot, red) / (float)H;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            float wv[8], g[8], xh[8], o[8];
            load8(w + c * 8, wv);
            unpack8(gq[i], g);
            unpack8(xq[i], xh);
#pragma unroll
            for (int j = 0; j < 8; ++j) { g[j] *= wv[j]; xh[j] *= rstd; o[j] = rstd * (g[j] - xh[j] * dot); }
            if (dres) {
                float rs[8];
                unpack8(rq[i], rs);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] += rs[j];
            }
            store8(dx + row * H + c * 8, o);
        }
    }
}

// --------------------------------------------------------------------------------------------- LayerNorm (+residual, +act)
// y = act( LN(x) * w + b  (+ residual) ).  Used for CLIP LN (act none) and timm LayerNorm2d on NHWC rows (+SiLU).
template <int THREADS, int MAXV>
__global__ void __launch_bounds__(THREADS) layernorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                const bf16* __restrict__ b, const bf16* __restrict__ residual,
                                                                bf16* __restrict__ y, int H, float eps, int act) {
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    float v[MAXV][8];
    float s = 0.f;
    const int nchunk = H / 8;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            load8(x + row * H + c * 8, v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) s += v[i][j];
        }
    }
    const float mean = block_sum(s, red) / (float)H;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
        }
    }
    const float rstd = rsqrtf(block_sum(ss, red) / (float)H + eps);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nchunk) {
            float wv[8], bv[8], o[8];
            load8(w + c * 8, wv);
            load8(b + c * 8, bv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * wv[j] + bv[j];
            if (residual) {
                float r[8]; load8(residual + row * H + c * 8, r);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = bf2f(f2bf(o[j])) + r[j];
            }
            if (act != ACT_NONE) apply_act8(o, act);
            store8(y + row * H + c * 8, o);
        }
    }
}

// Narrow rows (H <= 1024: the CLIP tower's 1024-wide LayerNorms, 47 calls per step over 36 928 rows): one WARP per row, the two
// reductions by shuffles only. With a 128-thread CTA per 2 KB row and four __syncthreads per row the kernel ran at 0.3 of the HBM
// rate (76 us for 151 MB): too few bytes in flight per SM and a barrier chain per row.
template <int CPL>
__global__ void __launch_bounds__(256) layernorm_fwd_warp_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                 const bf16* __restrict__ b, const bf16* __restrict__ residual,
                                                                 bf16* __restrict__ y, long long M, int H, float eps, int act) {
    const int lane = threadIdx.x & 31;
    const long long row = blockIdx.x * 8LL + (threadIdx.x >> 5);
    if (row >= M) return;
    const int nchunk = H / 8;
    float v[CPL][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        const int c = lane + i * 32;
        if (c < nchunk) {
            load8(x + row * H + c * 8, v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) s += v[i][j];
        }
    }
    const float mean = warp_sum(s) / (float)H;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        if (lane + i * 32 < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
        }
    }
    const float rstd = rsqrtf(warp_sum(ss) / (float)H + eps);
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        const int c = lane + i * 32;
        if (c < nchunk) {
            float wv[8], bv[8], o[8];
            load8(w + c * 8, wv);
            load8(b + c * 8, bv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * wv[j] + bv[j];
            if (residual) {
                float r[8]; load8(residual + row * H + c * 8, r);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = bf2f(f2bf(o[j])) + r[j];
            }
            if (act != ACT_NONE) apply_act8(o, act);
            store8(y + row * H + c * 8, o);
        }
    }
}

// --------------------------------------------------------------------------------------------- RoPE (in place on fused qkv rows)
// qkv row layout: [Hq*D | Hkv*D | Hkv*D]; rotates the first (Hq+Hkv) heads. cos/sin: bf16 tables [S, D/2].
// inverse=1 applies the transpose rotation (backward). One thread per (row, head, 8-element chunk of the first half).
__global__ void rope_kernel(bf16* __restrict__ qkv, int ld, const bf16* __restrict__ cosT, const bf16* __restrict__ sinT,
                            long long M, int S, int nheads, int D, int inverse) {
    const int half = D / 2, cph = half / 8;
    const long long total = M * nheads * cph;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cph);
        const int h = (int)((i / cph) % nheads);
        const long long row = i / ((long long)cph * nheads);
        const int pos = (int)(row % S);
        bf16* p = qkv + row * ld + h * D + c * 8;
        float a[8], b[8], cs[8], sn[8], oa[8], ob[8];
        load8(p, a); load8(p + half, b);
        load8(cosT + (size_t)pos * half + c * 8, cs);
        load8(sinT + (size_t)pos * half + c * 8, sn);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float s = inverse ? -sn[j] : sn[j];
            // bf16 products, bf16 sum (matches q*cos + rotate_half(q)*sin evaluated in bf16)
            oa[j] = bf2f(f2bf(a[j] * cs[j])) + bf2f(f2bf(-b[j] * s));
            ob[j] = bf2f(f2bf(b[j] * cs[j])) + bf2f(f2bf(a[j] * s));
        }
        store8(p, oa); store8(p + half, ob);
    }
}

// Backward variant fused with the fp32 -> bf16 conversion of attention's dq accumulator:
// reads dq fp32 [M, Hq*D], applies the inverse rotation, writes bf16 into the dqkv row (q part).
__global__ void rope_bwd_f32_kernel(const float* __restrict__ dq, bf16* __restrict__ dqkv, int ld,
                                    const bf16* __restrict__ cosT, const bf16* __restrict__ sinT,
                                    long long M, int S, int nheads, int D) {
    const int half = D / 2, cph = half / 8;
    const long long total = M * nheads * cph;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cph);
        const int h = (int)((i / cph) % nheads);
        const long long row = i / ((long long)cph * nheads);
        const int pos = (int)(row % S);
        const float* src = dq + row * (long long)nheads * D + h * D + c * 8;
        float a[8], b[8], cs[8], sn[8], oa[8], ob[8];
        load8(src, a); load8(src + half, b);
        load8(cosT + (size_t)pos * half + c * 8, cs);
        load8(sinT + (size_t)pos * half + c * 8, sn);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            oa[j] = a[j] * cs[j] + b[j] * sn[j];
            ob[j] = b[j] * cs[j] - a[j] * sn[j];
        }
        bf16* p = dqkv + row * ld + h * D + c * 8;
        store8(p, oa); store8(p + half, ob);
    }
}

// --------------------------------------------------------------------------------------------- SwiGLU
// gu: [M, 2I] = [gate | up];  act[M, I] = bf16(silu(gate)) * up
__global__ void swiglu_fwd_kernel(const bf16* __restrict__ gu, bf16* __restrict__ act, long long M, int I) {
    const int cpr = I / 8;
    const long long total = M * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / cpr; const int c = (int)(i % cpr);
        float g[8], u[8], o[8];
        load8(gu + row * 2 * I + c * 8, g);
        load8(gu + row * 2 * I + I + c * 8, u);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = bf2f(f2bf(g[j] / (1.f + __expf(-g[j])))) * u[j];
        store8(act + row * I + c * 8, o);
    }
}
// dgu[M, 2I] = [dact * up * silu'(gate) | dact * silu(gate)]
__global__ void swiglu_bwd_kernel(const bf16* __restrict__ gu, const bf16* __restrict__ dact, bf16* __restrict__ dgu,
                                  long long M, int I) {
    const int cpr = I / 8;
    const long long total = M * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / cpr; const int c = (int)(i % cpr);
        float g[8], u[8], d[8], dg[8], du[8];
        load8(gu + row * 2 * I + c * 8, g);
        load8(gu + row * 2 * I + I + c * 8, u);
        load8(dact + row * I + c * 8, d);
        swiglu_bwd8(g, u, d, dg, du);
        store8(dgu + row * 2 * I + c * 8, dg);
        store8(dgu + row * 2 * I + I + c * 8, du);
    }
}

// --------------------------------------------------------------------------------------------- LoRA input dropout
// peft lora.Linear computes B(A(dropout(x))) (reference lora_dropout = 0.1, VLB_vllama2_friends_lora.yaml:30). The keep
// mask is a counter-based hash of (seed, row * K + col): it is regenerated, never stored, so the backward pass applies the
// identical mask to the adapter's input gradient. One 64-bit mix per 8 consecutive elements (one byte each).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// keep bit of element j (0..7) of 8-chunk `chunk`: byte j of the hash compared with round(p * 256)
__device__ __forceinline__ unsigned long long drop_bits(unsigned long long seed, long long chunk) {
    return mix64(seed + (unsigned long long)chunk * 0x9E3779B97F4A7C15ull);
}
// out = x * keep / (1 - p)   (mask_out, if given, receives 0/1 bytes: test hook so an oracle can replay the mask)
__global__ void dropout_kernel(const bf16* __restrict__ x, int ldx, bf16* __restrict__ out, int ldo, uint8_t* __restrict__ mask_out,
                               long long M, int K, unsigned thr, float inv_keep, unsigned long long seed) {
    const int cpr = K / 8;
    const long long total = M * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / cpr; const int c = (int)(i % cpr);
        float v[8];
        load8(x + row * ldx + c * 8, v);
        const unsigned long long h = drop_bits(seed, i);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const bool keep = ((unsigned)(h >> (8 * j)) & 0xFFu) >= thr;
            v[j] = keep ? v[j] * inv_keep : 0.f;
            if (mask_out) mask_out[row * K + c * 8 + j] = keep ? 1 : 0;
        }
        store8(out + row * ldo + c * 8, v);
    }
}
// y += t * keep / (1 - p) with the SAME mask (adapter branch of the input gradient)
__global__ void dropout_add_kernel(const bf16* __restrict__ t, int ldt, bf16* __restrict__ y, int ldy, long long M, int K,
                                   unsigned thr, float inv_keep, unsigned long long seed) {
    const int cpr = K / 8;
    const long long total = M * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / cpr; const int c = (int)(i % cpr);
        float a[8], b[8];
        load8(t + row * ldt + c * 8, a);
        load8(y + row * ldy + c * 8, b);
        const unsigned long long h = drop_bits(seed, i);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (((unsigned)(h >> (8 * j)) & 0xFFu) >= thr) b[j] += a[j] * inv_keep;
        store8(y + row * ldy + c * 8, b);
    }
}

static inline int ew_grid(long long total, int threads) {
    long long b = ceil_div_ll(total, threads);
    long long cap = (long long)num_sms() * 16;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace vlb

using namespace vlb;

#define NORM_DISPATCH(KERNEL, ...)                                                                   \
    if (H <= 8 * 128 * 1)       KERNEL<128, 1><<<(unsigned)M, 128, 0, st>>>(__VA_ARGS__);            \
    else if (H <= 8 * 128 * 2)  KERNEL<128, 2><<<(unsigned)M, 128, 0, st>>>(__VA_ARGS__);            \
    else if (H <= 8 * 128 * 4)  KERNEL<128, 4><<<(unsigned)M, 128, 0, st>>>(__VA_ARGS__);            \
    else                        KERNEL<256, 4><<<(unsigned)M, 256, 0, st>>>(__VA_ARGS__);

extern "C" int vlb_rmsnorm_fwd(const void* x, const void* w, void* y, void* rstd, long long M, int H, float eps,
                               void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8192 && M > 0, "rmsnorm_fwd: bad shape M=%lld H=%d", M, H);
    cudaStream_t st = (cudaStream_t)stream;
    NORM_DISPATCH(rmsnorm_fwd_kernel, (const bf16*)x, (const bf16*)w, (bf16*)y, (float*)rstd, H, eps)
    return check_launch("rmsnorm_fwd");
}

extern "C" int vlb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd, const void* dres,
                               void* dx, long long M, int H, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8192 && M > 0, "rmsnorm_bwd: bad shape M=%lld H=%d", M, H);
    cudaStream_t st = (cudaStream_t)stream;
    NORM_DISPATCH(rmsnorm_bwd_kernel, (const bf16*)dy, (const bf16*)x, (const bf16*)w, (const float*)rstd,
                  (const bf16*)dres, (bf16*)dx, H)
    return check_launch("rmsnorm_bwd");
}

extern "C" int vlb_layernorm_fwd(const void* x, const void* w, const void* b, const void* residual, void* y,
                                 long long M, int H, float eps, int act, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8192 && M > 0, "layernorm_fwd: bad shape M=%lld H=%d", M, H);
    cudaStream_t st = (cudaStream_t)stream;
    if (H <= 1024 && M >= 1024) {
        const unsigned grid = (unsigned)ceil_div_ll(M, 8);
        if (H <= 512) layernorm_fwd_warp_kernel<2><<<grid, 256, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)b,
                                                                          (const bf16*)residual, (bf16*)y, M, H, eps, act);
        else layernorm_fwd_warp_kernel<4><<<grid, 256, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)b,
                                                                 (const bf16*)residual, (bf16*)y, M, H, eps, act);
        return check_launch("layernorm_fwd");
    }
    NORM_DISPATCH(layernorm_fwd_kernel, (const bf16*)x, (const bf16*)w, (const bf16*)b, (const bf16*)residual,
                  (bf16*)y, H, eps, act)
    return check_launch("layernorm_fwd");
}

extern "C" int vlb_rope(void* qkv, int ld, const void* cos_tab, const void* sin_tab, long long M, int S, int nheads,
                        int D, int inverse, void* stream) {
    VLB_REQUIRE(D % 16 == 0 && ld % 8 == 0, "rope: D=%d ld=%d", D, ld);
    const long long total = M * nheads * (D / 16);
    rope_kernel<<<ew_grid(total, 256), 256, 0, (cudaStream_t)stream>>>((bf16*)qkv, ld, (const bf16*)cos_tab,
                                                                      (const bf16*)sin_tab, M, S, nheads, D, inverse);
    return check_launch("rope");
}

extern "C" int vlb_rope_bwd_f32(const void* dq_f32, void* dqkv, int ld, const void* cos_tab, const void* sin_tab,
                                long long M, int S, int nheads, int D, void* stream) {
    VLB_REQUIRE(D % 16 == 0 && ld % 8 == 0, "rope_bwd_f32: D=%d ld=%d", D, ld);
    const long long total = M * nheads * (D / 16);
    rope_bwd_f32_kernel<<<ew_grid(total, 256), 256, 0, (cudaStream_t)stream>>>(
        (const float*)dq_f32, (bf16*)dqkv, ld, (const bf16*)cos_tab, (const bf16*)sin_tab, M, S, nheads, D);
    return check_launch("rope_bwd_f32");
}

extern "C" int vlb_swiglu_fwd(const void* gu, void* act, long long M, int I, void* stream) {
    VLB_REQUIRE(I % 8 == 0, "swiglu: I=%d", I);
    swiglu_fwd_kernel<<<ew_grid(M * (I / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)gu, (bf16*)act, M, I);
    return check_launch("swiglu_fwd");
}

extern "C" int vlb_swiglu_bwd(const void* gu, const void* dact, void* dgu, long long M, int I, void* stream) {
    VLB_REQUIRE(I % 8 == 0, "swiglu: I=%d", I);
    swiglu_bwd_kernel<<<ew_grid(M * (I / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)gu, (const bf16*)dact,
                                                                                  (bf16*)dgu, M, I);
    return check_launch("swiglu_bwd");
}

// p is quantised to 1/256 (0.1 -> 26/256 = 0.1016); the scale uses the quantised value so E[dropout(x)] = x exactly
static inline void drop_params(float p, unsigned* thr, float* inv_keep) {
    unsigned t = (unsigned)(p * 256.f + 0.5f);
    if (t > 255) t = 255;
    *thr = t; *inv_keep = 256.f / (float)(256 - t);
}

extern "C" int vlb_dropout(const void* x, int ldx, void* out, int ldo, void* mask_u8, long long M, int K, float p,
                           unsigned long long seed, void* stream) {
    VLB_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldo % 8 == 0 && p >= 0.f && p < 1.f, "dropout: K=%d ld=%d,%d p=%g", K, ldx, ldo, p);
    unsigned thr; float ik; drop_params(p, &thr, &ik);
    dropout_kernel<<<ew_grid(M * (K / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, ldx, (bf16*)out, ldo,
                                                                               (uint8_t*)mask_u8, M, K, thr, ik, seed);
    return check_launch("dropout");
}

extern "C" int vlb_dropout_add(const void* t, int ldt, void* y, int ldy, long long M, int K, float p,
                               unsigned long long seed, void* stream) {
    VLB_REQUIRE(K % 8 == 0 && ldt % 8 == 0 && ldy % 8 == 0 && p >= 0.f && p < 1.f, "dropout_add: K=%d ld=%d,%d p=%g", K, ldt, ldy, p);
    unsigned thr; float ik; drop_params(p, &thr, &ik);
    dropout_add_kernel<<<ew_grid(M * (K / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)t, ldt, (bf16*)y, ldy, M, K,
                                                                                   thr, ik, seed);
    return check_launch("dropout_add");
}
