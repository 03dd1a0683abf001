// Brain head: final RMSNorm -> LayerNorm1 -> HRF-weighted pooling over the token axis -> LayerNorm2 (-> dropout)
// -> [readout GEMM runs on the tcgen05 kernel] -> MSE + ridge penalty, and the matching backward sweep.
//   reference: src/utils.py:40-73 (HRFConvolveLayer einsum 'bse,bs->be', RidgeRegressionLayer),
//              src/litmodule/videollama2_vlb_litmodule.py:245-254 (layer_norm1/2, dropout), :302 (mse_loss + l2_reg)
// HBM-bound: every kernel reads its dominant operand exactly once with 16-byte accesses; rows whose HRF weight is 0
// (system prompt, instruction, padding: ~60 % of the 2048 tokens) are never fetched in the forward pass.
// Rounding points mirror bf16 autocast: LN outputs fp32 -> einsum operands bf16 -> pooled bf16 -> LN2 fp32 -> bf16.
#include "common.cuh"
#include <mutex>

namespace vlb {

constexpr int HT = 128;       // threads per CTA in the row kernels
constexpr int HV = 4;         // 8-element chunks per thread (H <= 8*HT*HV = 4096)

struct RowStats { float rstd_rms, mean1, rstd1; };

// this thread's 16-byte chunks of one row, still packed (issued early so that the next row is in flight during the reductions)
__device__ __forceinline__ void load_row_raw(const bf16* __restrict__ xr, int nchunk, uint4 (&raw)[HV]) {
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        raw[i] = c < nchunk ? __ldg(reinterpret_cast<const uint4*>(xr) + c) : make_uint4(0, 0, 0, 0);
    }
}

// first live (weight != 0) row at or after r in the CTA's staged weights
__device__ __forceinline__ int next_live(const float* wsm, int r, int n) {
    while (r < n && wsm[r] == 0.f) ++r;
    return r;
}

__device__ __forceinline__ void unpack_row(const uint4 (&raw)[HV], int nchunk, float (&z)[HV][8]) {
#pragma unroll
    for (int i = 0; i < HV; ++i)
        if (threadIdx.x + i * HT < nchunk) unpack8(raw[i], z[i]);
}

// z holds the row on entry: RMSNorm (bf16 rounding as MistralRMSNorm) -> z, then LN1 statistics over z.
// The arithmetic below is written on pairs (FFMA2 / FMUL2 / FADD2 and one packed convert per pair, sm_100): with three bf16 rounding
// points per element these row kernels are bound by instruction issue (ncu: 946 instructions per warp and row in the scalar form, 30 per
// element, issue slots 44 % busy at 3 warps per scheduler), not by HBM.
__device__ __forceinline__ float2 round_bf16_pair(float2 v) {          // float(bf16(v)) for both lanes: one F2FP + two shifts
    const uint32_t u = pack_bf16(v.x, v.y);
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xFFFF0000u));
}
__device__ __forceinline__ float2 unpack_pair(uint32_t u) { return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xFFFF0000u)); }

__device__ __forceinline__ RowStats row_norms_z(float (&z)[HV][8], const bf16* __restrict__ wn, int H,
                                                float eps_rms, float eps_ln, float* red) {
    const int nchunk = H / 8;
    float2 ss2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        if (threadIdx.x + i * HT < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; j += 2) { const float2 v = make_float2(z[i][j], z[i][j + 1]); ss2 = __ffma2_rn(v, v, ss2); }
        }
    }
    RowStats st;
    st.rstd_rms = rsqrtf(block_sum(ss2.x + ss2.y, red) / (float)H + eps_rms);
    const float2 r2 = make_float2(st.rstd_rms, st.rstd_rms);
    float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
            const uint4 wv = __ldg(reinterpret_cast<const uint4*>(wn) + c);
            const uint32_t wu[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                const float2 t = round_bf16_pair(__fmul2_rn(make_float2(z[i][j], z[i][j + 1]), r2));      // bf16(x * rstd)
                const float2 y = round_bf16_pair(__fmul2_rn(unpack_pair(wu[j >> 1]), t));                 // bf16(w * .)
                z[i][j] = y.x; z[i][j + 1] = y.y;
                s2 = __fadd2_rn(s2, y);
            }
        }
    }
    st.mean1 = block_sum(s2.x + s2.y, red) / (float)H;
    const float2 nm2 = make_float2(-st.mean1, -st.mean1);
    float2 sq2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        if (threadIdx.x + i * HT < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; j += 2) { const float2 d = __fadd2_rn(make_float2(z[i][j], z[i][j + 1]), nm2); sq2 = __ffma2_rn(d, d, sq2); }
        }
    }
    st.rstd1 = rsqrtf(block_sum(sq2.x + sq2.y, red) / (float)H + eps_ln);
    return st;
}

// loads row x, applies RMSNorm -> z, then LN1 statistics over z.
__device__ __forceinline__ RowStats row_norms(const bf16* __restrict__ xr, const bf16* __restrict__ wn, int H,
                                              float eps_rms, float eps_ln, float (&z)[HV][8], float* red) {
    uint4 raw[HV];
    load_row_raw(xr, H / 8, raw);
    unpack_row(raw, H / 8, z);
    return row_norms_z(z, wn, H, eps_rms, eps_ln, red);
}

// grid (nch, B): CTA accumulates sum_s w[b,s] * bf16(LN1(RMSNorm(x[b,s]))) over its slice of rows -> partial[b, ch, :]
// The CTA's HRF weights are staged in shared memory once (one coalesced load instead of a dependent global load per row) and the
// next live row is fetched before the three block reductions of the current one: the kernel is latency-bound (a few live rows
// per CTA, each a load -> reduce -> reduce -> reduce chain), not bandwidth-bound.
__global__ void __launch_bounds__(HT) hrf_pool_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ wn,
                                                          const bf16* __restrict__ g1, const bf16* __restrict__ b1,
                                                          const float* __restrict__ wmask, float* __restrict__ partial,
                                                          int S, int H, int rows_per, float eps_rms, float eps_ln) {
    __shared__ float red[32];
    __shared__ float wsm[HT];
    const int b = blockIdx.y, ch = blockIdx.x, nch = gridDim.x;
    const int nchunk = H / 8;
    float acc[HV][8];
#pragma unroll
    for (int i = 0; i < HV; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const int s0 = ch * rows_per, s1 = min(S, s0 + rows_per);
    for (int base = s0; base < s1; base += HT) {
        const int n = min(HT, s1 - base);
        __syncthreads();                                           // the previous batch of weights is no longer read
        if (threadIdx.x < n) wsm[threadIdx.x] = wmask[(size_t)b * S + base + threadIdx.x];
        __syncthreads();
        const bf16* xb = x + ((size_t)b * S + base) * H;
        int r = next_live(wsm, 0, n);                              // uniform across the CTA
        uint4 raw[HV];
        if (r < n) load_row_raw(xb + (size_t)r * H, nchunk, raw);
        while (r < n) {
            const float wgt = wsm[r];
            float z[HV][8];
            unpack_row(raw, nchunk, z);
            r = next_live(wsm, r + 1, n);
            if (r < n) load_row_raw(xb + (size_t)r * H, nchunk, raw);                  // in flight during the reductions below
            const RowStats st = row_norms_z(z, wn, H, eps_rms, eps_ln, red);
            const float2 nm2 = make_float2(-st.mean1, -st.mean1), rs2 = make_float2(st.rstd1, st.rstd1), w2 = make_float2(wgt, wgt);
#pragma unroll
            for (int i = 0; i < HV; ++i) {
                const int c = threadIdx.x + i * HT;
                if (c < nchunk) {
                    const uint4 gv = __ldg(reinterpret_cast<const uint4*>(g1) + c), bv = __ldg(reinterpret_cast<const uint4*>(b1) + c);
                    const uint32_t gu[4] = {gv.x, gv.y, gv.z, gv.w}, bu[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                    for (int j = 0; j < 8; j += 2) {
                        const float2 h = __fmul2_rn(__fadd2_rn(make_float2(z[i][j], z[i][j + 1]), nm2), rs2);         // (z - mean) * rstd
                        const float2 y = round_bf16_pair(__ffma2_rn(h, unpack_pair(gu[j >> 1]), unpack_pair(bu[j >> 1])));
                        const float2 a2 = __ffma2_rn(w2, y, make_float2(acc[i][j], acc[i][j + 1]));
                        acc[i][j] = a2.x; acc[i][j + 1] = a2.y;
                    }
                }
            }
        }
    }
    float* dst = partial + ((size_t)b * nch + ch) * H;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) store8(dst + c * 8, acc[i]);
    }
}

__device__ __forceinline__ float hash_uniform(unsigned long long seed, unsigned long long idx) {
    unsigned long long z = seed + idx * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

// one CTA per batch row: pooled = bf16(sum of partials); hrf = bf16(dropout(LN2(pooled)))
__global__ void __launch_bounds__(HT) hrf_pool_finalize_kernel(const float* __restrict__ partial, const bf16* __restrict__ g2,
                                                               const bf16* __restrict__ b2, bf16* __restrict__ pooled,
                                                               bf16* __restrict__ hrf, float* __restrict__ keep_scale,
                                                               int nch, int H, float eps, float drop_p,
                                                               unsigned long long seed) {
    __shared__ float red[32];
    const int b = blockIdx.x, nchunk = H / 8;
    float v[HV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
#pragma unroll 8
            for (int k = 0; k < nch; ++k) {                       // fixed summation order; the loads are independent
                float t[8]; load8(partial + ((size_t)b * nch + k) * H + c * 8, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[i][j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) { v[i][j] = bf2f(f2bf(v[i][j])); s += v[i][j]; }
            store8(pooled + (size_t)b * H + c * 8, v[i]);
        }
    }
    const float mean = block_sum(s, red) / (float)H;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; sq += d * d; }
        }
    }
    const float rstd = rsqrtf(block_sum(sq, red) / (float)H + eps);
    const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
            float g[8], bb[8], o[8], ks[8];
            load8(g2 + c * 8, g); load8(b2 + c * 8, bb);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                ks[j] = 1.f;
                if (drop_p > 0.f)
                    ks[j] = hash_uniform(seed, (unsigned long long)b * H + c * 8 + j) < drop_p ? 0.f : inv_keep;
                o[j] = ((v[i][j] - mean) * rstd * g[j] + bb[j]) * ks[j];
            }
            store8(hrf + (size_t)b * H + c * 8, o);
            if (keep_scale) store8(keep_scale + (size_t)b * H + c * 8, ks);
        }
    }
}

// d = pred - bf16(y); acc[0] += sum d^2; dpred = bf16(2 d / n)   (F.mse_loss, reduction mean over B*P)
__global__ void head_loss_kernel(const bf16* __restrict__ pred, const float* __restrict__ y, bf16* __restrict__ dpred,
                                 float* __restrict__ acc, long long n, long long n_total) {
    __shared__ float red[32];
    float s = 0.f;
    const float k = 2.f / (float)n_total;     // n_total = B * P over ALL voxel shards (== n when the head is not sharded)
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d = bf2f(pred[i]) - bf2f(f2bf(y[i]));
        s += d * d;
        if (dpred) dpred[i] = f2bf(k * d);
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(acc, s);
}

// grad = 2*lambda*W (fp32, written or accumulated) and acc[1] += sum W^2 : one sweep over the readout weight
__global__ void l2_prefill_kernel(const bf16* __restrict__ W, float* __restrict__ grad, float* __restrict__ acc,
                                  long long n8, float two_lambda, int write_grad) {
    __shared__ float red[32];
    float s = 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float w[8]; load8(W + i * 8, w);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s += w[j] * w[j]; w[j] *= two_lambda; }
        if (write_grad) store8(grad + i * 8, w);
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(acc + 1, s);
}

// loss = acc[0]/n + lambda*acc[1]  -> out[0]; out[1] = mse, out[2] = l2
__global__ void loss_finalize_kernel(const float* __restrict__ acc, float* __restrict__ out, float inv_n, float lambda) {
    const float mse = acc[0] * inv_n, l2 = lambda * acc[1];
    out[0] = mse + l2; out[1] = mse; out[2] = l2;
}

// dbias[p] = sum_b dpred[b, p]
__global__ void colsum_kernel(const bf16* __restrict__ d, float* __restrict__ out, int B, long long P) {
    for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < P; p += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += bf2f(d[(size_t)b * P + p]);
        out[p] = s;
    }
}

// LN2 backward on [B, H] (one CTA per row): dy fp32 (grad wrt hrf, before dropout scaling), x = pooled (bf16)
__global__ void __launch_bounds__(HT) ln2_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ keep_scale,
                                                     const bf16* __restrict__ pooled, const bf16* __restrict__ g2,
                                                     float* __restrict__ dpooled, float* __restrict__ dg2,
                                                     float* __restrict__ db2, int H, float eps) {
    __shared__ float red[32];
    const int b = blockIdx.x, nchunk = H / 8;
    float v[HV][8], g[HV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
            load8(pooled + (size_t)b * H + c * 8, v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) s += v[i][j];
        }
    }
    const float mean = block_sum(s, red) / (float)H;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { v[i][j] -= mean; sq += v[i][j] * v[i][j]; }
        }
    }
    const float rstd = rsqrtf(block_sum(sq, red) / (float)H + eps);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
            float gw[8];
            load8(dy + (size_t)b * H + c * 8, g[i]);
            load8(g2 + c * 8, gw);
            if (keep_scale) {
                float ks[8]; load8(keep_scale + (size_t)b * H + c * 8, ks);
#pragma unroll
                for (int j = 0; j < 8; ++j) g[i][j] *= ks[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = v[i][j] * rstd;
                atomicAdd(dg2 + c * 8 + j, g[i][j] * xh);
                atomicAdd(db2 + c * 8 + j, g[i][j]);
                g[i][j] *= gw[j];
                v[i][j] = xh;
                m1 += g[i][j]; m2 += g[i][j] * xh;
            }
        }
    }
    m1 = block_sum(m1, red) / (float)H;
    m2 = block_sum(m2, red) / (float)H;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
        if (c < nchunk) {
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - m1 - v[i][j] * m2);
            store8(dpooled + (size_t)b * H + c * 8, o);
        }
    }
}

// fire-and-forget 16-byte fp32 reduction (sm_90+)
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Backward of pooling + LN1 + final RMSNorm. grid (nch, B). For each row: dx = 0 if w == 0, else
//   gu = w * bf16?(dpooled)   (grad wrt LN1 output), LN1 backward -> gz, RMSNorm backward -> dx (bf16).
// dg1/db1 accumulate per CTA in registers and are flushed with one atomicAdd per element.
__global__ void __launch_bounds__(HT) hrf_pool_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ wn,
                                                          const bf16* __restrict__ g1, const float* __restrict__ wmask,
                                                          const float* __restrict__ dpooled, bf16* __restrict__ dx,
                                                          float* __restrict__ dg1, float* __restrict__ db1, int S, int H,
                                                          int rows_per, float eps_rms, float eps_ln) {
    __shared__ float red[32];
    const int b = blockIdx.y, ch = blockIdx.x, nchunk = H / 8;
    float ag[HV][8], ab[HV][8], dp[HV][8];
#pragma unroll
    for (int i = 0; i < HV; ++i) {
        const int c = threadIdx.x + i * HT;
#pragma unroll
        for (int j = 0; j < 8; ++j) { ag[i][j] = 0.f; ab[i][j] = 0.f; dp[i][j] = 0.f; }
        if (c < nchunk) load8(dpooled + (size_t)b * H + c * 8, dp[i]);
    }
    const int s0 = ch * rows_per, s1 = min(S, s0 + rows_per);
    bool any = false;
    for (int s = s0; s < s1; ++s) {
        const float wgt = wmask[(size_t)b * S + s];
        bf16* dxr = dx + ((size_t)b * S + s) * H;
        if (wgt == 0.f) {
            if (dx != nullptr) {
#pragma unroll
                for (int i = 0; i < HV; ++i) {
                    const int c = threadIdx.x + i * HT;
                    if (c < nchunk) *reinterpret_cast<uint4*>(dxr + c * 8) = make_uint4(0, 0, 0, 0);
                }
            }
            continue;
        }
        any = true;
        const bf16* xr = x + ((size_t)b * S + s) * H;
        float z[HV][8];
        const RowStats st = row_norms(xr, wn, H, eps_rms, eps_ln, z, red);
        // LN1 backward
        float2 m1p = make_float2(0.f, 0.f), m2p = make_float2(0.f, 0.f);
        const float2 nm2 = make_float2(-st.mean1, -st.mean1), rs2 = make_float2(st.rstd1, st.rstd1), w2 = make_float2(wgt, wgt);
        float gz[HV][8];
#pragma unroll
        for (int i = 0; i < HV; ++i) {
            const int c = threadIdx.x + i * HT;
            if (c < nchunk) {
                const uint4 gv = __ldg(reinterpret_cast<const uint4*>(g1) + c);
                const uint32_t gwu[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
                for (int j = 0; j < 8; j += 2) {                 // pairs: FMUL2 / FFMA2 / FADD2
                    const float2 gu = __fmul2_rn(w2, make_float2(dp[i][j], dp[i][j + 1]));
                    const float2 uh = __fmul2_rn(__fadd2_rn(make_float2(z[i][j], z[i][j + 1]), nm2), rs2);
                    const float2 a2 = __ffma2_rn(gu, uh, make_float2(ag[i][j], ag[i][j + 1]));
                    const float2 b2 = __fadd2_rn(make_float2(ab[i][j], ab[i][j + 1]), gu);
                    const float2 g2 = __fmul2_rn(gu, unpack_pair(gwu[j >> 1]));
                    ag[i][j] = a2.x; ag[i][j + 1] = a2.y; ab[i][j] = b2.x; ab[i][j + 1] = b2.y;
                    gz[i][j] = g2.x; gz[i][j + 1] = g2.y;
                    z[i][j] = uh.x; z[i][j + 1] = uh.y;
                    m1p = __fadd2_rn(m1p, g2); m2p = __ffma2_rn(g2, uh, m2p);
                }
            }
        }
        if (dx == nullptr) continue;          // frozen backbone: only the LN1 gradients are needed
        const float m1 = block_sum(m1p.x + m1p.y, red) / (float)H;
        const float m2 = block_sum(m2p.x + m2p.y, red) / (float)H;
        // RMSNorm backward: y = wn * (x * rstd); g = gz * wn; dx = rstd * (g - xh * mean(g * xh))
        float2 dotp = make_float2(0.f, 0.f);
        const float2 nm1 = make_float2(-m1, -m1), nm2b = make_float2(-m2, -m2), rr2 = make_float2(st.rstd_rms, st.rstd_rms);
        float xh[HV][8];
#pragma unroll
        for (int i = 0; i < HV; ++i) {
            const int c = threadIdx.x + i * HT;
            if (c < nchunk) {
                const uint4 wv = __ldg(reinterpret_cast<const uint4*>(wn) + c), xv = __ldg(reinterpret_cast<const uint4*>(xr) + c);
                const uint32_t wu[4] = {wv.x, wv.y, wv.z, wv.w}, xu[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    // gz = rstd1 * (gz - m1 - uh * m2) * w ; xh = x * rstd_rms ; dot += gz * xh
                    const float2 t = __ffma2_rn(make_float2(z[i][j], z[i][j + 1]), nm2b, __fadd2_rn(make_float2(gz[i][j], gz[i][j + 1]), nm1));
                    const float2 g2 = __fmul2_rn(__fmul2_rn(rs2, t), unpack_pair(wu[j >> 1]));
                    const float2 xh2 = __fmul2_rn(unpack_pair(xu[j >> 1]), rr2);
                    gz[i][j] = g2.x; gz[i][j + 1] = g2.y; xh[i][j] = xh2.x; xh[i][j + 1] = xh2.y;
                    dotp = __ffma2_rn(g2, xh2, dotp);
                }
            }
        }
        const float dot = block_sum(dotp.x + dotp.y, red) / (float)H;
        const float2 nd2 = make_float2(-dot, -dot);
#pragma unroll
        for (int i = 0; i < HV; ++i) {
            const int c = threadIdx.x + i * HT;
            if (c < nchunk) {
                uint4 o;
                uint32_t* ou = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    const float2 v = __fmul2_rn(rr2, __ffma2_rn(make_float2(xh[i][j], xh[i][j + 1]), nd2, make_float2(gz[i][j], gz[i][j + 1])));
                    ou[j >> 1] = pack_bf16(v.x, v.y);
                }
                *reinterpret_cast<uint4*>(dxr + c * 8) = o;
            }
        }
    }
    if (any) {                                 // 512 CTAs add into the same 2 x H floats: 16-byte reductions, a quarter of the operations
#pragma unroll
        for (int i = 0; i < HV; ++i) {
            const int c = threadIdx.x + i * HT;
            if (c < nchunk) {
                red_add_v4(dg1 + c * 8, ag[i][0], ag[i][1], ag[i][2], ag[i][3]);
                red_add_v4(dg1 + c * 8 + 4, ag[i][4], ag[i][5], ag[i][6], ag[i][7]);
                red_add_v4(db1 + c * 8, ab[i][0], ab[i][1], ab[i][2], ab[i][3]);
                red_add_v4(db1 + c * 8 + 4, ab[i][4], ab[i][5], ab[i][6], ab[i][7]);
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------- cached features
// With a frozen backbone the LN1 affine commutes with the pooling: sum_s w_s (g1 * n_s + b1) = g1 * u + b1 * sw with
// u = sum_s w_s n_s and sw = sum_s w_s, so (u, sw) is a per-sample feature that never changes while the head trains.
// pool_reduce: partial[b, nch, H] (hrf_pool_fwd run with g1 = 1, b1 = 0) -> u[b, H] fp32 ; sw[b] = sum_s wmask[b, s]
__global__ void __launch_bounds__(HT) pool_reduce_kernel(const float* __restrict__ partial, const float* __restrict__ wmask,
                                                         float* __restrict__ u, float* __restrict__ sw, int nch, int H, int S) {
    __shared__ float red[32];
    const int b = blockIdx.x, nchunk = H / 8;
    for (int c = threadIdx.x; c < nchunk; c += HT) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = 0.f;
        for (int k = 0; k < nch; ++k) {
            float t[8]; load8(partial + ((size_t)b * nch + k) * H + c * 8, t);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] += t[j];
        }
        store8(u + (size_t)b * H + c * 8, v);
    }
    float s = 0.f;
    for (int i = threadIdx.x; i < S; i += HT) s += wmask[(size_t)b * S + i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) sw[b] = s;
}

// pooled_pre[b, h] = g1[h] * u[b, h] + b1[h] * sw[b]   (fp32; consumed by hrf_pool_finalize with nch = 1)
__global__ void ln1_affine_fwd_kernel(const float* __restrict__ u, const float* __restrict__ sw, const bf16* __restrict__ g1,
                                      const bf16* __restrict__ b1, float* __restrict__ out, int B, int H) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)B * H) return;
    const int h = (int)(i % H); const int b = (int)(i / H);
    out[i] = bf2f(g1[h]) * u[i] + bf2f(b1[h]) * sw[b];
}

// dg1[h] += sum_b dpooled[b, h] * u[b, h] ; db1[h] += sum_b dpooled[b, h] * sw[b]
__global__ void ln1_affine_bwd_kernel(const float* __restrict__ dpooled, const float* __restrict__ u,
                                      const float* __restrict__ sw, float* __restrict__ dg1, float* __restrict__ db1,
                                      int B, int H) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float ag = 0.f, ab = 0.f;
    for (int b = 0; b < B; ++b) {
        const float d = dpooled[(size_t)b * H + h];
        ag += d * u[(size_t)b * H + h]; ab += d * sw[b];
    }
    dg1[h] += ag; db1[h] += ab;
}


// ||dW||^2 of the readout weight WITHOUT forming dW = dpred^T hrf + 2*lambda*W (rank-N structure):
//   ||dpred^T hrf||^2 = <Gd, Gh>,  Gd = dpred dpred^T, Gh = hrf hrf^T  (N x N Gram matrices from two small GEMMs)
//   2 <dpred^T hrf, 2 lambda W> = 4 lambda sum_{n,p} dpred[n,p] * (pred[n,p] - b[p])      (pred - b = hrf W^T)
//   ||2 lambda W||^2 = 4 lambda^2 * wsq
// sumsq += all three. grid-stride over the N x P elements for the cross term; block 0 adds the Gram and W terms.
__device__ DetSumWs g_gradnorm_ws;

__global__ void head_gradnorm_kernel(const bf16* __restrict__ dpred, const bf16* __restrict__ pred, const bf16* __restrict__ bias,
                                     const float* __restrict__ gd, const float* __restrict__ gh, const float* __restrict__ wsq,
                                     float* __restrict__ sumsq, int N, int ldg, long long P, float lambda) {
    __shared__ float red[32];
    float s = 0.f;
    const long long total = (long long)N * P;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long pcol = i % P;
        s += bf2f(dpred[i]) * (bf2f(pred[i]) - bf2f(bias[pcol]));
    }
    s *= 4.f * lambda;
    if (blockIdx.x == 0) {
        float g = 0.f;
        for (int i = threadIdx.x; i < N * N; i += blockDim.x) g += gd[(i / N) * ldg + i % N] * gh[(i / N) * ldg + i % N];
        s += g;
        if (threadIdx.x == 0) s += 4.f * lambda * lambda * wsq[0];
    }
    s = block_sum(s, red);
    det_grid_add(g_gradnorm_ws, s, sumsq, red);
}

// ---- HRFConvolveLayer on its own (src/utils.py:40-56): out[b, e] = sum_s w[b, s] * emb[b, s, e].
// grid (col blocks, nch, B): a thread owns 8 adjacent columns and walks its slice of the token axis (rows with weight 0 are
// skipped without being fetched); partial sums meet in the zeroed fp32 output through red.add.
__global__ void __launch_bounds__(HT) hrf_convolve_fwd_kernel(const bf16* __restrict__ emb, const float* __restrict__ w,
                                                              float* __restrict__ out, int S, int H, int rows_per) {
    const int b = blockIdx.z, ch = blockIdx.y;
    const int c = blockIdx.x * HT + threadIdx.x;
    if (c * 8 >= H) return;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const int s0 = ch * rows_per, s1 = min(S, s0 + rows_per);
    for (int s = s0; s < s1; ++s) {
        const float wgt = w[(size_t)b * S + s];
        if (wgt == 0.f) continue;
        float v[8]; load8(emb + ((size_t)b * S + s) * H + c * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += wgt * v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(out + (size_t)b * H + c * 8 + j, acc[j]);
}

// backward of the einsum, one CTA per token row: d_emb[b,s,:] = w[b,s] * g[b,:] (bf16), d_w[b,s] = <emb[b,s,:], g[b,:]>
__global__ void __launch_bounds__(HT) hrf_convolve_bwd_kernel(const bf16* __restrict__ emb, const float* __restrict__ w,
                                                              const float* __restrict__ g, bf16* __restrict__ demb,
                                                              float* __restrict__ dw, int S, int H) {
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const int b = (int)(row / S);
    const float wgt = w[row];
    float dot = 0.f;
    for (int c = threadIdx.x; c * 8 < H; c += HT) {
        float gv[8]; load8(g + (size_t)b * H + c * 8, gv);
        if (dw) {
            float v[8]; load8(emb + row * H + c * 8, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) dot += v[j] * gv[j];
        }
        if (demb) {
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = wgt * gv[j];
            store8(demb + row * H + c * 8, o);
        }
    }
    if (dw) {
        dot = block_sum(dot, red);
        if (threadIdx.x == 0) dw[row] = dot;
    }
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_hrf_convolve_fwd(const void* emb, const void* w_f32, void* out_f32, int B, int S, int H, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && B > 0 && S > 0, "hrf_convolve_fwd: B=%d S=%d H=%d", B, S, H);
    cudaError_t e = cudaMemsetAsync(out_f32, 0, (size_t)B * H * sizeof(float), (cudaStream_t)stream);
    VLB_REQUIRE(e == cudaSuccess, "hrf_convolve_fwd: memset failed (%s)", cudaGetErrorString(e));
    const int colb = ceil_div(H / 8, HT);
    int nch = ceil_div(num_sms() * 4, B * colb);
    if (nch > S) nch = S;
    if (nch < 1) nch = 1;
    hrf_convolve_fwd_kernel<<<dim3(colb, nch, B), HT, 0, (cudaStream_t)stream>>>((const bf16*)emb, (const float*)w_f32,
                                                                                (float*)out_f32, S, H, ceil_div(S, nch));
    return check_launch("hrf_convolve_fwd");
}

extern "C" int vlb_hrf_convolve_bwd(const void* emb, const void* w_f32, const void* dout_f32, void* demb, void* dw_f32,
                                    int B, int S, int H, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && B > 0 && S > 0 && (demb || dw_f32), "hrf_convolve_bwd: B=%d S=%d H=%d", B, S, H);
    hrf_convolve_bwd_kernel<<<(unsigned)((size_t)B * S), HT, 0, (cudaStream_t)stream>>>(
        (const bf16*)emb, (const float*)w_f32, (const float*)dout_f32, (bf16*)demb, (float*)dw_f32, S, H);
    return check_launch("hrf_convolve_bwd");
}

extern "C" int vlb_hrf_pool_fwd(const void* x, const void* norm_w, const void* ln1_w, const void* ln1_b,
                                const void* wmask_f32, void* partial_f32, int B, int S, int H, int nch, float eps_rms,
                                float eps_ln, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8 * HT * HV && nch > 0, "hrf_pool_fwd: H=%d nch=%d", H, nch);
    const int rows_per = ceil_div(S, nch);
    hrf_pool_fwd_kernel<<<dim3(nch, B), HT, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (const bf16*)norm_w, (const bf16*)ln1_w, (const bf16*)ln1_b, (const float*)wmask_f32,
        (float*)partial_f32, S, H, rows_per, eps_rms, eps_ln);
    return check_launch("hrf_pool_fwd");
}

extern "C" int vlb_hrf_pool_finalize(const void* partial_f32, const void* ln2_w, const void* ln2_b, void* pooled,
                                     void* hrf, void* keep_scale_f32, int B, int H, int nch, float eps, float drop_p,
                                     unsigned long long seed, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8 * HT * HV, "hrf_pool_finalize: H=%d", H);
    hrf_pool_finalize_kernel<<<B, HT, 0, (cudaStream_t)stream>>>((const float*)partial_f32, (const bf16*)ln2_w,
                                                                (const bf16*)ln2_b, (bf16*)pooled, (bf16*)hrf,
                                                                (float*)keep_scale_f32, nch, H, eps, drop_p, seed);
    return check_launch("hrf_pool_finalize");
}

extern "C" int vlb_head_loss(const void* pred, const void* y_f32, void* dpred, void* acc_f32, long long n,
                             long long n_total, void* stream) {
    VLB_REQUIRE(n_total >= n && n > 0, "head_loss: n=%lld n_total=%lld", n, n_total);
    int grid = (int)((n + 255) / 256); if (grid > num_sms() * 8) grid = num_sms() * 8;
    head_loss_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)pred, (const float*)y_f32, (bf16*)dpred,
                                                            (float*)acc_f32, n, n_total);
    return check_launch("head_loss");
}

extern "C" int vlb_l2_prefill(const void* W, void* grad_f32, void* acc_f32, long long n, float lambda, int write_grad,
                              void* stream) {
    VLB_REQUIRE(n % 8 == 0, "l2_prefill: n=%lld must be a multiple of 8", n);
    int grid = (int)((n / 8 + 255) / 256); if (grid > num_sms() * 8) grid = num_sms() * 8;
    l2_prefill_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)W, (float*)grad_f32, (float*)acc_f32, n / 8,
                                                             2.f * lambda, write_grad);
    return check_launch("l2_prefill");
}

extern "C" int vlb_loss_finalize(const void* acc_f32, void* out_f32, long long n, float lambda, void* stream) {
    loss_finalize_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const float*)acc_f32, (float*)out_f32, 1.f / (float)n, lambda);
    return check_launch("loss_finalize");
}

extern "C" int vlb_colsum(const void* d, void* out_f32, int B, long long P, void* stream) {
    int grid = (int)((P + 255) / 256); if (grid > num_sms() * 8) grid = num_sms() * 8;
    colsum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)d, (float*)out_f32, B, P);
    return check_launch("colsum");
}

extern "C" int vlb_ln2_bwd(const void* dy_f32, const void* keep_scale_f32, const void* pooled, const void* ln2_w,
                           void* dpooled_f32, void* dg2_f32, void* db2_f32, int B, int H, float eps, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8 * HT * HV, "ln2_bwd: H=%d", H);
    ln2_bwd_kernel<<<B, HT, 0, (cudaStream_t)stream>>>((const float*)dy_f32, (const float*)keep_scale_f32,
                                                      (const bf16*)pooled, (const bf16*)ln2_w, (float*)dpooled_f32,
                                                      (float*)dg2_f32, (float*)db2_f32, H, eps);
    return check_launch("ln2_bwd");
}

extern "C" int vlb_hrf_pool_bwd(const void* x, const void* norm_w, const void* ln1_w, const void* wmask_f32,
                                const void* dpooled_f32, void* dx, void* dg1_f32, void* db1_f32, int B, int S, int H,
                                int nch, float eps_rms, float eps_ln, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 8 * HT * HV && nch > 0, "hrf_pool_bwd: H=%d nch=%d", H, nch);
    const int rows_per = ceil_div(S, nch);
    hrf_pool_bwd_kernel<<<dim3(nch, B), HT, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (const bf16*)norm_w, (const bf16*)ln1_w, (const float*)wmask_f32, (const float*)dpooled_f32,
        (bf16*)dx, (float*)dg1_f32, (float*)db1_f32, S, H, rows_per, eps_rms, eps_ln);
    return check_launch("hrf_pool_bwd");
}

extern "C" int vlb_pool_reduce(const void* partial_f32, const void* wmask_f32, void* u_f32, void* sw_f32, int B, int S,
                               int H, int nch, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && B > 0 && nch > 0, "pool_reduce: B=%d H=%d nch=%d", B, H, nch);
    pool_reduce_kernel<<<B, HT, 0, (cudaStream_t)stream>>>((const float*)partial_f32, (const float*)wmask_f32, (float*)u_f32,
                                                          (float*)sw_f32, nch, H, S);
    return check_launch("pool_reduce");
}

extern "C" int vlb_ln1_affine_fwd(const void* u_f32, const void* sw_f32, const void* ln1_w, const void* ln1_b,
                                  void* out_f32, int B, int H, void* stream) {
    const long long n = (long long)B * H;
    ln1_affine_fwd_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(
        (const float*)u_f32, (const float*)sw_f32, (const bf16*)ln1_w, (const bf16*)ln1_b, (float*)out_f32, B, H);
    return check_launch("ln1_affine_fwd");
}

extern "C" int vlb_ln1_affine_bwd(const void* dpooled_f32, const void* u_f32, const void* sw_f32, void* dg1_f32,
                                  void* db1_f32, int B, int H, void* stream) {
    ln1_affine_bwd_kernel<<<ceil_div(H, 128), 128, 0, (cudaStream_t)stream>>>(
        (const float*)dpooled_f32, (const float*)u_f32, (const float*)sw_f32, (float*)dg1_f32, (float*)db1_f32, B, H);
    return check_launch("ln1_affine_bwd");
}

extern "C" int vlb_head_gradnorm(const void* dpred, const void* pred, const void* bias, const void* gram_dpred_f32,
                                 const void* gram_hrf_f32, int ld_gram, const void* wsq_f32, void* sumsq_f32, int N,
                                 long long P, float l2_lambda, void* stream) {
    VLB_REQUIRE(N > 0 && P > 0 && ld_gram >= N, "head_gradnorm: N=%d P=%lld", N, P);
    long long blocks = ceil_div_ll((long long)N * P, 256 * 8);
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    head_gradnorm_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)dpred, (const bf16*)pred, (const bf16*)bias, (const float*)gram_dpred_f32, (const float*)gram_hrf_f32,
        (const float*)wsq_f32, (float*)sumsq_f32, N, ld_gram, P, l2_lambda);
    return check_launch("head_gradnorm");
}

// ------------------------------------------------------------------ pooled head forward with NO allocation anywhere
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
// vlb_hrf_pool_fwd + vlb_hrf_pool_finalize in one call, with the partial[B, nch, H] scratch between them in a file-scope
// device array instead of a caller buffer. The array is part of the library image: every GPU that loads libvlb_b200 holds
// these 2 MiB until the process exits. All calls on a device share it, so two calls that overlap on the GPU (different
// streams, no dependency) race; calls on one stream are ordered and safe. Nothing has to start at zero: the pooling kernel
// overwrites every element of partial[0 : B*nch*H) that the finalize kernel of the same call reads.
constexpr long long POOL_STATIC_MAX_FLOATS = 512 * 1024;           // B * nch * H <= 524 288 floats (2 MiB)
static __device__ __align__(256) float g_pool_static_scratch[POOL_STATIC_MAX_FLOATS];

namespace {
constexpr int POOL_STATIC_MAX_DEVICES = 64;
float* g_pool_static_ptr[POOL_STATIC_MAX_DEVICES];                 // device address of the array, per device ordinal
std::mutex g_pool_static_mu;

// In host code `g_pool_static_scratch` is only a handle of the symbol, not a device address: resolve it once per device,
// with that device current, and hand the kernels the resolved pointer.
int pool_static_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("hrf_pool_fwd_static: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
    if (dev < 0 || dev >= POOL_STATIC_MAX_DEVICES) { set_error("hrf_pool_fwd_static: device ordinal %d out of range", dev); return 1; }
    std::lock_guard<std::mutex> lk(g_pool_static_mu);
    if (!g_pool_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_pool_static_scratch);
        if (e != cudaSuccess || !p) {
            set_error("hrf_pool_fwd_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : 1;
        }
        g_pool_static_ptr[dev] = (float*)p;
    }
    *out = g_pool_static_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int vlb_hrf_pool_static_max_floats(void) { return (int)POOL_STATIC_MAX_FLOATS; }

extern "C" int vlb_hrf_pool_fwd_static(const void* x, const void* norm_w, const void* ln1_w, const void* ln1_b,
                                       const void* wmask_f32, const void* ln2_w, const void* ln2_b, void* pooled,
                                       void* hrf, void* keep_scale_f32, int B, int S, int H, int nch, float eps_rms,
                                       float eps_ln, float drop_p, unsigned long long seed, void* stream) {
    VLB_REQUIRE(B > 0 && S > 0 && H % 8 == 0 && H > 0 && H <= 8 * HT * HV && nch > 0 && nch <= S,
                "hrf_pool_fwd_static: B=%d S=%d H=%d nch=%d", B, S, H, nch);
    const long long need = (long long)B * nch * H;
    VLB_REQUIRE(need <= POOL_STATIC_MAX_FLOATS,
                "hrf_pool_fwd_static: B*nch*H = %lld floats exceeds the built-in scratch of %lld floats (2 MiB); lower nch or "
                "use vlb_hrf_pool_fwd + vlb_hrf_pool_finalize with a caller buffer", need, POOL_STATIC_MAX_FLOATS);
    float* partial = nullptr;
    if (int rc = pool_static_scratch(&partial)) return rc;
    const int rows_per = ceil_div(S, nch);
    hrf_pool_fwd_kernel<<<dim3(nch, B), HT, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (const bf16*)norm_w, (const bf16*)ln1_w, (const bf16*)ln1_b, (const float*)wmask_f32, partial, S, H,
        rows_per, eps_rms, eps_ln);
    if (int rc = check_launch("hrf_pool_fwd_static/pool")) return rc;
    hrf_pool_finalize_kernel<<<B, HT, 0, (cudaStream_t)stream>>>(partial, (const bf16*)ln2_w, (const bf16*)ln2_b,
                                                                (bf16*)pooled, (bf16*)hrf, (float*)keep_scale_f32, nch, H,
                                                                eps_ln, drop_p, seed);
    return check_launch("hrf_pool_fwd_static/finalize");
}
