// tcgen05 / TMEM / TMA GEMM for sm_100a -- the tensor-core core of the decoder, ViT and STC linears.
//
//   C[M,N] = epi( alpha * ( sum_k A1[m,k] B1[n,k]  +  sum_k A2[m,k] B2[n,k] ) )
//   epi(v) = round_bf16( act(v + bias[n]) ) (+ residual[m,n])
//
// The second (A2,B2) operand pair is the fused LoRA update: with A2 = s*(x A^T) [M,R] and B2 = lora_B [N,R]
// the rank-R product is accumulated into the SAME TMEM accumulator as the frozen base weight product, as
// one extra 64-wide k-block after the K/64 base blocks. This replaces peft's three GEMMs + add per linear
// (reference: src/litmodule/videollama2_vlb_litmodule.py:113-120 wraps every decoder linear in lora.Linear).
//
// Structure (one CTA per SM, persistent over output tiles):
//   warp 0      : TMA producer (cp.async.bulk.tensor, 128B swizzle) -> STAGES-deep smem ring, mbarrier full/empty
//   warp 1      : MMA issuer (single thread, tcgen05.mma cta_group::1 kind::f16, M=128, N=BN, K=16)
//   warp 2      : TMEM allocator (2 accumulator stages x BN columns)
//   warps 4..7  : epilogue: tcgen05.ld -> registers -> bias/act/residual -> bf16 -> swizzled smem -> TMA store
// Operands may be K-major (row-major [rows, K]) or MN-major (row-major [K, rows]); the latter lets the
// backward contract over N (dX = dY W) or over M (dW-like products) without materialising transposes.
#include "common.cuh"
#include <mutex>
#include <unordered_map>
#include <string>
#include <string.h>
#include <math.h>
#include <stdlib.h>

namespace vlb {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 256;
constexpr int CD_STAGE_BYTES = BM * 128;          // 128 rows x 64 bf16
constexpr int SMEM_AB_BYTES = 192 * 1024;

template <int BN> struct GemmCfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = SMEM_AB_BYTES / STAGE_BYTES;
    static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * CD_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
    int M, N;
    int nk1, nk2;           // number of 64-wide k-blocks from operand pair 1 and 2
    float alpha;
    const bf16* bias;       // [N] or null
    const bf16* residual;   // [M, ldr] or null
    int ldr;
    int act;
    int mt, nt;             // tile counts
    float* c_f32;           // when non-null: fp32 output [M, ldc32] written straight from registers (gradients)
    int ldc32;
    int accum_f32;          // c_f32 += result instead of =
    int splits;             // split-K factor (fp32 accumulate outputs only): partial tiles are red.add'ed
    // fused rank-K gradient + clip + AdamW epilogue (readout head, config 3): when ad_p != null the accumulator D[m, n] is
    // the gradient of W[n, m] (W row-major [N, ad_ld]); it is consumed in registers and never written to memory.
    float* ad_p; float* ad_m; float* ad_v; bf16* ad_pb; const float* ad_sumsq; long long ad_ld;
    float ad_lr, ad_b1, ad_b2, ad_eps, ad_wd, ad_bc1, ad_bc2, ad_max_norm, ad_gscale, ad_two_lambda;
    int swiglu_I;           // pair kernel, SWIGLU variants: intermediate size I
    const bf16* sw_gu; int sw_ld;   // SWIGLU == 2: saved gate/up activations [M, 2I]
    int stages_used;        // pair kernel: ring depth actually used (<= G2_STAGES), tuning experiments
    int serp;               // pair kernel, long K: odd M-groups sweep the N-blocks backwards and every second tile of a CTA walks K
                            // backwards, so that a wave starts on the panels (and at the K end) the previous wave left in L2
    int group_m;            // M-blocks swept together before moving to the next N-block (L2 residency of the A panel)
    // pair kernel, LORA_EPI variants (dX under LoRA dropout): out[m, n] += lr_scale * sum_a keep_a[m, n] * (t_a[m, :] . A_a[:, n]),
    // t [M, 16 lr_na] bf16, A [16 lr_na, N] bf16, keep bits as 64-bit words [N / 64][M] per adapter (lora_dropout.cu)
    const bf16* lr_t; int lr_ldt; const bf16* lr_A; int lr_lda;
    const unsigned long long* lr_bits[3]; int lr_na; float lr_scale;
    // pair kernel, epilogue mode 3 (qkv projection): RoPE on the heads whose columns lie below rope_ncols, position = row % rope_S
    const bf16* rope_cos; const bf16* rope_sin; int rope_S; int rope_ncols;
};

__device__ __forceinline__ void tile_coords(int t, int mt, int nt, int GROUP_M, int& mb, int& nb, int snake = 0) {
    const int per_group = GROUP_M * nt;
    const int g = t / per_group;
    const int first_m = g * GROUP_M;
    const int gm = min(GROUP_M, mt - first_m);
    const int r = t - g * per_group;
    mb = first_m + (r % gm);
    nb = r / gm;
    if (snake && (g & 1)) nb = nt - 1 - nb;      // boustrophedon over N: the next group starts on the B panels this one ended with
}

// work item t -> (output tile, k-block range) when the contraction is split `splits` ways
__device__ __forceinline__ void split_range(int t, int splits, int nk, int& tile, int& kb0, int& kb1) {
    tile = t / splits;
    const int s = t - tile * splits;
    kb0 = (int)((long long)nk * s / splits);
    kb1 = (int)((long long)nk * (s + 1) / splits);
}

template <int BN, bool A_MN, bool B_MN, bool ADAM = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const GemmParams p,
                    const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                    const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                    const __grid_constant__ CUtensorMap tmC) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
    uint8_t* smem_cd = smem + STAGES * Cfg::STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_cd + 2 * CD_STAGE_BYTES);
    uint64_t* full_bar = bars;                    // [STAGES]
    uint64_t* empty_bar = bars + STAGES;          // [STAGES]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]
    uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nk = p.nk1 + p.nk2;
    const int num_tiles = p.mt * p.nt * p.splits;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA1); tma_prefetch_desc(&tmB1); tma_prefetch_desc(&tmC);
        if (p.nk2 > 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 128); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                int tile, kb0, kb1; split_range(t, p.splits, nk, tile, kb0, kb1);
                int mb, nb; tile_coords(tile, p.mt, p.nt, p.group_m, mb, nb);
                const int m0 = mb * BM, n0 = nb * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                    const bool second = kb >= p.nk1;
                    const CUtensorMap* ta = second ? &tmA2 : &tmA1;
                    const CUtensorMap* tb = second ? &tmB2 : &tmB1;
                    const int k0 = (second ? kb - p.nk1 : kb) * BK;
                    uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
                    uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
                    if constexpr (!A_MN) {
                        tma_load_2d(ta, &full_bar[stage], sa, k0, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            tma_load_2d(ta, &full_bar[stage], sa + j * (BK * 128), m0 + j * 64, k0);
                    }
                    if constexpr (!B_MN) {
                        tma_load_2d(tb, &full_bar[stage], sb, k0, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(tb, &full_bar[stage], sb + j * (BK * 128), n0 + j * 64, k0);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (one thread)
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
            // K-major: SBO = 1024 B (8 rows x 128 B), LBO unused. MN-major: SBO = 1024 B (8 k-rows), LBO = atom stride.
            constexpr uint32_t LBO = BK * 128;
            constexpr uint32_t KSTEP_A = A_MN ? (16 * 128) : 32;   // bytes to advance per UMMA_K=16
            constexpr uint32_t KSTEP_B = B_MN ? (16 * 128) : 32;
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tempty_bar[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                int tile, kb0, kb1; split_range(t, p.splits, nk, tile, kb0, kb1);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * Cfg::A_BYTES);
                    const uint32_t b_addr = smem_u32(smem_b + stage * Cfg::B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t ad = make_smem_desc(a_addr + k * KSTEP_A, A_MN ? LBO : 0, 1024);
                        const uint64_t bd = make_smem_desc(b_addr + k * KSTEP_B, B_MN ? LBO : 0, 1024);
                        umma_bf16(d_tmem, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[stage]);                 // frees this smem stage when the MMAs retire
                    if (kb == kb1 - 1) umma_commit(&tfull_bar[as]);  // accumulator ready for the epilogue
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue (128 threads)
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;
        int it = 0;
        uint32_t cd_idx = 0;
        const bool issuer = (warp == 4 && lane == 0);
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            int mb, nb; tile_coords(t / p.splits, p.mt, p.nt, p.group_m, mb, nb);
            const int m0 = mb * BM, n0 = nb * BN;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(&tfull_bar[as], aphase);
            tc_fence_after();
            const int m = m0 + row;
            const bool m_ok = m < p.M;
            if constexpr (ADAM) {
                // gradient + clip + AdamW: lane = m (input feature h, contiguous in W), register c = n (voxel): for a fixed
                // register the 32 lanes of a warp touch 128 contiguous bytes of W / m / v, so every access is one wavefront
                float clip = 1.f;
                if (p.ad_max_norm > 0.f) clip = fminf(1.f, p.ad_max_norm / (sqrtf(*p.ad_sumsq) * p.ad_gscale + 1e-6f));
                const float gs = p.ad_gscale * clip;
                const float step = p.ad_lr / p.ad_bc1, rbc2 = rsqrtf(p.ad_bc2), decay = 1.f - p.ad_lr * p.ad_wd;
                // software pipeline over 16-voxel slabs: the 48 loads (W, m, v) of slab s+1 are in flight while slab s is
                // updated and stored, so each SM keeps reads and writes streaming instead of alternating bursts
                auto ld_slab = [&](int c0, float (&w)[16], float (&mm)[16], float (&vv)[16]) {
                    const size_t base = (size_t)(n0 + c0) * (size_t)p.ad_ld + (size_t)m;
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const bool ok = m_ok && (n0 + c0 + c < p.N);
                        const size_t idx = base + (size_t)c * (size_t)p.ad_ld;
                        w[c] = ok ? p.ad_p[idx] : 0.f; mm[c] = ok ? p.ad_m[idx] : 0.f; vv[c] = ok ? p.ad_v[idx] : 0.f;
                    }
                };
                auto upd_slab = [&](int c0, const uint32_t* r, float (&w)[16], float (&mm)[16], float (&vv)[16]) {
                    const size_t base = (size_t)(n0 + c0) * (size_t)p.ad_ld + (size_t)m;
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const float gg = (__uint_as_float(r[c]) * p.alpha + p.ad_two_lambda * bf2f(f2bf(w[c]))) * gs;
                        const float m1 = p.ad_b1 * mm[c] + (1.f - p.ad_b1) * gg;
                        const float v1 = p.ad_b2 * vv[c] + (1.f - p.ad_b2) * gg * gg;
                        const float w1 = w[c] * decay - step * m1 / (sqrtf(v1) * rbc2 + p.ad_eps);
                        if (m_ok && (n0 + c0 + c < p.N)) {
                            const size_t idx = base + (size_t)c * (size_t)p.ad_ld;
                            p.ad_p[idx] = w1; p.ad_m[idx] = m1; p.ad_v[idx] = v1; p.ad_pb[idx] = f2bf(w1);
                        }
                    }
                };
                float wa[16], ma[16], va[16], wb[16], mb_[16], vb[16];
                ld_slab(0, wa, ma, va);
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c0), r);
                    tmem_ld_wait();
                    ld_slab(c0 + 16, wb, mb_, vb);
                    upd_slab(c0, r, wa, ma, va);
                    if (c0 + 32 < BN) ld_slab(c0 + 32, wa, ma, va);
                    upd_slab(c0 + 16, r + 16, wb, mb_, vb);
                }
                tc_fence_before();
                mbar_arrive(&tempty_bar[as]);
                continue;
            }
            if (p.c_f32 != nullptr) {
                // fp32 output path (LoRA / head gradients): registers -> global, no bf16 rounding
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c0), r);
                    tmem_ld_wait();
                    if (m_ok) {
                        float* dst = p.c_f32 + (size_t)m * p.ldc32 + n0 + c0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            if (n0 + c0 + c * 4 < p.N) {      // N is a multiple of 8
                                float4 v = make_float4(__uint_as_float(r[c * 4]) * p.alpha, __uint_as_float(r[c * 4 + 1]) * p.alpha,
                                                       __uint_as_float(r[c * 4 + 2]) * p.alpha, __uint_as_float(r[c * 4 + 3]) * p.alpha);
                                if (p.splits > 1) {
                                    atomicAdd(reinterpret_cast<float4*>(dst + c * 4), v);     // red.global.add.v4.f32
                                    continue;
                                }
                                if (p.accum_f32) {
                                    const float4 o = *reinterpret_cast<const float4*>(dst + c * 4);
                                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                                }
                                *reinterpret_cast<float4*>(dst + c * 4) = v;
                            }
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&tempty_bar[as]);
                continue;
            }
#pragma unroll 1
            for (int s = 0; s < BN / 64; ++s, ++cd_idx) {
                uint8_t* cd = smem_cd + (cd_idx & 1) * CD_STAGE_BYTES;
                if (issuer) tma_store_wait_read<1>();       // the store that last used this buffer has been read out
                asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t r[32];
                    const int c0 = s * 64 + h * 32;
                    const int n = n0 + c0;
                    // bias / residual for these 32 columns are requested BEFORE the TMEM wait: issued inside the per-chunk loop
                    // each load sat behind the previous chunk's math and exposed a full memory latency (32 + 32 per tile; the
                    // K = 1024 ViT GEMMs ran at 200-650 TF/s because of it)
                    uint4 bv[4], rv[4];
                    const bool has_bias = p.bias != nullptr, has_res = p.residual != nullptr && m_ok;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int nn = n + c * 8;
                        bv[c] = (has_bias && nn < p.N) ? __ldg(reinterpret_cast<const uint4*>(p.bias + nn)) : make_uint4(0u, 0u, 0u, 0u);
                        rv[c] = (has_res && nn < p.N) ? *reinterpret_cast<const uint4*>(p.residual + (size_t)m * p.ldr + nn)
                                                      : make_uint4(0u, 0u, 0u, 0u);
                    }
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c0), r);
                    tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 4; ++c) {           // 4 chunks of 8 columns (16 B of bf16)
                        float v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[c * 8 + j]) * p.alpha;
                        if (has_bias) {
                            float b[8]; unpack8(bv[c], b);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] += b[j];
                        }
                        if (p.act != ACT_NONE) apply_act8(v, p.act);
                        if (has_res) {
                            float rr[8]; unpack8(rv[c], rr);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] = bf2f(f2bf(v[j])) + rr[j];
                        }
                        uint4 u;
                        u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]);
                        u.z = pack_bf16(v[4], v[5]); u.w = pack_bf16(v[6], v[7]);
                        const int chunk = h * 4 + c;        // 16-byte chunk index inside the 128-byte row
                        sts128(smem_u32(cd) + row * 128 + ((chunk ^ (row & 7)) << 4), u);
                    }
                }
                if (s == BN / 64 - 1) {                     // all TMEM reads of this accumulator stage are done
                    tc_fence_before();
                    mbar_arrive(&tempty_bar[as]);
                }
                fence_proxy_async_smem();
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (issuer) {
                    tma_store_2d(&tmC, cd, n0 + s * 64, m0);
                    tma_store_commit();
                }
            }
        }
        if (issuer) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc<Cfg::TMEM_COLS>(tmem_base); }
}

}  // namespace vlb

#include "gemm2_kernel.cuh"

namespace vlb {

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(ptr);
    });
    return fn;
}

struct TmKey {
    const void* ptr; uint64_t d0, d1, stride; uint32_t b0, b1;
    bool operator==(const TmKey& o) const {
        return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && stride == o.stride && b0 == o.b0 && b1 == o.b1;
    }
};
struct TmHash {
    size_t operator()(const TmKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        h = h * 1000003u ^ k.d0; h = h * 1000003u ^ k.d1; h = h * 1000003u ^ k.stride;
        h = h * 1000003u ^ k.b0; h = h * 1000003u ^ k.b1;
        return h;
    }
};
static std::mutex g_tm_mutex;
static std::unordered_map<TmKey, CUtensorMap, TmHash> g_tm_cache;

// 2-D bf16 tensor map, row-major [d1 rows, d0 cols] with row stride `stride_elems`, box [b1 rows, b0 cols], SW128.
static int make_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride_elems,
                     uint32_t b0, uint32_t b1) {
    TmKey key{ptr, d0, d1, stride_elems, b0, b1};
    {
        std::lock_guard<std::mutex> lk(g_tm_mutex);
        auto it = g_tm_cache.find(key);
        if (it != g_tm_cache.end()) { *out = it->second; return 0; }
    }
    PFN_encodeTiled enc = get_encode();
    VLB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
    VLB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
    VLB_REQUIRE((stride_elems * 2) % 16 == 0, "TMA row stride %llu elements is not a multiple of 16 bytes",
                (unsigned long long)stride_elems);
    cuuint64_t dims[2] = {d0, d1};
    cuuint64_t strides[1] = {stride_elems * 2};
    cuuint32_t box[2] = {b0, b1};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VLB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) ptr=%p dims=%llu,%llu stride=%llu box=%u,%u",
                (int)r, ptr, (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)stride_elems, b0, b1);
    std::lock_guard<std::mutex> lk(g_tm_mutex);
    if (g_tm_cache.size() > 65536) g_tm_cache.clear();
    g_tm_cache.emplace(key, *out);
    return 0;
}

int tmap2d_cached(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride_elems, uint32_t b0, uint32_t b1) {
    return make_tmap(out, ptr, d0, d1, stride_elems, b0, b1);
}

int encode_tmap_bf16(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* dims,
                     const unsigned long long* strides_bytes, const unsigned int* box) {
    PFN_encodeTiled enc = get_encode();
    VLB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
    VLB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
    cuuint64_t d[5]; cuuint64_t st[5]; cuuint32_t bx[5]; cuuint32_t es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) {
        st[i] = strides_bytes[i];
        VLB_REQUIRE(st[i] % 16 == 0, "TMA stride %llu bytes is not a multiple of 16", (unsigned long long)st[i]);
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VLB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(rank %d) failed (%d)", rank, (int)r);
    return 0;
}

// fp32, no swizzle: plain row-major boxes (used by the attention backward's TMA reduce-add of dQ tiles)
int encode_tmap_f32_plain(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* dims,
                          const unsigned long long* strides_bytes, const unsigned int* box) {
    PFN_encodeTiled enc = get_encode();
    VLB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
    VLB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
    cuuint64_t d[5]; cuuint64_t st[5]; cuuint32_t bx[5]; cuuint32_t es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) {
        st[i] = strides_bytes[i];
        VLB_REQUIRE(st[i] % 16 == 0, "TMA stride %llu bytes is not a multiple of 16", (unsigned long long)st[i]);
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VLB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(f32, rank %d) failed (%d)", rank, (int)r);
    return 0;
}

void tmap_cache_clear() {
    std::lock_guard<std::mutex> lk(g_tm_mutex);
    g_tm_cache.clear();
}

struct GemmArgs {
    const void* A1; int lda1; int K1;
    const void* B1; int ldb1;
    const void* A2; int lda2; int K2;
    const void* B2; int ldb2;
    void* C; int ldc;
    int c_is_f32, accum;
    int M, N;
    float alpha;
    const void* bias; const void* residual; int ldr; int act;
    int a_mn, b_mn;     // operand storage: 0 = [rows, K] (K-major), 1 = [K, rows] (MN-major)
    int max_ctas;       // 0 = all SMs
    int split_k;        // 0 = auto (fp32 accumulate outputs only), 1 = off, >1 = forced
    // optional fused AdamW epilogue (see GemmParams)
    float* ad_p = nullptr; float* ad_m = nullptr; float* ad_v = nullptr; void* ad_pb = nullptr; const float* ad_sumsq = nullptr;
    long long ad_ld = 0;
    float ad_lr = 0, ad_b1 = 0, ad_b2 = 0, ad_eps = 0, ad_wd = 0, ad_bc1 = 1, ad_bc2 = 1, ad_max_norm = 0, ad_gscale = 1, ad_two_lambda = 0;
    // optional masked adapter term in the pair kernel's epilogue (LORA_EPI, see GemmParams)
    const void* lr_t = nullptr; int lr_ldt = 0; const void* lr_A = nullptr; int lr_lda = 0;
    const void* lr_bits[3] = {nullptr, nullptr, nullptr}; int lr_na = 0; float lr_scale = 1.f;
    // optional RoPE epilogue (pair kernel, mode 3)
    const void* rope_cos = nullptr; const void* rope_sin = nullptr; int rope_S = 1; int rope_ncols = 0;
};

template <int BN, bool A_MN, bool B_MN, bool ADAM = false>
static int launch_gemm(const GemmArgs& a, cudaStream_t stream) {
    using Cfg = GemmCfg<BN>;
    CUtensorMap tA1, tB1, tA2, tB2, tC;
    int rc;
    if (!A_MN) rc = make_tmap(&tA1, a.A1, a.K1, a.M, a.lda1, BK, BM);
    else       rc = make_tmap(&tA1, a.A1, a.M, a.K1, a.lda1, 64, BK);
    if (rc) return rc;
    if (!B_MN) rc = make_tmap(&tB1, a.B1, a.K1, a.N, a.ldb1, BK, BN);
    else       rc = make_tmap(&tB1, a.B1, a.N, a.K1, a.ldb1, 64, BK);
    if (rc) return rc;
    if (a.K2 > 0) {
        if (!A_MN) rc = make_tmap(&tA2, a.A2, a.K2, a.M, a.lda2, BK, BM);
        else       rc = make_tmap(&tA2, a.A2, a.M, a.K2, a.lda2, 64, BK);
        if (rc) return rc;
        if (!B_MN) rc = make_tmap(&tB2, a.B2, a.K2, a.N, a.ldb2, BK, BN);
        else       rc = make_tmap(&tB2, a.B2, a.N, a.K2, a.ldb2, 64, BK);
        if (rc) return rc;
    } else { tA2 = tA1; tB2 = tB1; }
    if (!a.c_is_f32) {
        rc = make_tmap(&tC, a.C, a.N, a.M, a.ldc, 64, BM);
        if (rc) return rc;
    } else { tC = tA1; }

    GemmParams p;
    p.M = a.M; p.N = a.N;
    p.nk1 = ceil_div(a.K1, BK); p.nk2 = a.K2 > 0 ? ceil_div(a.K2, BK) : 0;
    p.alpha = a.alpha;
    p.bias = reinterpret_cast<const bf16*>(a.bias);
    p.residual = reinterpret_cast<const bf16*>(a.residual);
    p.ldr = a.ldr; p.act = a.act;
    p.mt = ceil_div(a.M, BM); p.nt = ceil_div(a.N, BN);
    p.c_f32 = a.c_is_f32 ? reinterpret_cast<float*>(a.C) : nullptr;
    p.ldc32 = a.ldc; p.accum_f32 = a.accum;
    p.ad_p = a.ad_p; p.ad_m = a.ad_m; p.ad_v = a.ad_v; p.ad_pb = reinterpret_cast<bf16*>(a.ad_pb); p.ad_sumsq = a.ad_sumsq;
    p.ad_ld = a.ad_ld; p.ad_lr = a.ad_lr; p.ad_b1 = a.ad_b1; p.ad_b2 = a.ad_b2; p.ad_eps = a.ad_eps; p.ad_wd = a.ad_wd;
    p.ad_bc1 = a.ad_bc1; p.ad_bc2 = a.ad_bc2; p.ad_max_norm = a.ad_max_norm; p.ad_gscale = a.ad_gscale;
    p.ad_two_lambda = a.ad_two_lambda;
    p.stages_used = 0; p.serp = 0; p.swiglu_I = 0; p.sw_gu = nullptr; p.sw_ld = 0;
    p.group_m = 8;
    { const char* g = getenv("VLB_GEMM_GROUP_M"); if (g && atoi(g) > 0) p.group_m = atoi(g); }
    // split-K: gradient GEMMs contract over all tokens into a tiny [r, N] output; without it only mt*nt (< 148) CTAs
    // stream the activations. Partial tiles are red.add'ed, so the output must already hold the value to add to.
    p.splits = 1;
    if (a.c_is_f32 && a.accum && a.split_k != 1) {
        const int nk = p.nk1 + p.nk2;
        int sp = a.split_k > 1 ? a.split_k : num_sms() / (p.mt * p.nt);
        if (sp > nk / 4) sp = nk / 4;          // keep >= 4 k-blocks (256 contraction rows) per partial tile
        if (sp > 1) p.splits = sp;
    }
    VLB_REQUIRE(a.split_k <= 1 || (a.c_is_f32 && a.accum), "gemm: split_k=%d needs c_is_f32=1 and accumulate=1", a.split_k);

    VLB_REQUIRE(ADAM == (a.ad_p != nullptr), "gemm: AdamW epilogue arguments do not match the kernel variant");
    auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, ADAM>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        VLB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(gemm smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
        attr_set = true;
    }
    int grid = p.mt * p.nt * p.splits;
    int cap = a.max_ctas > 0 ? a.max_ctas : num_sms();
    if (grid > cap) grid = cap;
    kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(p, tA1, tB1, tA2, tB2, tC);
    return check_launch("gemm_tcgen05_kernel");
}

// CTA-pair launch: 256 x 256 tiles, bf16 output, K-major A (the decoder / ViT / STC linears and their dX GEMMs).
template <bool B_MN, int SWIGLU = 0, bool LORA = false>
static int launch_gemm2(const GemmArgs& a, cudaStream_t stream, void* act = nullptr, int ld_act = 0) {
    CUtensorMap tA1, tB1, tA2, tB2, tC, tAct;
    int rc;
    if ((rc = make_tmap(&tA1, a.A1, a.K1, a.M, a.lda1, BK, BM))) return rc;
    if (!B_MN) rc = make_tmap(&tB1, a.B1, a.K1, a.N, a.ldb1, BK, G2_HALF_N);
    else       rc = make_tmap(&tB1, a.B1, a.N, a.K1, a.ldb1, 64, BK);
    if (rc) return rc;
    if (a.K2 > 0) {
        if ((rc = make_tmap(&tA2, a.A2, a.K2, a.M, a.lda2, BK, BM))) return rc;
        if (!B_MN) rc = make_tmap(&tB2, a.B2, a.K2, a.N, a.ldb2, BK, G2_HALF_N);
        else       rc = make_tmap(&tB2, a.B2, a.N, a.K2, a.ldb2, 64, BK);
        if (rc) return rc;
    } else { tA2 = tA1; tB2 = tB1; }
    if ((rc = make_tmap(&tC, a.C, SWIGLU == 2 ? 2 * a.N : a.N, a.M, a.ldc, 64, BM))) return rc;
    if (SWIGLU == 1) { if ((rc = make_tmap(&tAct, act, a.N / 2, a.M, ld_act, 64, BM))) return rc; }
    else tAct = tC;
    GemmParams p;
    p.M = a.M; p.N = a.N;
    p.nk1 = ceil_div(a.K1, BK); p.nk2 = a.K2 > 0 ? ceil_div(a.K2, BK) : 0;
    p.alpha = a.alpha;
    p.bias = reinterpret_cast<const bf16*>(a.bias);
    p.residual = reinterpret_cast<const bf16*>(a.residual);
    p.ldr = a.ldr; p.act = a.act;
    p.swiglu_I = SWIGLU == 1 ? a.N / 2 : SWIGLU == 2 ? a.N : 0;
    p.sw_gu = reinterpret_cast<const bf16*>(act); p.sw_ld = ld_act;      // SWIGLU == 2: `act` carries gu
    p.mt = ceil_div(a.M, 2 * BM); p.nt = SWIGLU == 1 ? (a.N / 2) / G2_HALF_N : ceil_div(a.N, G2_BN);
    p.c_f32 = nullptr; p.ldc32 = 0; p.accum_f32 = 0; p.splits = 1;
    p.ad_p = nullptr; p.ad_m = nullptr; p.ad_v = nullptr; p.ad_pb = nullptr; p.ad_sumsq = nullptr; p.ad_ld = 0;
    // M-blocks (256 rows) swept together before moving to the next N-block: the A panel (group x 256 x K) stays in L2
    // while B streams past it. Measured sustained at M=16384: K=4096 shapes +1.5..3 % at 16, K >= 14336 shapes best at 8.
    p.group_m = a.K1 <= 8192 ? 16 : 8;       // 32 for the 28672-wide gate/up projection looked 2 % better standalone and was 1 % worse in the step
    { const char* g = getenv("VLB_GEMM_GROUP_M"); if (g && atoi(g) > 0) p.group_m = atoi(g); }
    // ring depth: 6 keeps the tensor pipe 95-99 % busy, but the deeper the ring the further the CTAs of a wave drift apart along K
    // and the less they share each other's A / B panels in L2. ncu at M = 16384 (profiles/r02r_ring_depth.md): K = 14336 reads
    // 3.49 GB per launch with 6 stages and 2.23 GB with 4 (and is 3 % faster), K = 28672 10.0 GB vs 4.47 GB (11 % faster); a
    // 3-deep ring moves the same bytes as 4 but starves the tensor pipe (71 % active). 2.23 / 4.47 GB is 3.1-3.4x the
    // algorithmic bytes: a wave of 74 256x256 tiles touches ~17 distinct 256-row panels of K columns each, and the panels of
    // consecutive waves (8 x 7-15 MB of A, all of B) no longer fit the 126 MB L2 at these K.
    // with the serpentine order (below) K = 14336 is 4 % faster again with 5 stages than with 4 (profiles/r02ba_stages_serp.log)
    p.stages_used = a.K1 >= 20480 ? 4 : a.K1 >= 8192 ? 5 : G2_STAGES;
    p.serp = (a.K1 >= 8192 && !getenv("VLB_GEMM_NO_SERP")) ? 1 : 0;
    if (SWIGLU == 1 || SWIGLU == 3) p.stages_used = G2_STAGES - 1;      // the last ring stage is epilogue staging (third buffer / second warpgroup's pair)
    p.rope_cos = reinterpret_cast<const bf16*>(a.rope_cos); p.rope_sin = reinterpret_cast<const bf16*>(a.rope_sin); p.rope_S = a.rope_S; p.rope_ncols = a.rope_ncols;
    p.lr_t = reinterpret_cast<const bf16*>(a.lr_t); p.lr_ldt = a.lr_ldt; p.lr_A = reinterpret_cast<const bf16*>(a.lr_A); p.lr_lda = a.lr_lda;
    for (int i = 0; i < 3; ++i) p.lr_bits[i] = reinterpret_cast<const unsigned long long*>(a.lr_bits[i]);
    p.lr_na = a.lr_na; p.lr_scale = a.lr_scale;
    if (LORA && p.stages_used > G2_STAGES - 1) p.stages_used = G2_STAGES - 1;   // the last ring stage holds the staged adapter columns / the second warpgroup's output staging
    { const char* g = getenv("VLB_GEMM_STAGES"); if (g && atoi(g) > 0 && atoi(g) <= G2_STAGES - ((LORA || SWIGLU == 1 || SWIGLU == 3) ? 1 : 0)) p.stages_used = atoi(g); }
    auto kern = gemm2_tcgen05_kernel<B_MN, SWIGLU, LORA>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G2_SMEM_BYTES);
        VLB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(gemm2 smem=%d): %s", G2_SMEM_BYTES, cudaGetErrorString(e));
        attr_set = true;
    }
    int pairs = p.mt * p.nt;
    int cap = (a.max_ctas > 0 ? a.max_ctas : num_sms()) / 2;
    if (pairs > cap) pairs = cap;
    if (pairs < 1) pairs = 1;
    kern<<<2 * pairs, g2_threads<SWIGLU, LORA>(), G2_SMEM_BYTES, stream>>>(p, tA1, tB1, tA2, tB2, tC, tAct);
    return check_launch("gemm2_tcgen05_kernel");
}

static bool use_pair_kernel(const GemmArgs& a) {
    static int mode = -1;                           // VLB_GEMM_2CTA=0 forces the single-CTA kernel (A/B comparisons)
    if (mode < 0) { const char* e = getenv("VLB_GEMM_2CTA"); mode = (e && atoi(e) == 0) ? 0 : 1; }
    return mode == 1 && !a.a_mn && !a.c_is_f32 && a.M >= 2 * BM && a.N >= G2_BN &&
           (long long)ceil_div(a.M, 2 * BM) * ceil_div(a.N, G2_BN) >= num_sms() / 2;
}

template <int BN>
static int dispatch_major(const GemmArgs& a, cudaStream_t s) {
    if (!a.a_mn && !a.b_mn) return launch_gemm<BN, false, false>(a, s);
    if (!a.a_mn && a.b_mn)  return launch_gemm<BN, false, true>(a, s);
    if (a.a_mn && !a.b_mn)  return launch_gemm<BN, true, false>(a, s);
    return launch_gemm<BN, true, true>(a, s);
}

int gemm_dispatch(const GemmArgs& a, int block_n, cudaStream_t s) {
    VLB_REQUIRE(a.M > 0 && a.N > 0 && a.K1 > 0, "gemm: bad shape M=%d N=%d K=%d", a.M, a.N, a.K1);
    VLB_REQUIRE(a.N % 8 == 0, "gemm: N=%d must be a multiple of 8", a.N);
    int bn = block_n;
    if (bn == 0) bn = (a.N >= 256 && a.N % 256 == 0) ? 256 : (a.N > 64 ? 128 : 64);
    if (bn == 256) {
        // prefer 128-wide tiles when 256-wide ones cannot fill the machine
        long long tiles = (long long)ceil_div(a.M, BM) * ceil_div(a.N, 256);
        // ... except the split-K rank-r gradient GEMMs over wide outputs: split-K fills the machine anyway and 512 contiguous
        // bytes per token row use DRAM better (N=14336: 105 -> 91 us, tools/bench_lora_grad.py)
        const bool wide_splitk = a.c_is_f32 && a.accum && a.split_k != 1 && a.N >= 8192;
        if (block_n == 0 && tiles < num_sms() && !wide_splitk) bn = 128;
    }
    if (a.ad_p != nullptr) return launch_gemm<256, true, true, true>(a, s);      // fused head optimiser (vlb_head_adamw_fused)
    if (bn == 256 && use_pair_kernel(a)) return a.b_mn ? launch_gemm2<true>(a, s) : launch_gemm2<false>(a, s);
    switch (bn) {
        case 256: return dispatch_major<256>(a, s);
        case 128: return dispatch_major<128>(a, s);
        case 64:  return dispatch_major<64>(a, s);
        default: VLB_REQUIRE(false, "gemm: unsupported block_n=%d", bn);
    }
    return 1;
}

}  // namespace vlb

extern "C" int vlb_gemm_bf16(const void* A, int lda, const void* B, int ldb, void* C, int ldc,
                             int M, int N, int K,
                             const void* A2, int lda2, const void* B2, int ldb2, int K2,
                             float alpha, const void* bias, const void* residual, int ldr, int act,
                             int a_mn_major, int b_mn_major, int c_is_f32, int accumulate,
                             int block_n, int max_ctas, int split_k, void* stream) {
    vlb::GemmArgs a;
    a.A1 = A; a.lda1 = lda; a.K1 = K; a.B1 = B; a.ldb1 = ldb;
    a.A2 = A2; a.lda2 = lda2; a.K2 = (A2 && B2) ? K2 : 0; a.B2 = B2; a.ldb2 = ldb2;
    a.C = C; a.ldc = ldc; a.c_is_f32 = c_is_f32; a.accum = accumulate; a.M = M; a.N = N; a.alpha = alpha;
    a.bias = bias; a.residual = residual; a.ldr = ldr; a.act = act;
    a.a_mn = a_mn_major; a.b_mn = b_mn_major; a.max_ctas = max_ctas; a.split_k = split_k;
    return vlb::gemm_dispatch(a, block_n, reinterpret_cast<cudaStream_t>(stream));
}

// Readout-head optimiser step without a materialised gradient (BASELINE configs[3], P ~ 250k voxels):
//   dW[p, h] = sum_n dpred[n, p] * hrf[n, h] + 2*lambda*W[p, h]   (rank-N, N = samples in the step)
// is formed tile by tile on the tensor cores (M = H on the TMEM lanes, N = voxels, K = samples) and consumed by the
// clip + AdamW epilogue in registers: 26 B/parameter of HBM traffic instead of 42 (no 4 GB fp32 dW write + re-read).
extern "C" int vlb_head_adamw_fused(const void* dpred, int ld_dpred, const void* hrf, int ld_hrf, int n_samples,
                                    long long P, int H, void* w_f32, void* m_f32, void* v_f32, void* w_bf16,
                                    const void* sumsq_f32, float lr, float beta1, float beta2, float eps,
                                    float weight_decay, int step, float max_norm, float grad_scale, float l2_lambda,
                                    void* stream) {
    VLB_REQUIRE(n_samples > 0 && P > 0 && P < (1ll << 31) && H % 8 == 0 && step >= 1, "head_adamw_fused: n=%d P=%lld H=%d step=%d",
                n_samples, P, H, step);
    vlb::GemmArgs a;
    a.A1 = hrf; a.lda1 = ld_hrf; a.K1 = n_samples; a.B1 = dpred; a.ldb1 = ld_dpred;      // both stored [samples, features]
    a.A2 = nullptr; a.lda2 = 0; a.K2 = 0; a.B2 = nullptr; a.ldb2 = 0;
    a.C = nullptr; a.ldc = 0; a.c_is_f32 = 1; a.accum = 0; a.M = H; a.N = (int)P; a.alpha = 1.f;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 1; a.b_mn = 1; a.max_ctas = 0; a.split_k = 1;
    a.ad_p = (float*)w_f32; a.ad_m = (float*)m_f32; a.ad_v = (float*)v_f32; a.ad_pb = w_bf16; a.ad_sumsq = (const float*)sumsq_f32;
    a.ad_ld = H;
    a.ad_lr = lr; a.ad_b1 = beta1; a.ad_b2 = beta2; a.ad_eps = eps; a.ad_wd = weight_decay;
    a.ad_bc1 = 1.f - powf(beta1, (float)step); a.ad_bc2 = 1.f - powf(beta2, (float)step);
    a.ad_max_norm = max_norm; a.ad_gscale = grad_scale; a.ad_two_lambda = 2.f * l2_lambda;
    return vlb::gemm_dispatch(a, 256, reinterpret_cast<cudaStream_t>(stream));
}

// q/k/v projection of a decoder layer with RoPE fused into the epilogue (pair kernel, head_dim 128): the two warpgroups of a CTA
// own one head each (128 of the tile's 256 columns), read both halves of the head from TMEM and write the rotated pair, with the
// rounding points of the stand-alone kernel (bf16 projection, bf16 products, bf16 sum). Shapes the pair kernel does not take
// (narrow test models) run the plain GEMM followed by vlb_rope.
extern "C" int vlb_rope(void* qkv, int ld, const void* cos_tab, const void* sin_tab, long long M, int S, int nheads, int D,
                        int inverse, void* stream);
extern "C" int vlb_gemm_qkv_rope(const void* x, int ldx, const void* wqkv, int ldw, void* qkv, int ldqkv, int M, int N, int K,
                                 const void* A2, int lda2, const void* B2, int ldb2, int K2, float alpha,
                                 const void* cos_tab, const void* sin_tab, int S, int n_rot_heads, int D, void* stream) {
    VLB_REQUIRE(M > 0 && N > 0 && K > 0 && S > 0 && D > 0 && n_rot_heads * D <= N, "gemm_qkv_rope: M=%d N=%d K=%d S=%d heads=%d D=%d",
                M, N, K, S, n_rot_heads, D);
    vlb::GemmArgs a;
    a.A1 = x; a.lda1 = ldx; a.K1 = K; a.B1 = wqkv; a.ldb1 = ldw;
    a.A2 = A2; a.lda2 = lda2; a.K2 = (A2 && B2) ? K2 : 0; a.B2 = B2; a.ldb2 = ldb2;
    a.C = qkv; a.ldc = ldqkv; a.c_is_f32 = 0; a.accum = 0; a.M = M; a.N = N; a.alpha = alpha;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 0; a.b_mn = 0; a.max_ctas = 0; a.split_k = 1;
    const bool fuse = D == 128 && N % vlb::G2_BN == 0 && vlb::use_pair_kernel(a) && !getenv("VLB_ROPE_UNFUSED");
    if (!fuse) {
        if (int rc = vlb::gemm_dispatch(a, 0, reinterpret_cast<cudaStream_t>(stream))) return rc;
        return vlb_rope(qkv, ldqkv, cos_tab, sin_tab, M, S, n_rot_heads, D, 0, stream);
    }
    a.rope_cos = cos_tab; a.rope_sin = sin_tab; a.rope_S = S; a.rope_ncols = n_rot_heads * D;
    return vlb::launch_gemm2<false, 3>(a, reinterpret_cast<cudaStream_t>(stream));
}

// Forward gate/up projection with SwiGLU fused into the epilogue (pair kernel): gu[M, 2I] = x Wgu^T (+ LoRA k-block) and
// act[M, I] = bf16(silu(gate)) * up in one launch. Wgu is [2I, K] (gate rows then up rows), I % 128 == 0.
extern "C" int vlb_gemm_gateup_swiglu(const void* x, int ldx, const void* wgu, int ldw, void* gu, int ldgu, void* act, int ldact,
                                      int M, int I, int K, const void* A2, int lda2, const void* B2, int ldb2, int K2,
                                      float alpha, void* stream) {
    VLB_REQUIRE(M > 0 && I > 0 && I % 128 == 0 && K > 0, "gemm_gateup_swiglu: M=%d I=%d K=%d (I must be a multiple of 128)", M, I, K);
    vlb::GemmArgs a;
    a.A1 = x; a.lda1 = ldx; a.K1 = K; a.B1 = wgu; a.ldb1 = ldw;
    a.A2 = A2; a.lda2 = lda2; a.K2 = (A2 && B2) ? K2 : 0; a.B2 = B2; a.ldb2 = ldb2;
    a.C = gu; a.ldc = ldgu; a.c_is_f32 = 0; a.accum = 0; a.M = M; a.N = 2 * I; a.alpha = alpha;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 0; a.b_mn = 0; a.max_ctas = 0; a.split_k = 1;
    return vlb::launch_gemm2<false, 1>(a, reinterpret_cast<cudaStream_t>(stream), act, ldact);
}

// Backward of the down projection with the SwiGLU backward in the epilogue (pair kernel): d(act) = dy Wd (+ LoRA k-block)
// stays in TMEM; dgu[M, 2I] = [d(act) up silu'(gate) | d(act) silu(gate)] is written directly. wd is [H, I] (the forward
// weight, read MN-major), gu [M, 2I] the saved forward activations, I % 256 == 0.
extern "C" int vlb_gemm_down_dx_swiglu_bwd(const void* dy, int lddy, const void* wd, int ldw, const void* gu, int ldgu,
                                           void* dgu, int lddgu, int M, int I, int H, const void* A2, int lda2, const void* B2,
                                           int ldb2, int K2, float alpha, void* stream) {
    VLB_REQUIRE(M > 0 && I > 0 && I % 256 == 0 && H > 0 && ldgu % 8 == 0, "gemm_down_dx_swiglu_bwd: M=%d I=%d H=%d ldgu=%d", M, I, H, ldgu);
    vlb::GemmArgs a;
    a.A1 = dy; a.lda1 = lddy; a.K1 = H; a.B1 = wd; a.ldb1 = ldw;
    a.A2 = A2; a.lda2 = lda2; a.K2 = (A2 && B2) ? K2 : 0; a.B2 = B2; a.ldb2 = ldb2;
    a.C = dgu; a.ldc = lddgu; a.c_is_f32 = 0; a.accum = 0; a.M = M; a.N = I; a.alpha = alpha;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 0; a.b_mn = 1; a.max_ctas = 0; a.split_k = 1;
    return vlb::launch_gemm2<true, 2>(a, reinterpret_cast<cudaStream_t>(stream), const_cast<void*>(gu), ldgu);
}

// dX GEMMs under LoRA dropout (pair kernel, LORA_EPI): out[M, N] = dy[M, K] W[K, N] + scale * sum_a keep_a . (t_a A_a), with
// W the forward weight [out_features = K, in_features = N] read MN-major, t [M, 16 n_adapters] = s dy B (one 16-column group per
// adapter), A [16 n_adapters, N] the stacked lora_A matrices and keep_a the adapter's dropout bits over its input [M, N].
static int lora_args(vlb::GemmArgs& a, int M, int N, const void* t, int ldt, const void* A, int lda, int n_adapters,
                     const void* b0, const void* b1, const void* b2, float scale) {
    VLB_REQUIRE(n_adapters >= 1 && n_adapters <= 3 && t && A && b0 && (n_adapters < 2 || b1) && (n_adapters < 3 || b2),
                "gemm lora epilogue: adapters=%d", n_adapters);
    VLB_REQUIRE(N % 64 == 0 && ldt % 8 == 0 && lda % 8 == 0 && M > 0, "gemm lora epilogue: M=%d N=%d ldt=%d lda=%d", M, N, ldt, lda);
    a.lr_t = t; a.lr_ldt = ldt; a.lr_A = A; a.lr_lda = lda; a.lr_bits[0] = b0; a.lr_bits[1] = b1; a.lr_bits[2] = b2;
    a.lr_na = n_adapters; a.lr_scale = scale;
    return 0;
}

extern "C" int vlb_gemm_dx_lora_masked(const void* dy, int lddy, const void* W, int ldw, void* out, int ldo, int M, int N, int K,
                                       const void* t, int ldt, const void* A, int lda, int n_adapters, const void* bits0,
                                       const void* bits1, const void* bits2, float scale, void* stream) {
    VLB_REQUIRE(M > 0 && N > 0 && K > 0 && N % 8 == 0, "gemm_dx_lora_masked: M=%d N=%d K=%d", M, N, K);
    vlb::GemmArgs a;
    a.A1 = dy; a.lda1 = lddy; a.K1 = K; a.B1 = W; a.ldb1 = ldw;
    a.A2 = nullptr; a.lda2 = 0; a.K2 = 0; a.B2 = nullptr; a.ldb2 = 0;
    a.C = out; a.ldc = ldo; a.c_is_f32 = 0; a.accum = 0; a.M = M; a.N = N; a.alpha = 1.f;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 0; a.b_mn = 1; a.max_ctas = 0; a.split_k = 1;
    if (lora_args(a, M, N, t, ldt, A, lda, n_adapters, bits0, bits1, bits2, scale)) return 1;
    return vlb::launch_gemm2<true, 0, true>(a, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vlb_gemm_down_dx_swiglu_bwd_masked(const void* dy, int lddy, const void* wd, int ldw, const void* gu, int ldgu,
                                                  void* dgu, int lddgu, int M, int I, int H, const void* t, int ldt, const void* A,
                                                  int lda, const void* bits, float scale, void* stream) {
    VLB_REQUIRE(M > 0 && I > 0 && I % 256 == 0 && H > 0 && ldgu % 8 == 0, "gemm_down_dx_swiglu_bwd_masked: M=%d I=%d H=%d ldgu=%d", M, I, H, ldgu);
    vlb::GemmArgs a;
    a.A1 = dy; a.lda1 = lddy; a.K1 = H; a.B1 = wd; a.ldb1 = ldw;
    a.A2 = nullptr; a.lda2 = 0; a.K2 = 0; a.B2 = nullptr; a.ldb2 = 0;
    a.C = dgu; a.ldc = lddgu; a.c_is_f32 = 0; a.accum = 0; a.M = M; a.N = I; a.alpha = 1.f;
    a.bias = nullptr; a.residual = nullptr; a.ldr = 0; a.act = 0;
    a.a_mn = 0; a.b_mn = 1; a.max_ctas = 0; a.split_k = 1;
    if (lora_args(a, M, I, t, ldt, A, lda, 1, bits, nullptr, nullptr, scale)) return 1;
    return vlb::launch_gemm2<true, 2, true>(a, reinterpret_cast<cudaStream_t>(stream), const_cast<void*>(gu), ldgu);
}

extern "C" void vlb_tmap_cache_clear() { vlb::tmap_cache_clear(); }
