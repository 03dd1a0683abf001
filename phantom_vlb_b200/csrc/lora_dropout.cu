// LoRA input dropout on the fused path (peft lora.Linear: y = W x + s B(A(dropout(x))), reference lora_dropout = 0.1,
// config/experiment/VLB_vllama2_friends_lora.yaml:30; src/litmodule/videollama2_vlb_litmodule.py:113-120).
//
// Every adapter (q, k, v, o, gate, up, down) draws its OWN Bernoulli keep mask over its input x [tokens, features], exactly as
// peft does, with the exact drop probability (32-bit uniform against p * 2^32, compared bit-serially). The masks are generated once per step as
// packed bits -- 1 bit per element, 1/16 of the bf16 activation they cover -- and are applied where the masked value is
// consumed, so no masked copy of an activation ever goes to memory:
//   forward   xa_a = s/(1-p) * (m_a . x) A_a^T      lora_masked_kernel<false>: the x tile lands in shared memory by TMA, eight
//                                                    transform warps build one masked copy per adapter in place / beside it and
//                                                    the tcgen05 MMAs (N = 16 per adapter) read those; x is streamed ONCE for
//                                                    the q/k/v (gate/up) adapters that share it
//   backward  dA_a += 1/(1-p) * t_a^T (m_a . x)      lora_masked_kernel<true>: same transform on the MN-major x tile (features on
//                                                    the MMA M axis, tokens contracted, split over CTAs), fp32 red.add
//             dx   += 1/(1-p) * m_a . (t_a A_a)      in the epilogue of the dX pair GEMM (gemm2_kernel.cuh, LORA_EPI)
// Bit layout: word64[fb * tokens + tok] holds features [64 fb, 64 fb + 64) of token `tok` (bit i <-> feature 64 fb + i): a
// 128-row x 128-byte shared-memory tile needs one coalesced 8-byte word per row, whichever operand major-ness it has.
#include "common.cuh"
#include <stdlib.h>

namespace vlb {

int tmap2d_cached(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride_elems, uint32_t b0, uint32_t b1);

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// 32-bit counter hash (two multiply-xorshift rounds) with the two halves of the step key folded in between the rounds
__device__ __forceinline__ unsigned int hash32(unsigned int x, unsigned int k1, unsigned int k2) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= k1;
    x ^= x >> 15; x *= 0x846ca68bU; x ^= k2;
    return x ^ (x >> 16);
}

// One thread per 64-bit word. keep_i = (u_i >= thresh) for 64 independent 32-bit uniforms u_i. The top DB_PLANES bits are
// evaluated bit-serially for all 64 lanes at once (random word number b holds bit (31 - b) of all 64 uniforms; a lane is
// decided as soon as its bit differs from thresh's), which leaves 64 / 2^DB_PLANES = 1 lane per word undecided on average;
// those draw their remaining 32 - DB_PLANES bits from one more hash each. The drop probability is exactly thresh / 2^32.
// (Running the planes until the whole warp is decided costs the MAXIMUM over 2048 lanes, ~12 planes; this costs 6 + a tail.)
constexpr int DB_PLANES = 6;
constexpr int DB_THREADS = 128;
__global__ void __launch_bounds__(DB_THREADS) dropout_bits_kernel(unsigned long long* __restrict__ bits, long long nwords,
                                                                  unsigned int thresh, unsigned long long seed) {
    const long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    const unsigned long long key = mix64(seed ^ 0xD1B54A32D192ED03ull);
    const unsigned int k1 = (unsigned int)key ^ (unsigned int)((unsigned long long)w >> 26) * 0x9E3779B1u, k2 = (unsigned int)(key >> 32);
    const unsigned int ctr = (unsigned int)w << 6;                 // 64 counters per word
    unsigned int keep0 = 0, keep1 = 0, und0 = 0xFFFFFFFFu, und1 = 0xFFFFFFFFu;
#pragma unroll
    for (int b = 0; b < DB_PLANES; ++b) {
        const unsigned int r0 = hash32(ctr + 2 * b, k1, k2), r1 = hash32(ctr + 2 * b + 1, k1, k2);
        if ((thresh >> (31 - b)) & 1u) { und0 &= r0; und1 &= r1; }                      // uniform bit 0 under a 1: below thresh, dropped
        else { keep0 |= und0 & r0; keep1 |= und1 & r1; und0 &= ~r0; und1 &= ~r1; }      // uniform bit 1 over a 0: above thresh, kept
    }
    // lanes whose top DB_PLANES bits equal thresh's: compare the remaining bits, drawn per lane from a second key stream
    const unsigned int tl = thresh & (0xFFFFFFFFu >> DB_PLANES), k3 = k1 ^ 0xA511E9B3u;
    while (und0) {
        const int i = __ffs(und0) - 1;
        und0 &= und0 - 1;
        if ((hash32(ctr + i, k3, k2) >> DB_PLANES) >= tl) keep0 |= 1u << i;
    }
    while (und1) {
        const int i = __ffs(und1) - 1;
        und1 &= und1 - 1;
        if ((hash32(ctr + 32 + i, k3, k2) >> DB_PLANES) >= tl) keep1 |= 1u << i;
    }
    bits[w] = (unsigned long long)keep0 | ((unsigned long long)keep1 << 32);
}

constexpr int LM_BM = 128;            // MMA M: tokens (forward) / features (backward)
constexpr int LM_BK = 64;
constexpr int LM_BN = 64;             // B tile rows: 16 per adapter, <= 48 valid, the rest zero-filled by TMA
constexpr int LM_A_BYTES = LM_BM * LM_BK * 2;     // 16 KB
constexpr int LM_B_BYTES = LM_BN * LM_BK * 2;     // 8 KB
constexpr int LM_W_BYTES = LM_BM * 8;             // 1 KB: one 64-bit keep word per tile row and adapter
constexpr int LM_THREADS = 512;       // warp 0 TMA, 1 MMA, 2 TMEM, 3 idle, 4-7 epilogue, 8-15 transform
constexpr int LM_XFORM_THREADS = 256;
#ifndef LM_GROUPS_N
#define LM_GROUPS_N 2
#endif
constexpr int LM_GROUPS = LM_GROUPS_N;  // transform warp groups taking k-blocks in turn (power of two)
constexpr int LM_MAX_STAGES = 8;
constexpr int LM_SMEM_BUDGET = 224 * 1024;

struct LoraMaskParams {
    int n_tok;                         // rows of the bit arrays (tokens)
    int nk, mt, splits, stages, n_ad;
    float alpha;
    const unsigned long long* bits[3];
    bf16* out_bf16;                    // forward: xa [tokens, ld_out]
    float* out_f32;                    // backward: dA [16 n_ad, ld_out], accumulated with red.add
    int ld_out;
    int n_rows;                        // valid MMA-M rows: tokens (forward) / features (backward)
    int n_fb;                          // 64-feature blocks per token in the bit arrays
    int bulk;                          // keep words travel in the TMA ring (full tiles only); 0: per-thread global loads
};

// keep-mask a 16-byte chunk (8 bf16) with 8 bits
__device__ __forceinline__ uint4 mask8(const uint4& v, unsigned int byte) {
    uint4 o;
    o.x = v.x & ((((byte     ) & 1u) | (((byte >> 1) & 1u) << 16)) * 0xFFFFu);
    o.y = v.y & ((((byte >> 2) & 1u) | (((byte >> 3) & 1u) << 16)) * 0xFFFFu);
    o.z = v.z & ((((byte >> 4) & 1u) | (((byte >> 5) & 1u) << 16)) * 0xFFFFu);
    o.w = v.w & ((((byte >> 6) & 1u) | (((byte >> 7) & 1u) << 16)) * 0xFFFFu);
    return o;
}

template <bool A_MN>
__global__ void __launch_bounds__(LM_THREADS, 1)
lora_masked_kernel(const LoraMaskParams p, const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    // stage: n_ad masked copies of the x tile (the raw tile lands in copy 0) | B tile | (bulk) 1 KB of keep words per adapter
    const int words_off = p.n_ad * LM_A_BYTES + LM_B_BYTES;
    const int stage_bytes = words_off + (p.bulk ? p.n_ad * LM_W_BYTES : 0);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
    uint64_t* full_bar = bars;                              // [stages] TMA landed
    uint64_t* ready_bar = bars + LM_MAX_STAGES;             // [stages] masked copies written
    uint64_t* empty_bar = bars + 2 * LM_MAX_STAGES;         // [stages] MMAs retired
    uint64_t* tfull_bar = bars + 3 * LM_MAX_STAGES;         // [2]
    uint64_t* tempty_bar = bars + 3 * LM_MAX_STAGES + 2;    // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 3 * LM_MAX_STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_items = p.mt * p.splits;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < p.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&ready_bar[i], LM_XFORM_THREADS / 32 / LM_GROUPS); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 128); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<2 * LM_BN>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    auto item_range = [&](int t, int& tile, int& kb0, int& kb1) {
        tile = t / p.splits;
        const int s = t - tile * p.splits;
        kb0 = (int)((long long)p.nk * s / p.splits);
        kb1 = (int)((long long)p.nk * (s + 1) / p.splits);
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_items; t += gridDim.x) {
                int tile, kb0, kb1; item_range(t, tile, kb0, kb1);
                const int m0 = tile * LM_BM;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], LM_A_BYTES + LM_B_BYTES + (p.bulk ? p.n_ad * LM_W_BYTES : 0));
                    uint8_t* sa = smem + stage * stage_bytes;
                    uint8_t* sb = sa + p.n_ad * LM_A_BYTES;
                    const int k0 = kb * LM_BK;
                    if (p.bulk) {
                        // the keep words ride in the same ring as the tile they mask: requested by per-thread loads they queued
                        // behind ~100 KB of in-flight tile traffic per SM and arrived several k-blocks late
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            if (a >= p.n_ad) break;
                            uint8_t* sw = sa + words_off + a * LM_W_BYTES;
                            if constexpr (!A_MN) {          // 128 tokens of feature block kb
                                bulk_load_1d(sw, p.bits[a] + (size_t)kb * p.n_tok + m0, LM_W_BYTES, &full_bar[stage]);
                            } else {                        // 64 tokens of the tile's two feature blocks
                                bulk_load_1d(sw, p.bits[a] + (size_t)(m0 >> 6) * p.n_tok + k0, LM_W_BYTES / 2, &full_bar[stage]);
                                bulk_load_1d(sw + LM_W_BYTES / 2, p.bits[a] + (size_t)((m0 >> 6) + 1) * p.n_tok + k0, LM_W_BYTES / 2, &full_bar[stage]);
                            }
                        }
                    }
                    if constexpr (!A_MN) {
                        tma_load_2d(&tmA, &full_bar[stage], sa, k0, m0);          // x [tokens, features]: rows = tokens
                        tma_load_2d(&tmB, &full_bar[stage], sb, k0, 0);           // A [16 n_ad, features]
                    } else {
#pragma unroll
                        for (int j = 0; j < LM_BM / 64; ++j)                      // x tile as two 64-feature atoms x 64 tokens
                            tma_load_2d(&tmA, &full_bar[stage], sa + j * (LM_BK * 128), m0 + j * 64, k0);
                        tma_load_2d(&tmB, &full_bar[stage], sb, 0, k0);           // t [tokens, 16 n_ad]: rows = tokens
                    }
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(LM_BM, 16, A_MN, A_MN);
            constexpr uint32_t LBO = LM_BK * 128;
            constexpr uint32_t KSTEP = A_MN ? (16 * 128) : 32;
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < num_items; t += gridDim.x, ++it) {
                const int as = it & 1;
                mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                int tile, kb0, kb1; item_range(t, tile, kb0, kb1);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&ready_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * stage_bytes);
                    const uint32_t b_addr = a_addr + p.n_ad * LM_A_BYTES;
                    for (int a = 0; a < p.n_ad; ++a) {
                        // K-major B (forward): adapter a's 16 rows of A start 16 * 128 B further; MN-major B (backward): adapter
                        // a's 16 columns of t start 32 B into every 128-byte token row
                        const uint32_t b_a = b_addr + (A_MN ? a * 32 : a * 16 * 128);
#pragma unroll
                        for (int k = 0; k < LM_BK / 16; ++k) {
                            const uint64_t ad = make_smem_desc(a_addr + a * LM_A_BYTES + k * KSTEP, A_MN ? LBO : 0, 1024);
                            const uint64_t bd = make_smem_desc(b_a + k * KSTEP, A_MN ? LBO : 0, 1024);
                            umma_bf16(tmem_base + as * LM_BN + a * 16, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                    }
                    umma_commit(&empty_bar[stage]);
                    if (kb == kb1 - 1) umma_commit(&tfull_bar[as]);
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 8) {
        // ------------------------------------------------------------------ transform: masked copies of the x tile
        // Two groups of four warps take alternate k-blocks: a stage's chain (wait for the tile -> LDS -> mask -> STS -> proxy fence
        // -> arrive) is ~1000 cycles long while HBM delivers a tile every ~600, so with all eight warps on the same stage the
        // chain, not the memory system, set the pace. A thread owns one 128-byte row (8 chunks) of its group's stage.
        constexpr int GT = LM_XFORM_THREADS / LM_GROUPS;     // threads per group, each owning LM_BM / GT rows of its stage
        const int tt = threadIdx.x - 256;
        const int grp = tt / GT, row0 = tt % GT;
        const uint32_t smem_base = smem_u32(smem);
        int cyc = 0;                                         // k-blocks this CTA has been through (all groups count all of them)
        for (int t = blockIdx.x; t < num_items; t += gridDim.x) {
            int tile, kb0, kb1; item_range(t, tile, kb0, kb1);
            const int m0 = tile * LM_BM;
            for (int kb = kb0; kb < kb1; ++kb, ++cyc) {
                if ((cyc & (LM_GROUPS - 1)) != grp) continue;
                const int stage = cyc % p.stages;
                const uint32_t phase = (uint32_t)(cyc / p.stages) & 1u;
                const uint32_t st = smem_base + stage * stage_bytes;
                bool waited = false;
#pragma unroll 1
                for (int row = row0; row < LM_BM; row += GT) {
                unsigned long long w[3] = {~0ull, ~0ull, ~0ull};
                if (!p.bulk) {                            // ragged shapes: this row's keep words straight from global memory
                    const int tok = A_MN ? kb * LM_BK + (row & 63) : m0 + row;
                    const int fb = A_MN ? (m0 >> 6) + (row >> 6) : kb;
#pragma unroll
                    for (int a = 0; a < 3; ++a)
                        if (a < p.n_ad && tok < p.n_tok && fb < p.n_fb) w[a] = __ldg(p.bits[a] + (size_t)fb * p.n_tok + tok);
                }
                if (!waited) { mbar_wait(&full_bar[stage], phase); waited = true; }
                if (p.bulk) {                             // ... or landed with the tile
#pragma unroll
                    for (int a = 0; a < 3; ++a)
                        if (a < p.n_ad) {
                            uint32_t lo, hi;
                            lds64(st + words_off + a * LM_W_BYTES + row * 8, lo, hi);
                            w[a] = (unsigned long long)lo | ((unsigned long long)hi << 32);
                        }
                }
                const uint32_t sa = st + row * 128;
                uint4 v[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) v[c] = lds128(sa + ((c ^ (row & 7)) << 4));
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int off = (c ^ (row & 7)) << 4;
#pragma unroll
                    for (int a = 2; a >= 0; --a)          // adapter 0 is masked in place, so it goes last
                        if (a < p.n_ad) sts128(sa + a * LM_A_BYTES + off, mask8(v[c], (unsigned int)(w[a] >> (8 * c)) & 0xFFu));
                }
                }
                fence_proxy_async_smem();                 // generic-proxy writes -> visible to the tensor core's async proxy
                __syncwarp();
                if (lane == 0) mbar_arrive(&ready_bar[stage]);      // one arrival per warp
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue
        const int q = warp & 3;
        const int row = q * 32 + lane;
        int it = 0;
        for (int t = blockIdx.x; t < num_items; t += gridDim.x, ++it) {
            const int tile = t / p.splits;
            const int as = it & 1;
            mbar_wait(&tfull_bar[as], (it >> 1) & 1);
            tc_fence_after();
            uint32_t r0[32], r1[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * LM_BN), r0);
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * LM_BN + 32), r1);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&tempty_bar[as]);
            const int m = tile * LM_BM + row;
            if (m >= p.n_rows) continue;
            if constexpr (!A_MN) {
                bf16* dst = p.out_bf16 + (size_t)m * p.ld_out;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    if (a >= p.n_ad) break;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int c = a * 16 + h * 8 + j;
                            v[j] = __uint_as_float(c < 32 ? r0[c & 31] : r1[c & 31]) * p.alpha;
                        }
                        store8(dst + a * 16 + h * 8, v);
                    }
                }
            } else {
                // D[feature, 16 a + j] = dA[16 a + j][feature]: for a fixed column the 32 lanes of a warp add 128 contiguous bytes
#pragma unroll
                for (int c = 0; c < 48; ++c) {
                    if (c >= 16 * p.n_ad) break;
                    const float v = __uint_as_float(c < 32 ? r0[c & 31] : r1[c & 31]) * p.alpha;
                    atomicAdd(p.out_f32 + (size_t)c * p.ld_out + m, v);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc<2 * LM_BN>(tmem_base); }
}

template <bool A_MN>
static int launch_lora_masked(const LoraMaskParams& p0, const CUtensorMap& tA, const CUtensorMap& tB, cudaStream_t stream) {
    LoraMaskParams p = p0;
    const int stage_bytes = p.n_ad * LM_A_BYTES + LM_B_BYTES + (p.bulk ? p.n_ad * LM_W_BYTES : 0);
    p.stages = LM_SMEM_BUDGET / stage_bytes;
    if (p.stages > LM_MAX_STAGES) p.stages = LM_MAX_STAGES;
    const int smem = p.stages * stage_bytes + 1024 + 512;
    auto kern = lora_masked_kernel<A_MN>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, LM_SMEM_BUDGET + 1024 + 512);
        VLB_REQUIRE(e == cudaSuccess, "cudaFuncSetAttribute(lora_masked): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    int grid = p.mt * p.splits;
    if (grid > num_sms()) grid = num_sms();
    kern<<<grid, LM_THREADS, smem, stream>>>(p, tA, tB);
    return check_launch("lora_masked_kernel");
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_dropout_bits(void* bits64, long long n_tok, int n_feat, float p, unsigned long long seed, void* stream) {
    VLB_REQUIRE(n_tok > 0 && n_feat > 0 && n_feat % 64 == 0 && p >= 0.f && p < 1.f, "dropout_bits: tokens=%lld features=%d p=%g",
                n_tok, n_feat, p);
    const long long nwords = n_tok * (n_feat / 64);
    double th = (double)p * 4294967296.0 + 0.5;
    const unsigned int thresh = th >= 4294967295.0 ? 4294967295u : (unsigned int)th;
    dropout_bits_kernel<<<(unsigned)ceil_div_ll(nwords, DB_THREADS), DB_THREADS, 0, (cudaStream_t)stream>>>((unsigned long long*)bits64, nwords,
                                                                                                thresh, seed);
    return check_launch("dropout_bits");
}

static int lm_common(LoraMaskParams& p, int n_tok, int n_feat, int n_ad, const void* b0, const void* b1, const void* b2, float alpha) {
    VLB_REQUIRE(n_ad >= 1 && n_ad <= 3 && n_feat % 64 == 0 && n_tok > 0, "lora_masked: adapters=%d features=%d tokens=%d", n_ad,
                n_feat, n_tok);
    VLB_REQUIRE(b0 && (n_ad < 2 || b1) && (n_ad < 3 || b2), "lora_masked: missing mask bits for %d adapters", n_ad);
    p.n_tok = n_tok; p.n_ad = n_ad; p.alpha = alpha; p.n_fb = n_feat / 64; p.bulk = 0;
    p.bits[0] = (const unsigned long long*)b0; p.bits[1] = (const unsigned long long*)b1; p.bits[2] = (const unsigned long long*)b2;
    p.out_bf16 = nullptr; p.out_f32 = nullptr;
    return 0;
}

static bool lm_words_aligned(const LoraMaskParams& p) {        // cp.async.bulk needs 16-byte aligned sources
    if (getenv("VLB_LORA_WORDS_LDG")) return false;
    for (int a = 0; a < p.n_ad; ++a)
        if (reinterpret_cast<uintptr_t>(p.bits[a]) % 16 != 0) return false;
    return p.n_tok % 2 == 0;
}

extern "C" int vlb_lora_down_masked(const void* x, int ldx, const void* A, int lda, void* xa, int ldxa, int n_tok, int n_feat,
                                    int n_adapters, const void* bits0, const void* bits1, const void* bits2, float alpha,
                                    void* stream) {
    LoraMaskParams p;
    if (lm_common(p, n_tok, n_feat, n_adapters, bits0, bits1, bits2, alpha)) return 1;
    VLB_REQUIRE(ldxa % 8 == 0, "lora_down_masked: ldxa=%d", ldxa);
    CUtensorMap tA, tB;
    int rc;
    if ((rc = tmap2d_cached(&tA, x, n_feat, n_tok, ldx, LM_BK, LM_BM))) return rc;
    if ((rc = tmap2d_cached(&tB, A, n_feat, 16 * n_adapters, lda, LM_BK, LM_BN))) return rc;
    p.nk = n_feat / LM_BK; p.mt = ceil_div(n_tok, LM_BM); p.splits = 1;
    p.out_bf16 = (bf16*)xa; p.ld_out = ldxa; p.n_rows = n_tok;
    p.bulk = n_tok % LM_BM == 0 && lm_words_aligned(p);          // full token tiles: 1 KB of words per (feature block, tile)
    return launch_lora_masked<false>(p, tA, tB, (cudaStream_t)stream);
}

extern "C" int vlb_lora_grad_a_masked(const void* x, int ldx, const void* t, int ldt, void* dA_f32, int ldda, int n_tok,
                                      int n_feat, int n_adapters, const void* bits0, const void* bits1, const void* bits2,
                                      float alpha, void* stream) {
    LoraMaskParams p;
    if (lm_common(p, n_tok, n_feat, n_adapters, bits0, bits1, bits2, alpha)) return 1;
    VLB_REQUIRE(n_tok % 8 == 0, "lora_grad_a_masked: tokens=%d must be a multiple of 8", n_tok);
    CUtensorMap tA, tB;
    int rc;
    if ((rc = tmap2d_cached(&tA, x, n_feat, n_tok, ldx, 64, LM_BK))) return rc;             // box: 64 features x 64 tokens
    if ((rc = tmap2d_cached(&tB, t, 16 * n_adapters, n_tok, ldt, 64, LM_BK))) return rc;     // box: 64 columns (zero-filled) x 64 tokens
    p.nk = ceil_div(n_tok, LM_BK); p.mt = ceil_div(n_feat, LM_BM);
    // split the token contraction so that the (feature tile, split) items fill whole waves of the persistent grid: 112 tiles
    // on 148 SMs leave a quarter of the machine idle, 112 x 5 = 560 items are 3.8 waves (95 %); fewer splits = fewer atomics
    const int sms = num_sms();
    int best = 1; double best_eff = 0.0;
    for (int sp = 1; sp <= 16 && sp <= (p.nk >= 4 ? p.nk / 4 : 1); ++sp) {
        const int items = p.mt * sp;
        const double eff = (double)items / ((double)ceil_div(items, sms) * sms);
        if (eff > best_eff + 0.02) { best_eff = eff; best = sp; }
    }
    p.splits = best;
    { const char* g = getenv("VLB_LORA_GRAD_SPLITS"); if (g && atoi(g) > 0) p.splits = atoi(g); }
    p.out_f32 = (float*)dA_f32; p.ld_out = ldda; p.n_rows = n_feat;
    p.bulk = n_tok % LM_BK == 0 && n_feat % LM_BM == 0 && lm_words_aligned(p);   // full 64-token x 128-feature tiles
    return launch_lora_masked<true>(p, tA, tB, (cudaStream_t)stream);
}
