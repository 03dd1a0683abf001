// Flash-style attention on tcgen05 / TMEM for sm_100a: causal GQA (Mistral decoder, d=128) and non-causal MHA
// (CLIP ViT, d=64).  Replaces flash-attn 2 reached through `_attn_implementation="flash_attention_2"`
// (src/litmodule/videollama2_vlb_litmodule.py:71-72, VideoLLaMA2/videollama2/model/encoder.py:24).
//
// Forward, one CTA per (128 query rows, head, batch element), two CTAs resident per SM:
//   warp 0 : TMA producer   Q tile once; K/V tiles of 64 keys through 2-stage rings (3-D tensor maps [cols, S, B],
//            so a tile never straddles two sequences and rows past S arrive as zeros)
//   warp 1 : MMA issuer     S = Q K^T (M128 N64 K=d) into TMEM;  O += P V (M128 N=d K64) with P from shared memory
//   warp 2 : TMEM allocator
//   warps 4-7 : softmax     tcgen05.ld gives each thread ONE FULL ROW of S (no shuffles): scale, mask, running max
//            with lazy rescale of the TMEM-resident O (only when the max grows by > 2^8), exp2, row sum,
//            P -> bf16 -> 128B-swizzled shared memory; final O/l -> bf16 -> shared -> TMA store; LSE (log2 domain).
// Right-padded text (attention_mask = ids.ne(0)) needs no extra mask under causality: valid rows never see padded
// keys, and padded rows carry zero HRF weight so their values are don't-care.
#include "common.cuh"

namespace vlb {

constexpr int ATT_BM = 128;   // query rows per CTA
constexpr int ATT_BN = 64;    // keys per inner block
constexpr int ATT_THREADS = 256;

template <int D> struct AttCfg {
    static constexpr int Q_BYTES = ATT_BM * D * 2;
    static constexpr int KV_BYTES = ATT_BN * D * 2;
    static constexpr int P_BYTES = ATT_BM * ATT_BN * 2;
    static constexpr int K_STAGES = 3;                         // K is needed two key blocks ahead of V (S is issued two ahead)
    static constexpr int OFF_K = Q_BYTES;
    static constexpr int OFF_V = OFF_K + K_STAGES * KV_BYTES;
    static constexpr int OFF_BAR = OFF_V + 2 * KV_BYTES;      // P lives in TMEM: no staging buffer
    static constexpr int SMEM_BYTES = OFF_BAR + 256;          // + barriers; two CTAs fit one SM
    static constexpr int TMEM_COLS = 256;                     // two S buffers (64 columns each) + O (D columns); x2 CTAs = 512
    static constexpr int O_COL = 2 * ATT_BN;                  // S0 at column 0, S1 at 64, O after them
};

struct AttParams {
    int S;              // sequence length (rows per batch element)
    int kv_len;         // keys that exist (== S)
    int group;          // q heads per kv head
    float scale_log2;   // softmax scale * log2(e)
    float* lse;         // [B, Hq, S] log2-domain logsumexp
    int Hq;
    long long* ts;      // debug (VLB_ATT_TS=1): clock64 stamps of the hand-offs of one heavy CTA, 8 per key block
};

// P stays in TMEM. The softmax threads write bf16 P over the S buffer they just read (tcgen05.st) and the PV MMA takes
// its A operand from TMEM, so P never crosses shared memory (-32 of the 128 KB of smem traffic per key block; the kernel is
// shared-memory-bandwidth bound) and P(j) no longer has to wait for PV(j-1) to release a staging buffer.
#define ATT_TS(slot) do { if (ts_on) p.ts[j * 8 + (slot)] = clock64(); } while (0)
template <int D, bool CAUSAL>
__global__ void __launch_bounds__(ATT_THREADS, 2)
attn_fwd_kernel(const AttParams p, const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO) {
    using Cfg = AttCfg<D>;
    extern __shared__ __align__(1024) uint8_t smem[];      // 128B swizzle atoms need 1024-byte alignment
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* sQ = smem;
    uint8_t* sK = smem + Cfg::OFF_K;
    uint8_t* sV = smem + Cfg::OFF_V;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
    uint64_t* q_full = bars;            // [1]
    uint64_t* k_full = bars + 1;        // [3]
    uint64_t* k_empty = bars + 4;       // [3]
    uint64_t* v_full = bars + 7;        // [2]
    uint64_t* v_empty = bars + 9;       // [2]
    uint64_t* s_full = bars + 11;       // [2] one per S buffer
    uint64_t* p_ready = bars + 13;      // [2] count 128, alternating: a softmax warp may be one key block ahead of its
                                        //     siblings (it no longer waits for PV(j-1)), and its arrival for block j+1 must
                                        //     not be counted into block j's phase
    uint64_t* pv_done = bars + 15;      // [2] alternating, so a waiter may lag one PV behind
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 17);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nqb = gridDim.x;
    const int qb = CAUSAL ? (nqb - 1 - (int)blockIdx.x) : (int)blockIdx.x;   // heavy (late) causal blocks first
    const int h = blockIdx.y, b = blockIdx.z;
    const int hk = h / p.group;
    const int q0 = qb * ATT_BM;
    const int kv_end = CAUSAL ? min(p.kv_len, q0 + ATT_BM) : p.kv_len;
    const int nblk = (kv_end + ATT_BN - 1) / ATT_BN;
    const bool ts_on = p.ts != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && ((threadIdx.x & 31) == 0 || (threadIdx.x >> 5) == 1);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < Cfg::K_STAGES; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1); mbar_init(&p_ready[0], 128); mbar_init(&p_ready[1], 128); mbar_init(&pv_done[0], 1); mbar_init(&pv_done[1], 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(q_full, Cfg::Q_BYTES);
#pragma unroll
            for (int c = 0; c < D / 64; ++c) tma_load_3d(&tmQ, q_full, sQ + c * (ATT_BM * 128), h * D + c * 64, q0, b);
            int ks = 0; uint32_t kph = 0;
            for (int j = 0; j < nblk; ++j) {
                const int st = j & 1; const uint32_t ph = (j >> 1) & 1;
                mbar_wait(&k_empty[ks], kph ^ 1);
                mbar_expect_tx(&k_full[ks], Cfg::KV_BYTES);
#pragma unroll
                for (int c = 0; c < D / 64; ++c)
                    tma_load_3d(&tmK, &k_full[ks], sK + ks * Cfg::KV_BYTES + c * (ATT_BN * 128), hk * D + c * 64, j * ATT_BN, b);
                if (++ks == Cfg::K_STAGES) { ks = 0; kph ^= 1; }
                mbar_wait(&v_empty[st], ph ^ 1);
                mbar_expect_tx(&v_full[st], Cfg::KV_BYTES);
#pragma unroll
                for (int c = 0; c < D / 64; ++c)
                    tma_load_3d(&tmV, &v_full[st], sV + st * Cfg::KV_BYTES + c * (ATT_BN * 128), hk * D + c * 64, j * ATT_BN, b);
            }
        }
    } else if (warp == 1) {
        // The whole warp runs the issue loop (warp-uniform control flow and descriptor arithmetic: no per-lane waterfall loops
        // around the tcgen05 instructions); one elected lane issues the MMAs and commits.
        {
            const bool ldr = elect_one();
            constexpr uint32_t idesc_qk = make_idesc_bf16(ATT_BM, ATT_BN, false, false);
            constexpr uint32_t idesc_pv = make_idesc_bf16(ATT_BM, D, false, true);
            const uint32_t tS = tmem_base, tO = tmem_base + Cfg::O_COL;
            int ks = 0; uint32_t kph = 0;                   // K ring position of the next QK to issue (issued in order)
            auto issue_qk = [&](int j) {
                const int st = j & 1;
                mbar_wait(&k_full[ks], kph);
                tc_fence_after();
                const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK + ks * Cfg::KV_BYTES);
#pragma unroll
                for (int k = 0; k < D / 16; ++k) {
                    const uint32_t off_q = (k >> 2) * (ATT_BM * 128) + (k & 3) * 32;
                    const uint32_t off_k = (k >> 2) * (ATT_BN * 128) + (k & 3) * 32;
                    if (ldr) umma_bf16(tS + (uint32_t)(st * ATT_BN), make_smem_desc(qa + off_q, 0, 1024),
                                       make_smem_desc(ka + off_k, 0, 1024), idesc_qk, k > 0 ? 1u : 0u);
                }
                if (ldr) { umma_commit(&k_empty[ks]); umma_commit(&s_full[st]); }
                __syncwarp();
                if (++ks == Cfg::K_STAGES) { ks = 0; kph ^= 1; }
            };
            // S is double-buffered in TMEM and issued two key blocks ahead, so the tensor pipe computes S(j+1) while the
            // softmax threads are still working on S(j); PV(j) (on the critical path to the next P write) goes first.
            mbar_wait(q_full, 0);
            issue_qk(0);
            if (nblk > 1) issue_qk(1);
            for (int j = 0; j < nblk; ++j) {
                const int st = j & 1;
                mbar_wait(&p_ready[j & 1], (j >> 1) & 1);  // P(j) in TMEM (over S buffer j & 1)
                tc_fence_after();
                if (ldr) ATT_TS(4);
                mbar_wait(&v_full[st], (j >> 1) & 1);
                tc_fence_after();
                if (ldr) ATT_TS(5);
                const uint32_t va = smem_u32(sV + st * Cfg::KV_BYTES);
#pragma unroll
                for (int k = 0; k < ATT_BN / 16; ++k) {
                    // A = P [128 x 64] from TMEM (bf16 pairs, 8 columns per K=16); B = V tile [64 keys x D], MN-major operand
                    if (ldr) umma_bf16_ts(tO, tS + (uint32_t)(st * ATT_BN + k * 8),
                                          make_smem_desc(va + k * (16 * 128), ATT_BN * 128, 1024), idesc_pv, (j > 0 || k > 0) ? 1u : 0u);
                }
                if (ldr) { umma_commit(&v_empty[st]); umma_commit(&pv_done[j & 1]); }
                __syncwarp();
                if (ldr) ATT_TS(6);
                if (j + 2 < nblk) issue_qk(j + 2);
                if (ldr) ATT_TS(7);
            }
        }
    } else if (warp >= 4) {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const uint32_t tS = tmem_base + lane_off, tO = tmem_base + lane_off + Cfg::O_COL;
        const int qrow = q0 + row;
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < nblk; ++j) {
            mbar_wait(&s_full[j & 1], (j >> 1) & 1);
            tc_fence_after();
            if (warp == 4) ATT_TS(0);
            uint32_t r0[32], r1[32];
            tmem_ld32(tS + (uint32_t)((j & 1) * ATT_BN), r0);
            tmem_ld32(tS + (uint32_t)((j & 1) * ATT_BN) + 32, r1);
            tmem_ld_wait();
            if (warp == 4) ATT_TS(1);
            // raw scores: the softmax scale is folded into the exponent's FMA (max commutes with the positive scale)
            float s[64];
#pragma unroll
            for (int c = 0; c < 32; ++c) { s[c] = __uint_as_float(r0[c]); s[32 + c] = __uint_as_float(r1[c]); }
            const int k0 = j * ATT_BN;
            const bool need_mask = (CAUSAL && k0 + ATT_BN - 1 > q0) || (k0 + ATT_BN > p.kv_len);
            if (need_mask) {                                  // block-uniform: a real branch, diagonal / ragged blocks only
                asm volatile("" ::: "memory");
                const int nvalid = (CAUSAL ? min(qrow, p.kv_len - 1) : (p.kv_len - 1)) - k0 + 1;
#pragma unroll
                for (int c = 0; c < 64; ++c) s[c] = c < nvalid ? s[c] : -INFINITY;
            }
            float mx0 = fmaxf(s[0], s[1]), mx1 = fmaxf(s[2], s[3]);
#pragma unroll
            for (int c = 4; c < 64; c += 4) { mx0 = fmaxf(mx0, fmaxf(s[c], s[c + 1])); mx1 = fmaxf(mx1, fmaxf(s[c + 2], s[c + 3])); }
            const float mx = fmaxf(mx0, mx1) * p.scale_log2;
            const bool grow = mx > m_used + 8.f;
            const bool any_grow = j > 0 && __any_sync(0xffffffffu, grow);
            // P(j) goes to its own S buffer: only a rescale of O needs PV(j-1) retired. PV(j-2) is consumed every
            // iteration so the parity of each of the two alternating barriers never runs two phases ahead.
            if (j > 1) mbar_wait(&pv_done[j & 1], ((j - 2) >> 1) & 1);
            if (any_grow) mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);
            if (j == 0) {
                m_used = (mx == -INFINITY) ? 0.f : mx;
            } else if (any_grow) {
                tc_fence_after();
                const float m_new = grow ? mx : m_used;
                const float f = exp2f(m_used - m_new);
#pragma unroll 1
                for (int c0 = 0; c0 < D; c0 += 32) {
                    uint32_t o[32];
                    tmem_ld32(tO + c0, o);
                    tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * f);
                    tmem_st32(tO + c0, o);
                }
                tmem_st_wait();
                l *= f;
                m_used = m_new;
            }
            // packed fp32 (FFMA2 / FADD2, sm_100) for the exponent's FMA and the row sum. Two experiments of this round are NOT here
            // because they were slower (profiles/r02aq_*, r02ar_*): Q as a TMEM A operand with a single S buffer
            // (682 vs 929 TF/s) and part of the exponentials as polynomials on the FMA pipe (655 with 2 of 8).
            float2 sum2 = make_float2(0.f, 0.f);
            const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(-m_used, -m_used);
            uint32_t pk[32];                                  // PT: the row's 64 probabilities as bf16 pairs
#pragma unroll
            for (int c2 = 0; c2 < 32; ++c2) {
                const float2 x = __ffma2_rn(make_float2(s[2 * c2], s[2 * c2 + 1]), sc2, nm2);
                const float2 e = make_float2(exp2f(x.x), exp2f(x.y));
                sum2 = __fadd2_rn(sum2, e);                   // fp32 row sum (as flash-attn)
                pk[c2] = pack_bf16(e.x, e.y);
            }
            const float sum = sum2.x + sum2.y;
            l += sum;
            if (warp == 4) ATT_TS(2);
            tmem_st32(tS + (uint32_t)((j & 1) * ATT_BN), pk);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&p_ready[j & 1]);
            if (warp == 4) ATT_TS(3);
        }
        // ---- epilogue: O / l -> bf16 -> smem (Q region, free: every QK MMA has retired) -> TMA store
        mbar_wait(&pv_done[(nblk - 1) & 1], ((nblk - 1) >> 1) & 1);
        tc_fence_after();
        const float inv_l = l > 0.f ? 1.f / l : 0.f;
#pragma unroll 1
        for (int c0 = 0; c0 < D; c0 += 32) {
            uint32_t o[32];
            tmem_ld32(tO + c0, o);
            tmem_ld_wait();
            uint8_t* dst = sQ + (c0 / 64) * (ATT_BM * 128) + row * 128;
#pragma unroll
            for (int c8 = 0; c8 < 4; ++c8) {
                uint4 u;
                u.x = pack_bf16(__uint_as_float(o[c8 * 8 + 0]) * inv_l, __uint_as_float(o[c8 * 8 + 1]) * inv_l);
                u.y = pack_bf16(__uint_as_float(o[c8 * 8 + 2]) * inv_l, __uint_as_float(o[c8 * 8 + 3]) * inv_l);
                u.z = pack_bf16(__uint_as_float(o[c8 * 8 + 4]) * inv_l, __uint_as_float(o[c8 * 8 + 5]) * inv_l);
                u.w = pack_bf16(__uint_as_float(o[c8 * 8 + 6]) * inv_l, __uint_as_float(o[c8 * 8 + 7]) * inv_l);
                const int chunk = ((c0 & 63) >> 3) + c8;
                sts128(smem_u32(dst) + ((chunk ^ (row & 7)) << 4), u);
            }
        }
        if (p.lse != nullptr && qrow < p.S)
            p.lse[((size_t)b * p.Hq + h) * p.S + qrow] = (l > 0.f) ? (m_used + log2f(l)) : -INFINITY;
        tc_fence_before();
        fence_proxy_async_smem();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (warp == 4 && lane == 0) {
#pragma unroll
            for (int c = 0; c < D / 64; ++c) tma_store_3d(&tmO, sQ + c * (ATT_BM * 128), h * D + c * 64, q0, b);
            tma_store_commit();
            tma_store_wait_all<0>();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc<Cfg::TMEM_COLS>(tmem_base); }
}

// 3-D map over a [B, S, cols] bf16 view with row stride ld: dims (cols, S, B), box (64, rows, 1)
static int att_tmap(CUtensorMap* m, const void* ptr, int cols, int S, int B, int ld, int box_rows) {
    unsigned long long dims[3] = {(unsigned long long)cols, (unsigned long long)S, (unsigned long long)B};
    unsigned long long strides[2] = {(unsigned long long)ld * 2ull, (unsigned long long)S * ld * 2ull};
    unsigned int box[3] = {64u, (unsigned int)box_rows, 1u};
    return encode_tmap_bf16(m, ptr, 3, dims, strides, box);
}

// debug timeline buffer (VLB_ATT_TS=1): 64 key blocks x 8 stamps, read back with vlb_attn_debug_ts
static long long* att_ts_buffer() {
    static long long* buf = [] {
        const char* e = getenv("VLB_ATT_TS");
        long long* b = nullptr;
        if (e && e[0] == '1') { cudaMalloc(&b, 64 * 8 * sizeof(long long)); cudaMemset(b, 0, 64 * 8 * sizeof(long long)); }
        return b;
    }();
    return buf;
}

template <int D, bool CAUSAL>
static int launch_attn_fwd(const void* q, const void* k, const void* v, int ld_qkv, void* o, int ld_o, float* lse, int B,
                           int S, int Hq, int Hkv, float scale, cudaStream_t st) {
    using Cfg = AttCfg<D>;
    CUtensorMap tq, tk, tv, to;
    int rc;
    if ((rc = att_tmap(&tq, q, Hq * D, S, B, ld_qkv, ATT_BM))) return rc;
    if ((rc = att_tmap(&tk, k, Hkv * D, S, B, ld_qkv, ATT_BN))) return rc;
    if ((rc = att_tmap(&tv, v, Hkv * D, S, B, ld_qkv, ATT_BN))) return rc;
    if ((rc = att_tmap(&to, o, Hq * D, S, B, ld_o, ATT_BM))) return rc;
    AttParams p;
    p.S = S; p.kv_len = S; p.group = Hq / Hkv; p.scale_log2 = scale * 1.4426950408889634f; p.lse = lse; p.Hq = Hq;
    p.ts = att_ts_buffer();
    auto kern = attn_fwd_kernel<D, CAUSAL>;
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        VLB_REQUIRE(e == cudaSuccess, "attn_fwd smem attr (%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
        attr = true;
    }
    dim3 grid(ceil_div(S, ATT_BM), Hq, B);
    kern<<<grid, ATT_THREADS, Cfg::SMEM_BYTES, st>>>(p, tq, tk, tv, to);
    return check_launch("attn_fwd_kernel");
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_attn_debug_ts(long long* host_out, int n) {
    long long* b = att_ts_buffer();
    VLB_REQUIRE(b != nullptr && n <= 64 * 8, "attn_debug_ts: set VLB_ATT_TS=1 (n=%d)", n);
    return cudaMemcpy(host_out, b, n * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1;
}

extern "C" int vlb_attn_fwd(const void* q, const void* k, const void* v, int ld_qkv, void* o, int ld_o, void* lse_f32,
                            int B, int S, int Hq, int Hkv, int D, int causal, float scale, void* stream) {
    VLB_REQUIRE(Hkv > 0 && Hq % Hkv == 0 && (D == 64 || D == 128), "attn_fwd: Hq=%d Hkv=%d D=%d", Hq, Hkv, D);
    VLB_REQUIRE(ld_qkv % 8 == 0 && ld_o % 8 == 0, "attn_fwd: leading dims must be multiples of 8");
    cudaStream_t st = (cudaStream_t)stream;
    if (D == 128 && causal) return launch_attn_fwd<128, true>(q, k, v, ld_qkv, o, ld_o, (float*)lse_f32, B, S, Hq, Hkv, scale, st);
    if (D == 128) return launch_attn_fwd<128, false>(q, k, v, ld_qkv, o, ld_o, (float*)lse_f32, B, S, Hq, Hkv, scale, st);
    if (causal) return launch_attn_fwd<64, true>(q, k, v, ld_qkv, o, ld_o, (float*)lse_f32, B, S, Hq, Hkv, scale, st);
    return launch_attn_fwd<64, false>(q, k, v, ld_qkv, o, ld_o, (float*)lse_f32, B, S, Hq, Hkv, scale, st);
}
