// Flash attention backward on tcgen05 / TMEM (causal GQA, d = 128) -- the decoder's attention gradient.
// Replaces flash-attn 2's backward reached by autograd through `_attn_implementation="flash_attention_2"`
// (src/litmodule/videollama2_vlb_litmodule.py:71-72).
//
// One CTA per (128-key block j, kv head, batch element); it keeps K_j, V_j in shared memory and dK_j, dV_j in TMEM
// while it sweeps every query head of the GQA group and every 64-row query block i that can see the keys.
// All score-shaped tiles are computed TRANSPOSED (keys on the 128 TMEM lanes, queries on 64 columns) so that each
// softmax thread owns one key row, P^T / dS^T land in shared memory already in the K-major layout the dV / dK MMAs
// want, and LSE / Delta are per-column broadcasts:
//   (1) S^T  = K_j Q_i^T          (2) dP^T = V_j dO_i^T           [M128 N64 K128]  -> TMEM
//   threads : P^T = exp2(S^T*c - LSE), dS^T = P^T o (dP^T - Delta)  -> bf16 -> swizzled smem
//   (3) dV_j += P^T dO_i          (4) dK_j += dS^T Q_i            [M128 N128 K64]  accumulate in TMEM over the sweep
//   (5) dQ_i^T = K_j^T dS^T       [M128(d) N64(q) K128(keys)]  -> TMEM -> fp32 red.global (coalesced along d)
//
// Software pipeline (r01b): the tensor pipe and the softmax threads never wait on each other's full step.
//   * S^T is double-buffered in TMEM and issued TWO steps ahead, dP^T one step ahead, so while the 8 softmax warps
//     work on step t the tensor core runs dP(t+1), dV/dK/dQ(t-1..t) and S(t+2);
//   * P^T / dS^T are double-buffered in shared memory, Q/dO/LSE/Delta ride a 3-stage TMA ring;
//   * the softmax side is two warpgroups splitting the 64 query columns (a warp can only reach its own 32 TMEM lanes,
//     so the split has to be along columns); the dQ^T drain of step t-1 happens after step t's P/dS hand-off.
// TMEM: S0 | S1 | dP | dQ (64 columns each) | dV | dK (128 each) = 512 columns.
// warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-11 softmax / dQ reduction / epilogue.
#include "common.cuh"

#include <stdlib.h>

namespace vlb {

constexpr int BW_D = 128;
constexpr int BW_BK = 128;      // keys per CTA
constexpr int BW_BQ = 64;       // query rows per inner step
constexpr int BW_NST = 3;       // Q/dO ring stages
constexpr int BW_THREADS = 512;      // warp 0 TMA, 1 MMA issuer, 2 TMEM alloc, 4-11 softmax (two warpgroups), 12-15 dQ drain

struct BwCfg {
    static constexpr int KV_BYTES = BW_BK * BW_D * 2;          // 32 KB each
    static constexpr int QDO_BYTES = BW_BQ * BW_D * 2;         // 16 KB each
    static constexpr int PS_BYTES = BW_BK * BW_BQ * 2;         // 16 KB each
    static constexpr int OFF_K = 0;
    static constexpr int OFF_V = KV_BYTES;
    static constexpr int OFF_Q = 2 * KV_BYTES;                         // [3 stages]
    static constexpr int OFF_DO = OFF_Q + BW_NST * QDO_BYTES;          // [3 stages]
    static constexpr int OFF_P = OFF_DO + BW_NST * QDO_BYTES;          // [2 buffers]
    static constexpr int OFF_DS = OFF_P + 2 * PS_BYTES;                // [2 buffers]
    static constexpr int OFF_VEC = OFF_DS + 2 * PS_BYTES;              // [3 stages][lse 64 | delta 64] floats
    static constexpr int OFF_BAR = OFF_VEC + BW_NST * 2 * BW_BQ * 4;
    static constexpr int SMEM_BYTES = OFF_BAR + 256;
    static constexpr int T_S = 0, T_DP = 128, T_DQ = 192, T_DV = 256, T_DK = 384;   // TMEM columns (S: two 64-col buffers)
    static constexpr int TMEM_COLS = 512;
};
static_assert(BwCfg::SMEM_BYTES <= 232448, "attention backward shared memory exceeds the 227 KB per-CTA limit");

struct BwParams {
    int S, Hq, Hkv, group, causal;
    float scale, scale_log2;
    const float* lse;       // [B, Hq, S] log2 domain
    const float* delta;     // [B, Hq, S]
    float* dq;              // [B*S, Hq*D] fp32, pre-zeroed
    bf16* dk; bf16* dv;     // rows [B*S], row stride ld_dkv, head hk at column hk*D
    int ld_dkv;
    int nkb, nbh;           // key blocks per sequence, Hkv * B
    int chunk;              // batch elements per dispatch chunk
    int ts_cta;             // which CTA records the timeline (VLB_ATT_TS_CTA, default 0)
    long long* ts;          // debug (VLB_ATT_TS=1): clock64 stamps of the hand-offs of CTA 0, 16 per step (tools/attn_timeline.py bwd)
    int dbg;                // VLB_BW_DBG timing experiments (bit0: skip dQ reduction, bit1: skip exp2) -- never set in product runs
};

#define BW_TS(slot) do { if (ts_on && t < 128) p.ts[t * 16 + (slot)] = clock64(); } while (0)
__global__ void __launch_bounds__(BW_THREADS, 1)
attn_bwd_kernel(const BwParams p, const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQ) {
    using Cfg = BwCfg;
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* sK = smem + Cfg::OFF_K;
    uint8_t* sV = smem + Cfg::OFF_V;
    uint8_t* sQ = smem + Cfg::OFF_Q;
    uint8_t* sDO = smem + Cfg::OFF_DO;
    uint8_t* sP = smem + Cfg::OFF_P;
    uint8_t* sDS = smem + Cfg::OFF_DS;
    float* sVec = reinterpret_cast<float*>(smem + Cfg::OFF_VEC);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
    uint64_t* kv_full = bars;            // [1]
    uint64_t* qdo_full = bars + 1;       // [3]
    uint64_t* qdo_empty = bars + 4;      // [3]
    uint64_t* s_full = bars + 7;         // [2] S^T buffer (t & 1) ready
    uint64_t* dp_full = bars + 9;        // dP^T(t) ready
    uint64_t* pds_ready = bars + 10;     // count 256: P^T, dS^T(t) in smem; S(t) / dP(t) TMEM drained
    uint64_t* mma_done = bars + 11;      // [2] alternating: (3)(4)(5) of step t retired -> dS buffer t & 1 free, dQ^T(t) readable
    uint64_t* dq_free = bars + 13;       // count 128: dQ^T TMEM drained by the drain warpgroup
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // (batch, kv head)-major dispatch keeps the fp32 dQ rows that co-resident CTAs reduce into (4 heads x S x 128 x 4 B
    // = 4 MB per group) inside L2; a key-block-major order touched the whole 268 MB dQ buffer per wave and turned the
    // red.adds into DRAM read-modify-writes (ncu r01b: 3.4 GB of DRAM traffic per launch at B=8). Within a group the
    // heaviest key blocks (j = 0 sees every query block) go first.
    // order: batch chunks of p.chunk elements; inside a chunk key-block-major (heaviest first), then kv head, batch
    const int per_chunk = p.nkb * p.Hkv * p.chunk;
    const int c_ = blockIdx.x / per_chunk, r_ = blockIdx.x % per_chunk;
    const int nb_c = min(p.chunk, p.nbh / p.Hkv - c_ * p.chunk);      // batch elements in this chunk (last may be short)
    const int j = r_ / (p.Hkv * nb_c);
    const int hk = (r_ % (p.Hkv * nb_c)) % p.Hkv, b = c_ * p.chunk + (r_ % (p.Hkv * nb_c)) / p.Hkv;
    const int k0 = j * BW_BK;
    const int nq = p.S / BW_BQ;
    const int i_start = p.causal ? (k0 / BW_BQ) : 0;
    const int per_head = nq - i_start;
    const int T = p.group * per_head;
    const bool ts_on = p.ts != nullptr && (int)blockIdx.x == p.ts_cta;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmDQ);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(kv_full, 1);
        for (int i = 0; i < BW_NST; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 1); }
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1); mbar_init(dp_full, 1);
        mbar_init(pds_ready, 256); mbar_init(&mma_done[0], 1); mbar_init(&mma_done[1], 1); mbar_init(dq_free, 128);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(kv_full, 2 * Cfg::KV_BYTES);
#pragma unroll
            for (int c = 0; c < BW_D / 64; ++c) {
                tma_load_3d(&tmK, kv_full, sK + c * (BW_BK * 128), hk * BW_D + c * 64, k0, b);
                tma_load_3d(&tmV, kv_full, sV + c * (BW_BK * 128), hk * BW_D + c * 64, k0, b);
            }
            int st = 0; uint32_t ph = 0;
            int hh = 0, qi = 0;                             // step t = (head hh of the group, query block qi): no division per step
            for (int t = 0; t < T; ++t) {
                const int h = hk * p.group + hh;
                const int q0 = (i_start + qi) * BW_BQ;
                if (++qi == per_head) { qi = 0; ++hh; }
                mbar_wait(&qdo_empty[st], ph ^ 1);
                mbar_expect_tx(&qdo_full[st], 2 * Cfg::QDO_BYTES + 2 * BW_BQ * 4);
#pragma unroll
                for (int c = 0; c < BW_D / 64; ++c) {
                    tma_load_3d(&tmQ, &qdo_full[st], sQ + st * Cfg::QDO_BYTES + c * (BW_BQ * 128), h * BW_D + c * 64, q0, b);
                    tma_load_3d(&tmDO, &qdo_full[st], sDO + st * Cfg::QDO_BYTES + c * (BW_BQ * 128), h * BW_D + c * 64, q0, b);
                }
                const size_t voff = ((size_t)b * p.Hq + h) * p.S + q0;
                bulk_load_1d(sVec + st * 2 * BW_BQ, p.lse + voff, BW_BQ * 4, &qdo_full[st]);
                bulk_load_1d(sVec + st * 2 * BW_BQ + BW_BQ, p.delta + voff, BW_BQ * 4, &qdo_full[st]);
                if (++st == BW_NST) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // The whole warp runs the issue loop (warp-uniform control flow and descriptor arithmetic, so the compiler keeps the
        // operands in uniform registers instead of wrapping every tcgen05.mma in a per-lane waterfall loop); one elected
        // lane issues the MMAs and commits.
        {
            const bool ldr = elect_one();
            constexpr uint32_t id_s = make_idesc_bf16(128, BW_BQ, false, false);    // (1)(2): A,B K-major
            constexpr uint32_t id_acc = make_idesc_bf16(128, BW_D, false, true);    // (3)(4): A K-major, B MN-major
            constexpr uint32_t id_dq = make_idesc_bf16(128, BW_BQ, true, true);     // (5): A,B MN-major
            const uint32_t ka = smem_u32(sK), va = smem_u32(sV);
            auto issue_s = [&](int t) {            // (1) S^T(t) -> S buffer t & 1 ; waits for the Q/dO stage of step t
                const int st = t % BW_NST;
                mbar_wait(&qdo_full[st], (t / BW_NST) & 1);
                tc_fence_after();
                const uint32_t qa = smem_u32(sQ + st * Cfg::QDO_BYTES);
#pragma unroll
                for (int k = 0; k < BW_D / 16; ++k) {
                    const uint32_t offA = (k >> 2) * (BW_BK * 128) + (k & 3) * 32;
                    const uint32_t offB = (k >> 2) * (BW_BQ * 128) + (k & 3) * 32;
                    if (ldr) umma_bf16(tmem_base + Cfg::T_S + (t & 1) * BW_BQ, make_smem_desc(ka + offA, 0, 1024),
                              make_smem_desc(qa + offB, 0, 1024), id_s, k > 0 ? 1u : 0u);
                }
                if (ldr) umma_commit(&s_full[t & 1]);
            };
            auto issue_dp = [&](int t) {           // (2) dP^T(t); the stage was already waited for by issue_s(t)
                const uint32_t doa = smem_u32(sDO + (t % BW_NST) * Cfg::QDO_BYTES);
#pragma unroll
                for (int k = 0; k < BW_D / 16; ++k) {
                    const uint32_t offA = (k >> 2) * (BW_BK * 128) + (k & 3) * 32;
                    const uint32_t offB = (k >> 2) * (BW_BQ * 128) + (k & 3) * 32;
                    if (ldr) umma_bf16(tmem_base + Cfg::T_DP, make_smem_desc(va + offA, 0, 1024), make_smem_desc(doa + offB, 0, 1024),
                              id_s, k > 0 ? 1u : 0u);
                }
                if (ldr) umma_commit(dp_full);
            };
            mbar_wait(kv_full, 0);
            issue_s(0);
            issue_dp(0);
            if (T > 1) issue_s(1);
            for (int t = 0; t < T; ++t) {
                const int st = t % BW_NST;
                const uint32_t qa = smem_u32(sQ + st * Cfg::QDO_BYTES), doa = smem_u32(sDO + st * Cfg::QDO_BYTES);
                const uint32_t dsa = smem_u32(sDS + (t & 1) * Cfg::PS_BYTES);
                mbar_wait(pds_ready, t & 1);       // P/dS(t) in smem / TMEM; S(t), dP(t) TMEM columns drained
                tc_fence_after();
                if (ldr) BW_TS(7);
                if (t + 1 < T) issue_dp(t + 1);
                if (ldr) BW_TS(8);
                {
                    // (3)(4) with the A operand from TMEM: P^T / dS^T were packed (bf16 pairs along the query axis) over the
                    // S buffer by the softmax threads -- warpgroup g: P^T at columns 32g..32g+15, dS^T at 32g+16..32g+31
                    const uint32_t ta = tmem_base + Cfg::T_S + (t & 1) * BW_BQ;
#pragma unroll
                    for (int k = 0; k < BW_BQ / 16; ++k) {
                        const uint32_t ca = (uint32_t)((k >> 1) * 32 + (k & 1) * 8);
                        if (ldr) umma_bf16_ts(tmem_base + Cfg::T_DV, ta + ca,
                                     make_smem_desc(doa + k * (16 * 128), BW_BQ * 128, 1024), id_acc, (t > 0 || k > 0) ? 1u : 0u);
                        if (ldr) umma_bf16_ts(tmem_base + Cfg::T_DK, ta + ca + 16,
                                     make_smem_desc(qa + k * (16 * 128), BW_BQ * 128, 1024), id_acc, (t > 0 || k > 0) ? 1u : 0u);
                    }
                }
                if (ldr) BW_TS(9);
                if (t > 0) { mbar_wait(dq_free, (t - 1) & 1); tc_fence_after(); }
                if (ldr) BW_TS(10);
#pragma unroll
                for (int k = 0; k < BW_BK / 16; ++k) {          // (5) dQ^T = K^T dS^T, contraction over the 128 keys
                    if (ldr) umma_bf16(tmem_base + Cfg::T_DQ, make_smem_desc(ka + k * (16 * 128), BW_BK * 128, 1024),
                              make_smem_desc(dsa + k * (16 * 128), BW_BK * 128, 1024), id_dq, k > 0 ? 1u : 0u);
                }
                if (ldr) umma_commit(&qdo_empty[st]);
                if (ldr) umma_commit(&mma_done[t & 1]);
                if (ldr) BW_TS(11);
                if (t + 2 < T) issue_s(t + 2);
                if (ldr) BW_TS(12);     // S buffer t & 1 was drained before pds_ready(t)
            }
        }
    } else if (warp >= 12) {
        // ------------------------------------------------------------------ dQ drain warpgroup (r01d)
        // dQ^T(t) [128 d x 64 q] leaves through shared memory: each thread holds the 64 query columns of its d row, writes
        // them transposed into a plain [64 q][128 d] fp32 tile (a warp stores 32 consecutive floats: conflict-free), and ONE
        // TMA reduce-add (cp.reduce.async.bulk.tensor ... .add) folds the 32 KB tile into dq. It used to be 32 red.global per
        // softmax thread (~1000 cycles of every 2900-cycle step, timeline r01d); in its own warpgroup the drain also leaves
        // the softmax warps' serial chain.
        const int qd = warp & 3;
        const int row = qd * 32 + lane;                  // d index
        const uint32_t tb = tmem_base + ((uint32_t)(qd * 32) << 16);
        float* stg = reinterpret_cast<float*>(sP);       // the former smem P buffers (P lives in TMEM)
        const bool red_issuer = (warp == 12 && lane == 0);
        int hh = 0, qi = 0;
        for (int t = 0; t < T; ++t) {
            const int h = hk * p.group + hh;
            const int q0 = (i_start + qi) * BW_BQ;
            if (++qi == per_head) { qi = 0; ++hh; }
            mbar_wait(&mma_done[t & 1], (t >> 1) & 1);
            tc_fence_after();
            if (red_issuer) tma_store_wait_read<0>();    // the previous tile has been read out of the staging buffer
            uint32_t r0[32], r1[32];
            tmem_ld32(tb + Cfg::T_DQ, r0);
            tmem_ld32(tb + Cfg::T_DQ + 32, r1);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(dq_free);                        // registers hold the tile: the next dQ^T MMA may overwrite TMEM
            asm volatile("bar.sync 2, 128;" ::: "memory");
#pragma unroll
            for (int c = 0; c < 32; ++c) stg[c * BW_D + row] = __uint_as_float(r0[c]);
#pragma unroll
            for (int c = 0; c < 32; ++c) stg[(32 + c) * BW_D + row] = __uint_as_float(r1[c]);
            fence_proxy_async_smem();
            asm volatile("bar.sync 2, 128;" ::: "memory");
            if (red_issuer && !(p.dbg & 1)) {
                tma_reduce_add_2d(&tmDQ, stg, h * BW_D, b * p.S + q0);
                tma_store_commit();
            }
            if (warp == 12 && lane == 0) BW_TS(6);
        }
        if (red_issuer) tma_store_wait_all<0>();
    } else if (warp >= 4) {
        const int qd = warp & 3;                         // TMEM lane quadrant this warp may touch
        const int wg = (warp - 4) >> 2;                  // which 32 of the 64 query columns
        const int row = qd * 32 + lane;                  // key row (steps 1-4) / d index (step 5)
        const uint32_t tb = tmem_base + ((uint32_t)(qd * 32) << 16);
        const int key = k0 + row;
        const int cb = wg * 32;
        int qi = 0, st = 0;
        uint32_t st_ph = 0;
        for (int t = 0; t < T; ++t) {
            const int q0 = (i_start + qi) * BW_BQ;
            mbar_wait(&qdo_full[st], st_ph);             // lse/delta vectors landed (bulk copies on the same barrier)
            if (warp == 4 && lane == 0) BW_TS(0);
            const float* lse = sVec + st * 2 * BW_BQ + cb;
            const float* dl = lse + BW_BQ;
            const bool diag = p.causal && (q0 < k0 + BW_BK);      // some (key > query) pairs in this tile
            uint8_t* pDS = sDS + (t & 1) * Cfg::PS_BYTES + row * 128;
            float pv[32];
            {   // phase A: P^T = exp2(S^T * c - LSE) for this thread's key row and 32 query columns
                uint32_t rs[32];
                mbar_wait(&s_full[t & 1], (t >> 1) & 1);
                tc_fence_after();
                if (warp == 4 && lane == 0) BW_TS(1);
                tmem_ld32(tb + Cfg::T_S + (t & 1) * BW_BQ + cb, rs);
                tmem_ld_wait();
#pragma unroll
                for (int c4 = 0; c4 < 8; ++c4) {
                    const float4 l4 = *reinterpret_cast<const float4*>(lse + c4 * 4);
                    pv[c4 * 4 + 0] = exp2f(fmaf(__uint_as_float(rs[c4 * 4 + 0]), p.scale_log2, -l4.x));
                    pv[c4 * 4 + 1] = exp2f(fmaf(__uint_as_float(rs[c4 * 4 + 1]), p.scale_log2, -l4.y));
                    pv[c4 * 4 + 2] = exp2f(fmaf(__uint_as_float(rs[c4 * 4 + 2]), p.scale_log2, -l4.z));
                    pv[c4 * 4 + 3] = exp2f(fmaf(__uint_as_float(rs[c4 * 4 + 3]), p.scale_log2, -l4.w));
                }
                if (diag) {                                  // CTA-uniform: a real branch, diagonal tiles only
                    asm volatile("" ::: "memory");
                    const int nz = key - (q0 + cb);          // query columns c < nz lie above the diagonal (key > query)
#pragma unroll
                    for (int c = 0; c < 32; ++c) pv[c] = c < nz ? 0.f : pv[c];
                }
                uint32_t pp[16];
#pragma unroll
                for (int c = 0; c < 16; ++c) pp[c] = pack_bf16(pv[2 * c], pv[2 * c + 1]);
                tmem_st16(tb + Cfg::T_S + (t & 1) * BW_BQ + cb, pp);      // over this thread's own (already read) S columns
            }
            if (warp == 4 && lane == 0) BW_TS(2);
            {   // phase B: dS^T = P^T o (dP^T - Delta)
                uint32_t rp[32];
                if (t > 1) mbar_wait(&mma_done[t & 1], ((t - 2) >> 1) & 1);      // dS smem buffer t & 1: dK/dQ^T(t-2) have read it
                mbar_wait(dp_full, t & 1);
                tc_fence_after();
                if (warp == 4 && lane == 0) BW_TS(3);
                tmem_ld32(tb + Cfg::T_DP + cb, rp);
                tmem_ld_wait();
                uint32_t pd[16];
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    float d[8];
                    const float4 d0 = *reinterpret_cast<const float4*>(dl + c8 * 8), d1 = *reinterpret_cast<const float4*>(dl + c8 * 8 + 4);
                    const float dlv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
                    for (int e = 0; e < 8; ++e) d[e] = pv[c8 * 8 + e] * fmaf(__uint_as_float(rp[c8 * 8 + e]), p.scale, -dlv[e]);   // scale * dS
                    uint4 u;
                    u.x = pack_bf16(d[0], d[1]); u.y = pack_bf16(d[2], d[3]); u.z = pack_bf16(d[4], d[5]); u.w = pack_bf16(d[6], d[7]);
                    *reinterpret_cast<uint4*>(pDS + (((wg * 4 + c8) ^ (row & 7)) << 4)) = u;      // B operand of dQ^T
                    pd[c8 * 4] = u.x; pd[c8 * 4 + 1] = u.y; pd[c8 * 4 + 2] = u.z; pd[c8 * 4 + 3] = u.w;
                }
                tmem_st16(tb + Cfg::T_S + (t & 1) * BW_BQ + cb + 16, pd);   // A operand of dK
                tmem_st_wait();
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(pds_ready);
            if (warp == 4 && lane == 0) BW_TS(4);
            if (++qi == per_head) qi = 0;
            if (++st == BW_NST) { st = 0; st_ph ^= 1; }
        }
        mbar_wait(&mma_done[(T - 1) & 1], ((T - 1) >> 1) & 1);      // every MMA of this CTA has retired
        tc_fence_after();
        // ---- epilogue: dV_j, dK_j (the softmax scale is already inside dS^T) -> bf16 -> global rows [b*S + key], head hk; each warpgroup 64 columns
        bf16* dvr = p.dv + ((size_t)b * p.S + key) * p.ld_dkv + hk * BW_D;
        bf16* dkr = p.dk + ((size_t)b * p.S + key) * p.ld_dkv + hk * BW_D;
#pragma unroll 1
        for (int c0 = wg * 64; c0 < wg * 64 + 64; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tb + Cfg::T_DV + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int c8 = 0; c8 < 4; ++c8) {
                uint4 u;
                u.x = pack_bf16(__uint_as_float(r[c8 * 8 + 0]), __uint_as_float(r[c8 * 8 + 1]));
                u.y = pack_bf16(__uint_as_float(r[c8 * 8 + 2]), __uint_as_float(r[c8 * 8 + 3]));
                u.z = pack_bf16(__uint_as_float(r[c8 * 8 + 4]), __uint_as_float(r[c8 * 8 + 5]));
                u.w = pack_bf16(__uint_as_float(r[c8 * 8 + 6]), __uint_as_float(r[c8 * 8 + 7]));
                *reinterpret_cast<uint4*>(dvr + c0 + c8 * 8) = u;
            }
            tmem_ld32(tb + Cfg::T_DK + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int c8 = 0; c8 < 4; ++c8) {
                uint4 u;
                u.x = pack_bf16(__uint_as_float(r[c8 * 8 + 0]), __uint_as_float(r[c8 * 8 + 1]));
                u.y = pack_bf16(__uint_as_float(r[c8 * 8 + 2]), __uint_as_float(r[c8 * 8 + 3]));
                u.z = pack_bf16(__uint_as_float(r[c8 * 8 + 4]), __uint_as_float(r[c8 * 8 + 5]));
                u.w = pack_bf16(__uint_as_float(r[c8 * 8 + 6]), __uint_as_float(r[c8 * 8 + 7]));
                *reinterpret_cast<uint4*>(dkr + c0 + c8 * 8) = u;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc<Cfg::TMEM_COLS>(tmem_base); }
}

static long long* bw_ts_buffer() {
    static long long* buf = [] {
        const char* e = getenv("VLB_ATT_TS");
        long long* b = nullptr;
        if (e && e[0] == '1') { cudaMalloc(&b, 128 * 16 * sizeof(long long)); cudaMemset(b, 0, 128 * 16 * sizeof(long long)); }
        return b;
    }();
    return buf;
}

// Delta[b,h,s] = scale * sum_d dO[b,s,h,d] * O[b,s,h,d]; 16 lanes per (row, head), 8 elements (one 16-byte load of each operand)
// per lane, so a warp covers two heads of a row = 512 contiguous bytes of o and of dO
__global__ void attn_delta_kernel(const bf16* __restrict__ o, const bf16* __restrict__ dO, int ld, float* __restrict__ delta,
                                  long long rows, int S, int Hq, float scale) {
    const long long g = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 4;      // (row, head) index
    const int sub = threadIdx.x & 15;
    const bool ok = g < rows * Hq;
    const long long r = ok ? g / Hq : 0; const int h = ok ? (int)(g % Hq) : 0;
    const size_t off = (size_t)r * ld + h * BW_D + sub * 8;
    float a[8], d[8];
    load8(o + off, a); load8(dO + off, d);
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) s = fmaf(a[e], d[e], s);
#pragma unroll
    for (int m = 8; m > 0; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
    if (ok && sub == 0) delta[((size_t)(r / S) * Hq + h) * S + (r % S)] = s * scale;      // pre-scaled: dS^T leaves phase B as scale * dS^T
}

static int bw_tmap(CUtensorMap* m, const void* ptr, int cols, int S, int B, int ld, int box_rows) {
    unsigned long long dims[3] = {(unsigned long long)cols, (unsigned long long)S, (unsigned long long)B};
    unsigned long long strides[2] = {(unsigned long long)ld * 2ull, (unsigned long long)S * ld * 2ull};
    unsigned int box[3] = {64u, (unsigned int)box_rows, 1u};
    return encode_tmap_bf16(m, ptr, 3, dims, strides, box);
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_attn_bwd_debug_ts(long long* host_out, int n) {
    long long* b = vlb::bw_ts_buffer();
    VLB_REQUIRE(b != nullptr && n <= 128 * 16, "attn_bwd_debug_ts: set VLB_ATT_TS=1 (n=%d)", n);
    return cudaMemcpy(host_out, b, n * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1;
}

extern "C" int vlb_attn_bwd(const void* q, const void* k, const void* v, int ld_qkv, const void* o, const void* d_o,
                            int ld_o, const void* lse_f32, void* delta_f32, void* dq_f32, void* dk, void* dv, int ld_dkv,
                            int B, int S, int Hq, int Hkv, int D, int causal, float scale, void* stream) {
    VLB_REQUIRE(D == BW_D, "attn_bwd: only head_dim 128 (decoder) is implemented, got %d", D);
    VLB_REQUIRE(Hkv > 0 && Hq % Hkv == 0 && S % BW_BK == 0, "attn_bwd: Hq=%d Hkv=%d S=%d (S must be a multiple of 128)",
                Hq, Hkv, S);
    cudaStream_t st = (cudaStream_t)stream;
    const long long rows = (long long)B * S;
    {
        const long long threads = rows * Hq * 16;
        attn_delta_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>((const bf16*)o, (const bf16*)d_o, ld_o,
                                                                             (float*)delta_f32, rows, S, Hq, scale);
        int rc = check_launch("attn_delta_kernel");
        if (rc) return rc;
    }
    cudaError_t e = cudaMemsetAsync(dq_f32, 0, (size_t)rows * Hq * D * sizeof(float), st);
    VLB_REQUIRE(e == cudaSuccess, "attn_bwd memset: %s", cudaGetErrorString(e));
    CUtensorMap tq, tk, tv, tdo;
    int rc;
    if ((rc = bw_tmap(&tq, q, Hq * D, S, B, ld_qkv, BW_BQ))) return rc;
    if ((rc = bw_tmap(&tk, k, Hkv * D, S, B, ld_qkv, BW_BK))) return rc;
    if ((rc = bw_tmap(&tv, v, Hkv * D, S, B, ld_qkv, BW_BK))) return rc;
    if ((rc = bw_tmap(&tdo, d_o, Hq * D, S, B, ld_o, BW_BQ))) return rc;
    CUtensorMap tdq;
    {   // dq fp32 [B*S, Hq*D]: box = one dQ tile (64 query rows x 128 d), plain layout, target of the TMA reduce-add
        unsigned long long dims[2] = {(unsigned long long)Hq * D, (unsigned long long)B * S};
        unsigned long long strides[1] = {(unsigned long long)Hq * D * sizeof(float)};
        unsigned int box[2] = {(unsigned)BW_D, (unsigned)BW_BQ};
        if ((rc = encode_tmap_f32_plain(&tdq, dq_f32, 2, dims, strides, box))) return rc;
    }
    BwParams p;
    p.S = S; p.Hq = Hq; p.Hkv = Hkv; p.group = Hq / Hkv; p.causal = causal;
    p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
    p.lse = (const float*)lse_f32; p.delta = (const float*)delta_f32; p.dq = (float*)dq_f32;
    p.dk = (bf16*)dk; p.dv = (bf16*)dv; p.ld_dkv = ld_dkv;
    { const char* d = getenv("VLB_BW_DBG"); p.dbg = d ? atoi(d) : 0; }
    p.ts = bw_ts_buffer();
    { const char* c = getenv("VLB_ATT_TS_CTA"); p.ts_cta = c ? atoi(c) : 0; }
    static bool attr = false;
    if (!attr) {
        e = cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BwCfg::SMEM_BYTES);
        VLB_REQUIRE(e == cudaSuccess, "attn_bwd smem attr: %s", cudaGetErrorString(e));
        attr = true;
    }
    p.nkb = S / BW_BK; p.nbh = Hkv * B;
    // two batch elements per dispatch chunk: their fp32 dQ rows (2 x 32 heads x S x 128 x 4 B = 67 MB at S=2048) stay in
    // the 126 MB L2 while all key blocks reduce into them; measured at B=8: chunk 1/2/4/8 -> 1.078/1.065/1.071/1.118 ms
    p.chunk = B < 2 ? B : 2;
    dim3 grid((unsigned)(p.nkb * p.nbh));
    attn_bwd_kernel<<<grid, BW_THREADS, BwCfg::SMEM_BYTES, st>>>(p, tq, tk, tv, tdo, tdq);
    return check_launch("attn_bwd_kernel");
}
