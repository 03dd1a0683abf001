// CTA-pair (cta_group::2) variant of the tcgen05 GEMM for the big decoder / ViT / STC linears (bf16 output, K-major A).
// Included by gemm_sm100.cu after the single-CTA kernel (shares GemmParams, tile_coords, BM, BK, CD_STAGE_BYTES).
//
// A cluster of two CTAs (one TPC) owns a 256 x 256 output tile: CTA r holds rows [m0 + 128 r, +128) of A and columns
// [n0 + 128 r, +128) of B in its shared memory, so each k-block costs 32 KB of smem traffic per SM instead of 48 KB and
// the ring is 6 deep instead of 4. The leader's single MMA thread issues tcgen05.mma.cta_group::2 (M = 256, N = 256,
// K = 16); the fp32 accumulator rows land in each CTA's own TMEM (2 stages x 256 columns) and each CTA runs its own
// epilogue on its 128 rows. Barrier protocol:
//   full[s]   (leader only)  1 arrival (leader producer's expect_tx of BOTH CTAs' bytes) + TMA complete_tx from both CTAs
//   empty[s]  (both)         tcgen05.commit multicast from the leader's MMA thread
//   tfull[a]  (both)         tcgen05.commit multicast
//   tempty[a] (leader only)  one arrival per epilogue thread of BOTH CTAs (512 in the plain variant, 256 in the SwiGLU ones; remote mbarrier.arrive from the peer)
#pragma once

namespace vlb {

constexpr int G2_BN = 256;                 // pair tile columns
constexpr int G2_HALF_N = G2_BN / 2;       // B columns held by each CTA
constexpr int G2_A_BYTES = BM * BK * 2;    // 16 KB
constexpr int G2_B_BYTES = G2_HALF_N * BK * 2;   // 16 KB
constexpr int G2_STAGE_BYTES = G2_A_BYTES + G2_B_BYTES;
constexpr int G2_STAGES = SMEM_AB_BYTES / G2_STAGE_BYTES;    // 6
constexpr int G2_SMEM_BYTES = G2_STAGES * G2_STAGE_BYTES + 2 * CD_STAGE_BYTES + 1024 + 256;
constexpr int G2_TMEM_COLS = 512;

// SWIGLU (forward gate/up projection, B = [gate rows | up rows] of length I each): tile column block j pairs gate features
// [128 j, +128) (leader's B half) with the SAME up features (peer's B half, rows I + 128 j), so every epilogue thread holds g
// and u of 128 features of its row: it writes gu in the standard [M, 2I] layout (two column offsets) AND act = silu(g) * u
// through tmAct -- the separate SwiGLU pass (read 2I, write I per token) disappears.
// SWIGLU == 2 (backward of the down projection, N = I): the accumulator is d(act); the epilogue reads gate/up of its row
// from gu [M, 2I] and writes dgu = [d(act) up silu'(gate) | d(act) silu(gate)] at column offsets n and I + n (I % 256 == 0),
// so d(act) never goes to memory and the separate SwiGLU backward pass (read 3I, write 2I per token) disappears.
// The plain variant (SWIGLU == 0) runs EIGHT epilogue warps (384 threads): two warpgroups, each owning 128 of the tile's 256
// columns with its own staging buffer, named barrier and TMA-store issuer. With four warps the K = 1024 ViT GEMMs whose
// epilogue carries an activation or a residual were epilogue-bound (fc1 650, o-proj 500 TF/s against 1350 plain).
template <int SWIGLU, bool LORA = false> __host__ __device__ constexpr int g2_threads() { return (SWIGLU == 0 || SWIGLU == 3 || (SWIGLU == 2 && LORA)) ? 384 : 256; }

// LORA_EPI (dX GEMMs under LoRA dropout): the adapter branch of the input gradient, 1/(1-p) * keep_a . (t_a A_a), cannot ride in
// the accumulator as a k-block (the keep mask sits between the rank-16 product and the sum), so the epilogue forms it beside the
// tcgen05 main loop with warp-level MMAs: per 8 columns and adapter a warp computes P[32 rows, 8] = T_a[32, 16] A_a[16, 8] with
// two mma.sync m16n8k16 (T fragments in registers for the whole tile, the B fragment one conflict-free LDS.64 from A_a's columns
// staged in the ring stage this mode leaves unused), hands P from the MMA fragment layout to one-row-per-lane (the TMEM layout)
// through a 512-byte warp-private buffer, and adds the kept elements. An earlier version did the product on the CUDA cores from
// broadcast LDS.128 (16 loads = 64 shared-memory wavefronts + 256 issue slots per 8 columns and adapter): with the main loop
// already using ~3/4 of the shared-memory bandwidth (TMA writes + operand reads) that epilogue slowed the MMAs down.
//   v[i] += scale * keep_i * sum_j t[j] A[j][col + i]
constexpr int LA_REC_BYTES = 32;                          // one column of one adapter: 16 bf16 in B-fragment order
constexpr int LA_XBUF_BYTES = 1024;                       // per warp: two 32 x 16 B hand-off buffers
__device__ __forceinline__ void lora_rank16_add8(float (&v)[8], const uint32_t (&tf)[8], uint32_t brec, uint32_t xbuf, int lane,
                                                 unsigned int byte, float scale) {
    uint32_t b0, b1;
    lds64(brec, b0, b1);
    const uint32_t ta[4] = {tf[0], tf[1], tf[2], tf[3]}, tb[4] = {tf[4], tf[5], tf[6], tf[7]};
    float d0[4], d1[4];
    mma16816_bf16(d0, ta, b0, b1);                        // rows g, g + 8 of the warp's 32
    mma16816_bf16(d1, tb, b0, b1);                        // rows 16 + g, 24 + g
    const uint32_t w = xbuf + (uint32_t)(lane >> 2) * 16 + (uint32_t)(lane & 3) * 4;
    sts32(w, pack_bf16(d0[0], d0[1]));
    sts32(w + 8 * 16, pack_bf16(d0[2], d0[3]));
    sts32(w + 16 * 16, pack_bf16(d1[0], d1[1]));
    sts32(w + 24 * 16, pack_bf16(d1[2], d1[3]));
    __syncwarp();
    float acc[8];
    unpack8(lds128_sync(xbuf + (uint32_t)lane * 16), acc);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += ((byte >> i) & 1u) ? acc[i] * scale : 0.f;
}

// Same product with the B fragment already in registers and the hand-off through the 16-byte slots of the output staging buffer
// that this chunk's results are about to overwrite anyway (row r, chunk ch -> r * 128 + ((ch ^ (r & 7)) << 4), the 128B-swizzle
// pattern): no shared memory of its own, which is what lets the SwiGLU-backward variant keep a 5-deep ring.
__device__ __forceinline__ void lora_rank16_add8_slots(float (&v)[8], const uint32_t (&tf)[8], uint32_t b0, uint32_t b1, uint32_t cd,
                                                       int q, int lane, int ch, unsigned int byte, float scale) {
    const uint32_t ta[4] = {tf[0], tf[1], tf[2], tf[3]}, tb[4] = {tf[4], tf[5], tf[6], tf[7]};
    float d0[4], d1[4];
    mma16816_bf16(d0, ta, b0, b1);
    mma16816_bf16(d1, tb, b0, b1);
    const int g = lane >> 2;
    const uint32_t w = cd + (uint32_t)(q * 32 + g) * 128 + (uint32_t)((ch ^ g) << 4) + (uint32_t)(lane & 3) * 4;   // (row & 7) == g for all four rows
    sts32(w, pack_bf16(d0[0], d0[1]));
    sts32(w + 8 * 128, pack_bf16(d0[2], d0[3]));
    sts32(w + 16 * 128, pack_bf16(d1[0], d1[1]));
    sts32(w + 24 * 128, pack_bf16(d1[2], d1[3]));
    __syncwarp();
    const int row = q * 32 + lane;
    float acc[8];
    unpack8(lds128_sync(cd + (uint32_t)row * 128 + (uint32_t)((ch ^ (row & 7)) << 4)), acc);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += ((byte >> i) & 1u) ? acc[i] * scale : 0.f;
}

// B fragment (k = 2c, 2c+1 | 2c+8, 2c+9 of column `col`) of one adapter straight from A [16, N] in global memory
__device__ __forceinline__ void lora_bfrag_ldg(const unsigned short* __restrict__ A, int lda, int col, int c, uint32_t& b0, uint32_t& b1) {
    const unsigned short* a = A + (size_t)(2 * c) * lda + col;
    b0 = (uint32_t)__ldg(a) | ((uint32_t)__ldg(a + lda) << 16);
    b1 = (uint32_t)__ldg(a + 8 * (size_t)lda) | ((uint32_t)__ldg(a + 9 * (size_t)lda) << 16);
}

template <bool B_MN, int SWIGLU = 0, bool LORA = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(g2_threads<SWIGLU, LORA>(), 1)
gemm2_tcgen05_kernel(const GemmParams p,
                     const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                     const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                     const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAct) {
    constexpr int STAGES = G2_STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * G2_A_BYTES;
    uint8_t* smem_cd = smem + STAGES * G2_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_cd + 2 * CD_STAGE_BYTES);
    uint64_t* full_bar = bars;                    // [STAGES]
    uint64_t* empty_bar = bars + STAGES;          // [STAGES]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]
    uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    const int nk = p.nk1 + p.nk2;
    const int num_tiles = p.mt * p.nt;            // p.mt counts 256-row blocks here

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA1); tma_prefetch_desc(&tmB1); tma_prefetch_desc(&tmC);
        if (p.nk2 > 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 2 * (g2_threads<SWIGLU, LORA>() - 128)); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_2sm<G2_TMEM_COLS>(tmem_ptr);
    tc_fence_before();
    cluster_sync_all();                            // barriers of BOTH CTAs are initialised before any remote arrive / TMA
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
                int mb, nb; tile_coords(t, p.mt, p.nt, p.group_m, mb, nb, p.serp);
                const int m0 = mb * (2 * BM) + (int)rank * BM;
                const int n0 = SWIGLU == 1 ? nb * G2_HALF_N + (int)rank * p.swiglu_I : nb * G2_BN + (int)rank * G2_HALF_N;
                const bool rev = p.serp && (it & 1);              // every CTA of a wave agrees on `it`, so a wave still walks K in step
                for (int kq = 0; kq < nk; ++kq) {
                    const int kb = rev ? nk - 1 - kq : kq;        // the MMA issuer consumes ring stages in order, whatever k they hold
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * G2_STAGE_BYTES);
                    const bool second = kb >= p.nk1;
                    const CUtensorMap* ta = second ? &tmA2 : &tmA1;
                    const CUtensorMap* tb = second ? &tmB2 : &tmB1;
                    const int k0 = (second ? kb - p.nk1 : kb) * BK;
                    uint8_t* sa = smem_a + stage * G2_A_BYTES;
                    uint8_t* sb = smem_b + stage * G2_B_BYTES;
                    tma_load_2d_2sm(ta, &full_bar[stage], sa, k0, m0);
                    if constexpr (!B_MN) {
                        tma_load_2d_2sm(tb, &full_bar[stage], sb, k0, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < G2_HALF_N / 64; ++j)
                            tma_load_2d_2sm(tb, &full_bar[stage], sb + j * (BK * 128), n0 + j * 64, k0);
                    }
                    if (++stage == p.stages_used) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA, one thread)
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(2 * BM, G2_BN, false, B_MN);
            constexpr uint32_t LBO = BK * 128;
            constexpr uint32_t KSTEP_B = B_MN ? (16 * 128) : 32;
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tempty_bar[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * G2_BN;
                for (int kb = 0; kb < nk; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * G2_A_BYTES);
                    const uint32_t b_addr = smem_u32(smem_b + stage * G2_B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t ad = make_smem_desc(a_addr + k * 32, 0, 1024);
                        const uint64_t bd = make_smem_desc(b_addr + k * KSTEP_B, B_MN ? LBO : 0, 1024);
                        umma_bf16_2sm(d_tmem, ad, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit_2sm(&empty_bar[stage]);
                    if (kb == nk - 1) umma_commit_2sm(&tfull_bar[as]);
                    if (++stage == p.stages_used) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue (128 threads per CTA, own 128 rows)
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int wg = (warp - 4) >> 2;                     // epilogue warpgroup (always 0 in the 256-thread SwiGLU variants)
        int it = 0;
        const bool issuer = (q == 0 && lane == 0);          // one TMA-store issuer per warpgroup
        for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
            int mb, nb; tile_coords(t, p.mt, p.nt, p.group_m, mb, nb, p.serp);
            const int m0 = mb * (2 * BM) + (int)rank * BM, n0 = nb * G2_BN;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            if constexpr (SWIGLU == 2) {                    // pull this row's gate/up segments towards L2 while the MMAs run
                if (m0 + row < p.M) {
                    const bf16* gr = p.sw_gu + (size_t)(m0 + row) * p.sw_ld + n0;
#pragma unroll
                    for (int c = 0; c < G2_BN; c += 64) {
                        asm volatile("prefetch.global.L2 [%0];" :: "l"(gr + c));
                        asm volatile("prefetch.global.L2 [%0];" :: "l"(gr + p.swiglu_I + c));
                    }
                }
            }
            const int m = m0 + row;
            const bool m_ok = m < p.M;
            constexpr int LA_COLS = G2_BN / 2;                           // tile columns one epilogue warpgroup owns
            constexpr int LA_AD_BYTES = LA_COLS * LA_REC_BYTES;          // 4 KB per adapter and warpgroup
            uint32_t tf[LORA ? 3 : 1][8];                   // mma.sync A fragments of this warp's 32 rows of t, per adapter
            uint32_t la = 0;                                // this warpgroup's staged columns of A (shared-space address)
            uint32_t xb = 0;                                // this warp's hand-off buffers
            uint32_t xsel = 0;
            if constexpr (LORA) {
                // every thread of the warpgroup is past the last bar.sync of its previous tile, so nobody still reads the old A.
                // Both variants run a 5-deep ring and use ring stage 5. Plain: its A / B slot hold the two warpgroups' regions
                // [adapter][column of its 128][32 B record] + 4 x 1 KB hand-off buffers. SwiGLU backward: the two slots are the
                // second warpgroup's output staging buffers; its single adapter's B fragments come straight from global memory
                // and the hand-off goes through the staging slots (lora_rank16_add8_slots), so nothing is staged here.
                uint8_t* law = (wg == 0 ? smem_a : smem_b) + (STAGES - 1) * G2_A_BYTES;
                la = smem_u32(law);
                xb = la + 3 * LA_AD_BYTES + q * LA_XBUF_BYTES;
                const unsigned short* Ag = reinterpret_cast<const unsigned short*>(p.lr_A);
                for (int idx = q * 32 + lane; SWIGLU != 2 && idx < p.lr_na * LA_COLS; idx += 128) {      // one column of one adapter per thread
                    const int a = idx / LA_COLS, tid = idx - a * LA_COLS;
                    const int ncol = n0 + wg * LA_COLS + tid;
                    uint32_t rec[8];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {           // record order: k = 2c, 2c+1, 2c+8, 2c+9
                        uint32_t kv[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int k = 2 * c + (e & 1) + (e >> 1) * 8;
                            kv[e] = ncol < p.N ? (uint32_t)__ldg(Ag + (size_t)(a * 16 + k) * p.lr_lda + ncol) : 0u;
                        }
                        rec[2 * c] = kv[0] | (kv[1] << 16);
                        rec[2 * c + 1] = kv[2] | (kv[3] << 16);
                    }
                    const uint32_t dst = la + a * LA_AD_BYTES + tid * LA_REC_BYTES;
                    sts128(dst, make_uint4(rec[0], rec[1], rec[2], rec[3]));
                    sts128(dst + 16, make_uint4(rec[4], rec[5], rec[6], rec[7]));
                }
                const int g = lane >> 2, c2 = (lane & 3) * 2;
                const uint32_t* tg = reinterpret_cast<const uint32_t*>(p.lr_t);
#pragma unroll
                for (int a = 0; a < 3; ++a)
#pragma unroll
                    for (int e = 0; e < 8; ++e) {           // e = 4 h + 2 (k half) + (row half)
                        const int r = m0 + q * 32 + (e >> 2) * 16 + g + (e & 1) * 8;
                        const int k = a * 16 + c2 + ((e >> 1) & 1) * 8;
                        tf[a][e] = (a < p.lr_na && r < p.M) ? __ldg(tg + ((size_t)r * p.lr_ldt + k) / 2) : 0u;
                    }
            }
            mbar_wait(&tfull_bar[as], aphase);
            tc_fence_after();
            if constexpr (SWIGLU == 2 && LORA) {
                // Two warpgroups (384 threads), each owning 128 of the tile's 256 columns with its own pair of staging buffers:
                // with four warps (one per scheduler, latency-bound) the SwiGLU backward plus the adapter term made the epilogue
                // longer than the K = 4096 main loop (ncu: tensor pipe 54 % active, against 86 % for the variant without adapters). Rolled over 8-column chunks: unrolled, this code overflowed the
                // instruction cache and the warps stalled on instruction fetch.
                uint8_t* cdg = wg == 0 ? smem_cd : smem_a + (STAGES - 1) * G2_A_BYTES;
                uint8_t* cdu = wg == 0 ? smem_cd + CD_STAGE_BYTES : smem_b + (STAGES - 1) * G2_B_BYTES;
                const unsigned short* Ag = reinterpret_cast<const unsigned short*>(p.lr_A);
                uint32_t b0n, b1n;
                lora_bfrag_ldg(Ag, p.lr_lda, n0 + wg * (G2_BN / 2) + (lane >> 2), lane & 3, b0n, b1n);
                const int I = p.swiglu_I;
                const bf16* gurow = p.sw_gu + (size_t)(m_ok ? m : 0) * p.sw_ld;
                const int s_first = wg * (G2_BN / 128);
                uint4 gvn = __ldg(reinterpret_cast<const uint4*>(gurow + n0 + s_first * 64));
                uint4 uvn = __ldg(reinterpret_cast<const uint4*>(gurow + I + n0 + s_first * 64));
                unsigned long long kwn = ~0ull;
                if (m_ok) kwn = __ldg(p.lr_bits[0] + (size_t)((n0 >> 6) + s_first) * p.M + m);
#pragma unroll 1
                for (int sl = 0; sl < G2_BN / 128; ++sl) {
                    const int s = s_first + sl;
                    const unsigned long long kw = kwn;
                    if (m_ok && sl + 1 < G2_BN / 128) kwn = __ldg(p.lr_bits[0] + (size_t)((n0 >> 6) + s + 1) * p.M + m);
                    if (issuer) tma_store_wait_read<0>();
                    asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
#pragma unroll 1
                    for (int ch = 0; ch < 8; ++ch) {
                        const int c0 = s * 64 + ch * 8;
                        const uint4 gv = gvn, uv = uvn;
                        const uint32_t b0 = b0n, b1 = b1n;
                        {   // gate / up and the adapter's B fragment of the NEXT chunk: the global round trip overlaps this chunk's math
                            const int cn = s_first * 64 + ((sl * 64 + ch * 8 + 8) & (G2_BN / 2 - 1));
                            gvn = __ldg(reinterpret_cast<const uint4*>(gurow + n0 + cn));
                            uvn = __ldg(reinterpret_cast<const uint4*>(gurow + I + n0 + cn));
                            lora_bfrag_ldg(Ag, p.lr_lda, n0 + cn + (lane >> 2), lane & 3, b0n, b1n);
                        }
                        uint32_t r8[8];
                        tmem_ld8(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + c0), r8);
                        tmem_ld_wait();
                        float g[8], u[8], dg[8], du[8], dv[8];
                        unpack8(gv, g); unpack8(uv, u);
#pragma unroll
                        for (int j = 0; j < 8; ++j) dv[j] = __uint_as_float(r8[j]) * p.alpha;
                        lora_rank16_add8_slots(dv, tf[0], b0, b1, smem_u32(cdg), q, lane, ch, (unsigned int)(kw >> (ch * 8)) & 0xFFu, p.lr_scale);
#pragma unroll
                        for (int j = 0; j < 8; ++j) dv[j] = bf2f(f2bf(dv[j]));            // d(act) as the unfused path stores it
                        swiglu_bwd8(g, u, dv, dg, du);
                        uint4 vg, vu;
                        vg.x = pack_bf16(dg[0], dg[1]); vg.y = pack_bf16(dg[2], dg[3]); vg.z = pack_bf16(dg[4], dg[5]); vg.w = pack_bf16(dg[6], dg[7]);
                        vu.x = pack_bf16(du[0], du[1]); vu.y = pack_bf16(du[2], du[3]); vu.z = pack_bf16(du[4], du[5]); vu.w = pack_bf16(du[6], du[7]);
                        const int off = row * 128 + ((ch ^ (row & 7)) << 4);
                        sts128(smem_u32(cdg) + off, vg);
                        sts128(smem_u32(cdu) + off, vu);
                    }
                    if (sl == G2_BN / 128 - 1) {
                        tc_fence_before();
                        mbar_arrive_cluster(&tempty_bar[as], 0);
                    }
                    fence_proxy_async_smem();
                    asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
                    if (issuer) {
                        tma_store_2d(&tmC, cdg, n0 + s * 64, m0);
                        tma_store_2d(&tmC, cdu, I + n0 + s * 64, m0);
                        tma_store_commit();
                    }
                }
                continue;
            }
            if constexpr (SWIGLU == 2) {
                uint8_t* cdg = smem_cd;
                uint8_t* cdu = smem_cd + CD_STAGE_BYTES;
                const int I = p.swiglu_I;
                const bf16* gurow = p.sw_gu + (size_t)(m_ok ? m : 0) * p.sw_ld;
#pragma unroll 1
                for (int s = 0; s < G2_BN / 64; ++s) {
                    if (issuer) tma_store_wait_read<0>();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int c0 = s * 64 + h * 32;
                        const int n = n0 + c0;
                        uint4 gv[4], uv[4];
#pragma unroll
                        for (int c = 0; c < 4; ++c) {           // this row's gate/up values: issued before the TMEM wait
                            gv[c] = __ldg(reinterpret_cast<const uint4*>(gurow + n + c * 8));
                            uv[c] = __ldg(reinterpret_cast<const uint4*>(gurow + I + n + c * 8));
                        }
                        uint32_t r[32];
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + c0), r);
                        tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            float g[8], u[8], dg[8], du[8], dv[8];
                            unpack8(gv[c], g); unpack8(uv[c], u);
#pragma unroll
                            for (int j = 0; j < 8; ++j) dv[j] = __uint_as_float(r[c * 8 + j]);
                            scale_round_bf16_8(dv, p.alpha);                              // d(act) as the unfused path stores it
                            swiglu_bwd8(g, u, dv, dg, du);
                            uint4 vg, vu;
                            vg.x = pack_bf16(dg[0], dg[1]); vg.y = pack_bf16(dg[2], dg[3]); vg.z = pack_bf16(dg[4], dg[5]); vg.w = pack_bf16(dg[6], dg[7]);
                            vu.x = pack_bf16(du[0], du[1]); vu.y = pack_bf16(du[2], du[3]); vu.z = pack_bf16(du[4], du[5]); vu.w = pack_bf16(du[6], du[7]);
                            const int off = row * 128 + (((h * 4 + c) ^ (row & 7)) << 4);
                            sts128(smem_u32(cdg) + off, vg);
                            sts128(smem_u32(cdu) + off, vu);
                        }
                    }
                    if (s == G2_BN / 64 - 1) {
                        tc_fence_before();
                        mbar_arrive_cluster(&tempty_bar[as], 0);
                    }
                    fence_proxy_async_smem();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (issuer) {
                        tma_store_2d(&tmC, cdg, n0 + s * 64, m0);
                        tma_store_2d(&tmC, cdu, I + n0 + s * 64, m0);
                        tma_store_commit();
                    }
                }
                continue;
            }
            if constexpr (SWIGLU == 1) {
                // staging: cd0 = gate chunk, cd1 = up chunk, cd2 = act chunk (ring stage 5 is unused in this mode)
                uint8_t* cdg = smem_cd;
                uint8_t* cdu = smem_cd + CD_STAGE_BYTES;
                uint8_t* cda = smem_a + (STAGES - 1) * G2_A_BYTES;
                const int f0 = nb * G2_HALF_N;                  // first gate/up feature of this tile
#pragma unroll 1
                for (int s = 0; s < G2_HALF_N / 64; ++s) {
                    if (issuer) tma_store_wait_read<0>();       // the three staging buffers have been read out
                    asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t rg[32], ru[32];
                        const int c0 = s * 64 + h * 32;
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + c0), rg);
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + G2_HALF_N + c0), ru);
                        tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            float g[8], u[8], a[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                g[j] = bf2f(f2bf(__uint_as_float(rg[c * 8 + j]) * p.alpha));
                                u[j] = bf2f(f2bf(__uint_as_float(ru[c * 8 + j]) * p.alpha));
                                a[j] = bf2f(f2bf(g[j] / (1.f + __expf(-g[j])))) * u[j];      // same rounding points as swiglu_fwd
                            }
                            uint4 vg, vu, va;
                            vg.x = pack_bf16(g[0], g[1]); vg.y = pack_bf16(g[2], g[3]); vg.z = pack_bf16(g[4], g[5]); vg.w = pack_bf16(g[6], g[7]);
                            vu.x = pack_bf16(u[0], u[1]); vu.y = pack_bf16(u[2], u[3]); vu.z = pack_bf16(u[4], u[5]); vu.w = pack_bf16(u[6], u[7]);
                            va.x = pack_bf16(a[0], a[1]); va.y = pack_bf16(a[2], a[3]); va.z = pack_bf16(a[4], a[5]); va.w = pack_bf16(a[6], a[7]);
                            const int off = row * 128 + (((h * 4 + c) ^ (row & 7)) << 4);
                            sts128(smem_u32(cdg) + off, vg);
                            sts128(smem_u32(cdu) + off, vu);
                            sts128(smem_u32(cda) + off, va);
                        }
                    }
                    if (s == G2_HALF_N / 64 - 1) {
                        tc_fence_before();
                        mbar_arrive_cluster(&tempty_bar[as], 0);
                    }
                    fence_proxy_async_smem();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (issuer) {
                        tma_store_2d(&tmC, cdg, f0 + s * 64, m0);
                        tma_store_2d(&tmC, cdu, p.swiglu_I + f0 + s * 64, m0);
                        tma_store_2d(&tmAct, cda, f0 + s * 64, m0);
                        tma_store_commit();
                    }
                }
                continue;
            }
            if constexpr (SWIGLU == 3) {
                // RoPE (q/k/v projection, head_dim 128): this warpgroup owns one head = two 64-column chunks, x1 = columns [0, 64) and
                // x2 = [64, 128) of the head; out1 = x1 cos - x2 sin, out2 = x2 cos + x1 sin with the stand-alone kernel's rounding
                // (projection -> bf16, products -> bf16, sum -> bf16). Heads at or beyond rope_ncols (v) pass through.
                uint8_t* cd0 = wg == 0 ? smem_cd : smem_a + (STAGES - 1) * G2_A_BYTES;
                uint8_t* cd1 = wg == 0 ? smem_cd + CD_STAGE_BYTES : smem_b + (STAGES - 1) * G2_B_BYTES;
                const int nh0 = n0 + wg * (G2_BN / 2);
                const bool rot = nh0 < p.rope_ncols;
                const size_t pos = m_ok ? (size_t)(m % p.rope_S) : 0;
                const bf16* cr = p.rope_cos + pos * 64;
                const bf16* sr = p.rope_sin + pos * 64;
                if (issuer) tma_store_wait_read<0>();
                asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint4 cv[4], sv[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {               // requested before the TMEM wait
                        cv[c] = __ldg(reinterpret_cast<const uint4*>(cr + h * 32 + c * 8));
                        sv[c] = __ldg(reinterpret_cast<const uint4*>(sr + h * 32 + c * 8));
                    }
                    uint32_t ra[32], rb[32];
                    const uint32_t tcol = (uint32_t)(as * G2_BN + wg * (G2_BN / 2) + h * 32);
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + tcol, ra);
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + tcol + 64, rb);
                    tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        float cs[8], sn[8], oa[8], ob[8];
                        unpack8(cv[c], cs); unpack8(sv[c], sn);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float a = bf2f(f2bf(__uint_as_float(ra[c * 8 + j]) * p.alpha));
                            const float b = bf2f(f2bf(__uint_as_float(rb[c * 8 + j]) * p.alpha));
                            oa[j] = rot ? bf2f(f2bf(a * cs[j])) + bf2f(f2bf(-b * sn[j])) : a;
                            ob[j] = rot ? bf2f(f2bf(b * cs[j])) + bf2f(f2bf(a * sn[j])) : b;
                        }
                        uint4 ua, ub;
                        ua.x = pack_bf16(oa[0], oa[1]); ua.y = pack_bf16(oa[2], oa[3]); ua.z = pack_bf16(oa[4], oa[5]); ua.w = pack_bf16(oa[6], oa[7]);
                        ub.x = pack_bf16(ob[0], ob[1]); ub.y = pack_bf16(ob[2], ob[3]); ub.z = pack_bf16(ob[4], ob[5]); ub.w = pack_bf16(ob[6], ob[7]);
                        const int off = row * 128 + (((h * 4 + c) ^ (row & 7)) << 4);
                        sts128(smem_u32(cd0) + off, ua);
                        sts128(smem_u32(cd1) + off, ub);
                    }
                }
                tc_fence_before();
                mbar_arrive_cluster(&tempty_bar[as], 0);
                fence_proxy_async_smem();
                asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
                if (issuer) {
                    tma_store_2d(&tmC, cd0, nh0, m0);
                    tma_store_2d(&tmC, cd1, nh0 + 64, m0);
                    tma_store_commit();
                }
                continue;
            }
            uint8_t* cd = smem_cd + wg * CD_STAGE_BYTES;      // this warpgroup's staging buffer
            if constexpr (LORA) {
                // rolled over 8-column chunks (see the SwiGLU variant): ~0.9 k instructions per chunk body instead of 7 k unrolled
#pragma unroll 1
                for (int sl = 0; sl < G2_BN / 128; ++sl) {
                    const int s = wg * (G2_BN / 128) + sl;
                    unsigned long long kw[3];
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        kw[a] = ~0ull;
                        if (a < p.lr_na && m_ok) kw[a] = __ldg(p.lr_bits[a] + (size_t)((n0 >> 6) + s) * p.M + m);
                    }
                    if (issuer) tma_store_wait_read<0>();
                    asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
#pragma unroll 1
                    for (int ch = 0; ch < 8; ++ch) {
                        uint32_t r8[8];
                        tmem_ld8(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + s * 64 + ch * 8), r8);
                        tmem_ld_wait();
                        float v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r8[j]) * p.alpha;
#pragma unroll
                        for (int a = 0; a < 3; ++a)
                            if (a < p.lr_na) {
                                lora_rank16_add8(v, tf[a], la + a * LA_AD_BYTES + (sl * 64 + ch * 8 + (lane >> 2)) * LA_REC_BYTES + (lane & 3) * 8,
                                                 xb + xsel, lane, (unsigned int)(kw[a] >> (ch * 8)) & 0xFFu, p.lr_scale);
                                xsel ^= 512u;
                            }
                        uint4 u;
                        u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]);
                        u.z = pack_bf16(v[4], v[5]); u.w = pack_bf16(v[6], v[7]);
                        sts128(smem_u32(cd) + row * 128 + ((ch ^ (row & 7)) << 4), u);
                    }
                    if (sl == G2_BN / 128 - 1) {
                        tc_fence_before();
                        mbar_arrive_cluster(&tempty_bar[as], 0);
                    }
                    fence_proxy_async_smem();
                    asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
                    if (issuer) {
                        tma_store_2d(&tmC, cd, n0 + s * 64, m0);
                        tma_store_commit();
                    }
                }
                continue;
            }
            if constexpr (!LORA)                              // plain epilogue (alpha, bias, activation, residual)
#pragma unroll 1
            for (int sl = 0; sl < G2_BN / 128; ++sl) {
                const int s = wg * (G2_BN / 128) + sl;        // 64-column chunk index inside the tile
                if (issuer) tma_store_wait_read<0>();
                asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t r[32];
                    const int c0 = s * 64 + h * 32;
                    const int n = n0 + c0;
                    // bias / residual for these 32 columns are requested BEFORE the TMEM wait: issued inside the per-chunk loop
                    // each load sat behind the previous chunk's math and exposed a full memory latency
                    uint4 bv[4], rv[4];
                    const bool has_bias = p.bias != nullptr, has_res = p.residual != nullptr && m_ok;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int nn = n + c * 8;
                        bv[c] = (has_bias && nn < p.N) ? __ldg(reinterpret_cast<const uint4*>(p.bias + nn)) : make_uint4(0u, 0u, 0u, 0u);
                        rv[c] = (has_res && nn < p.N) ? *reinterpret_cast<const uint4*>(p.residual + (size_t)m * p.ldr + nn)
                                                      : make_uint4(0u, 0u, 0u, 0u);
                    }
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * G2_BN + c0), r);
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
                        const int chunk = h * 4 + c;
                        sts128(smem_u32(cd) + row * 128 + ((chunk ^ (row & 7)) << 4), u);
                    }
                }
                if (sl == G2_BN / 128 - 1) {                // this thread's TMEM reads of the accumulator stage are done
                    tc_fence_before();
                    mbar_arrive_cluster(&tempty_bar[as], 0);
                }
                fence_proxy_async_smem();
                asm volatile("bar.sync %0, 128;" :: "r"(1 + wg) : "memory");
                if (issuer) {
                    tma_store_2d(&tmC, cd, n0 + s * 64, m0);
                    tma_store_commit();
                }
            }
        }
        if (issuer) tma_store_wait_all<0>();
    }

    tc_fence_before();
    cluster_sync_all();       // the peer's smem / TMEM stay alive until the leader's last MMA and remote arrive have landed
    if (warp == 2) { tc_fence_after(); tmem_dealloc_2sm<G2_TMEM_COLS>(tmem_base); }
}

}  // namespace vlb
