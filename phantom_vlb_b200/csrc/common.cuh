// Shared device/host helpers for libvlb_b200 (sm_100a only).
// Hand-written PTX wrappers for mbarrier / TMA / tcgen05 + small math utilities.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

namespace vlb {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int  check_launch(const char* what);   // returns 0 or cudaError_t and records message

#define VLB_REQUIRE(cond, ...)                               \
    do { if (!(cond)) { ::vlb::set_error(__VA_ARGS__); return 1; } } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }
int num_sms();

// bf16 tensor map of rank 2 or 3 (dims/box innermost first, strides in bytes for dims 1..rank-1), 128B swizzle.
// Implemented in gemm_sm100.cu (uncached; the GEMM keeps its own cache).
int encode_tmap_bf16(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* dims,
                     const unsigned long long* strides_bytes, const unsigned int* box);

int encode_tmap_f32_plain(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* dims,
                          const unsigned long long* strides_bytes, const unsigned int* box);

// ---------------------------------------------------------------- bf16 utils
typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

__device__ __forceinline__ float bf2f(bf16 v) { return __bfloat162float(v); }
// through the packed convert (cvt.rn.bf16x2.f32 -> F2FP on the ALU pipe): the scalar cvt.rn.bf16.f32 compiles to F2F, which shares
// the quarter-rate conversion/XU pipe -- the row kernels that emulate the reference's bf16 rounding points (3 roundings per element
// in hrf_pool) were bound by it. Same round-to-nearest-even result.
__device__ __forceinline__ bf16 f2bf(float v) { return __floats2bfloat162_rn(v, 0.f).x; }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    bf162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t u) {
    bf162 v = *reinterpret_cast<bf162*>(&u);
    return __bfloat1622float2(v);
}

// 8 bf16 <-> 8 floats through one 16-byte access
__device__ __forceinline__ void load8(const bf16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
    float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void store8(bf16* p, const float (&v)[8]) {
    uint4 u;
    u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]);
    u.z = pack_bf16(v[4], v[5]); u.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p)     = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; `red` must hold >= 32 floats of shared memory. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();               // protect `red` from a previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : 0.f;
    r = warp_sum(r);
    return r;
}

// Deterministic grid-wide sum into *acc: every block parks its partial, the LAST block to finish adds them in block order. A
// plain atomicAdd per block makes the low bits depend on the arrival order -- data-parallel replicas then compute clip
// coefficients that differ in the last place and their parameters drift apart bit by bit.
struct DetSumWs { float partial[4096]; unsigned int count; };
__device__ __forceinline__ void det_grid_add(DetSumWs& ws, float s /* valid in thread 0 */, float* acc, float* red) {
    __shared__ bool last;
    if (threadIdx.x == 0) {
        ws.partial[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(&ws.count, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    float t = 0.f;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) t += reinterpret_cast<volatile float*>(ws.partial)[i];
    t = block_sum(t, red);
    if (threadIdx.x == 0) { *acc += t; ws.count = 0; }
}

// ---------------------------------------------------------------- activations
// v[i] = float(bf16(s * v[i])) for 8 values: packed multiply, one packed convert per pair, the bf16 halves moved back up by shifts
__device__ __forceinline__ void scale_round_bf16_8(float (&v)[8], float s) {
    const float2 s2 = make_float2(s, s);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        const float2 m = __fmul2_rn(make_float2(v[j], v[j + 1]), s2);
        const uint32_t u = pack_bf16(m.x, m.y);
        v[j] = __uint_as_float(u << 16);
        v[j + 1] = __uint_as_float(u & 0xFFFF0000u);
    }
}
__device__ __forceinline__ float rcp_approx(float x) {           // what 1.f / x compiles to under --use_fast_math (MUFU.RCP)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// SwiGLU backward of 8 features, shared by the stand-alone kernel and the GEMM epilogues (which must agree bit for bit):
//   dg = (d u) (sg (1 + g (1 - sg))),  du = d (g sg),  sg = sigmoid(g), with the fp32 arithmetic issued as packed pairs (FMUL2 / FFMA2 /
// FADD2, sm_100): 6.5 instead of 11 arithmetic instructions per feature in epilogues that sit right at the length of a K = 4096 main loop.
__device__ __forceinline__ void swiglu_bwd8(const float (&g)[8], const float (&u)[8], const float (&d)[8], float (&dg)[8], float (&du)[8]) {
    const float2 one2 = make_float2(1.f, 1.f), m1 = make_float2(-1.f, -1.f), nl2e = make_float2(-1.4426950408889634f, -1.4426950408889634f);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        const float2 g2 = make_float2(g[j], g[j + 1]);
        const float2 a = __fmul2_rn(g2, nl2e);
        const float2 den = __fadd2_rn(make_float2(exp2f(a.x), exp2f(a.y)), one2);
        const float2 sg = make_float2(rcp_approx(den.x), rcp_approx(den.y));
        const float2 d2 = make_float2(d[j], d[j + 1]);
        const float2 du2 = __fmul2_rn(d2, __fmul2_rn(g2, sg));
        const float2 om = __ffma2_rn(sg, m1, one2);                   // 1 - sg
        const float2 t = __fmul2_rn(sg, __ffma2_rn(g2, om, one2));    // sg (1 + g (1 - sg))
        const float2 dg2 = __fmul2_rn(__fmul2_rn(d2, make_float2(u[j], u[j + 1])), t);
        dg[j] = dg2.x; dg[j + 1] = dg2.y; du[j] = du2.x; du[j + 1] = du2.y;
    }
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
__device__ __forceinline__ float siluf_(float x) { return x * sigmoidf_(x); }
// x * sigmoid(1.702 x) with sigmoid(t) = 0.5 + 0.5 tanh(t / 2): ONE MUFU op (tanh.approx, rel. error ~2^-11, far below the bf16
// rounding of the result) instead of ex2 + rcp -- the ViT fc1 epilogue (K = 1024) is MUFU-bound
__device__ __forceinline__ float tanh_approx_(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float quick_geluf_(float x) { return x * fmaf(0.5f, tanh_approx_(0.851f * x), 0.5f); }
__device__ __forceinline__ float gelu_erff_(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }

enum Act { ACT_NONE = 0, ACT_QUICK_GELU = 1, ACT_GELU = 2, ACT_SILU = 3, ACT_SIGMOID = 4 };
__device__ __forceinline__ float apply_act(float x, int act) {
    switch (act) {
        case ACT_QUICK_GELU: return quick_geluf_(x);
        case ACT_GELU:       return gelu_erff_(x);
        case ACT_SILU:       return siluf_(x);
        case ACT_SIGMOID:    return sigmoidf_(x);
        default:             return x;
    }
}

// Activation of 8 values with the (kernel-uniform) selector branched ONCE per group: the per-element switch above was
// if-converted by the compiler into "compute every variant (erff included), then select" -- the ViT fc1 GEMM (quick-GELU
// epilogue) ran at 200 TF/s instead of ~1000 because of it (r01d, tools/gemm_shapes_in_step.py).
__device__ __forceinline__ void apply_act8(float (&v)[8], int act) {
    if (act == ACT_QUICK_GELU) {
        asm volatile("" ::: "memory");
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = quick_geluf_(v[j]);
    } else if (act == ACT_SILU) {
        asm volatile("" ::: "memory");
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = siluf_(v[j]);
    } else if (act == ACT_GELU) {
        asm volatile("" ::: "memory");
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = gelu_erff_(v[j]);
    } else if (act == ACT_SIGMOID) {
        asm volatile("" ::: "memory");
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = sigmoidf_(v[j]);
    }
}

// ---------------------------------------------------------------- PTX: mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// explicit shared-space 16-byte accesses: pointers derived from the 1024-byte-aligned dynamic smem base lose their address
// space (the compiler emits generic LD.E / ST.E: long-scoreboard latency instead of LDS / STS)
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, const uint4& v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void lds64(uint32_t saddr, uint32_t& a, uint32_t& b) {
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(saddr) : "memory");
}
__device__ __forceinline__ uint4 lds128_sync(uint32_t saddr) {      // ordered against other shared-memory traffic (hand-offs)
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(uint32_t saddr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(saddr), "r"(v) : "memory");
}
// warp-level bf16 MMA (m16n8k16, fp32 accumulate from zero) for rank-16 side products inside epilogues: runs on the legacy
// tensor-core path next to the tcgen05 main loop. Fragments: a0..a3 = rows (g, g+8) x k (2c, 2c+8) pairs, b0/b1 = k (2c, 2c+8)
// pairs of column g, d0..d3 = rows (g, g+8) x columns (2c, 2c+1), with g = lane / 4, c = lane % 4.
__device__ __forceinline__ void mma16816_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.f));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
#ifndef VLB_MBAR_SUSPEND_NS
#define VLB_MBAR_SUSPEND_NS 0x989680u
#endif
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(VLB_MBAR_SUSPEND_NS)      // suspend-time hint: waiting warps sleep instead of polling
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_try_wait_poll(uint64_t* bar, uint32_t parity) {      // no suspend hint: lowest wake-up latency
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded spin: a protocol bug becomes a trap (error) instead of a hung GPU box. Waiters sleep by default; -DVLB_MBAR_POLL
// switches to plain polling (measured r01d on both attention kernels: polling is 5-10 % SLOWER, the pollers take issue slots).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
#ifdef VLB_MBAR_POLL
    while (!mbar_try_wait_poll(bar, parity)) {
#else
    while (!mbar_try_wait(bar, parity)) {
#endif
        if (++spins > (1u << 28)) { __trap(); }
    }
}

// Polling wait (no suspend hint) for latency-critical hand-offs of short, memory-bound pipelines: a sleeping waiter resumes
// ~1 us after the barrier completes, which is longer than the work between two hand-offs there.
__device__ __forceinline__ void mbar_wait_spin(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait_poll(bar, parity)) {
        if (++spins > (1u << 30)) { __trap(); }
    }
}

// ---------------------------------------------------------------- PTX: TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(const void* tmap, uint64_t* bar, void* smem, int x, int y) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(x), "r"(y)
        : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* tmap, const void* smem, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(smem)), "r"(x), "r"(y)
                 : "memory");
}
// TMA reduce-add of a shared-memory tile into global memory (element-wise fp32 add done by the copy engine / L2)
__device__ __forceinline__ void tma_reduce_add_2d(const void* tmap, const void* smem, int x, int y) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(smem)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// 1-D bulk copy global -> shared (bytes % 16 == 0, both addresses 16-byte aligned), completes on an mbarrier
__device__ __forceinline__ void bulk_load_1d(void* smem, const void* gptr, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(gptr)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_load_3d(const void* tmap, uint64_t* bar, void* smem, int x, int y, int z) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z)
        : "memory");
}
__device__ __forceinline__ void tma_store_3d(const void* tmap, const void* smem, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(smem)), "r"(x), "r"(y), "r"(z)
                 : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------- PTX: tcgen05
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(smem_dst)), "n"(NCOLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t addr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "n"(NCOLS) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16/fp16 inputs with fp32 accumulate
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// A operand from TMEM (K-major, two bf16 per 32-bit column: 8 columns per K=16 step), B from shared memory
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread `lane` receives row (lane_base+lane), cols [c, c+32)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// 8 consecutive fp32 columns of this thread's lane
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr)
        : "memory");
}
// store 32 lanes x 32 columns (thread `lane` writes its row)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------- PTX: CTA pair (cta_group::2) helpers
// Two CTAs of a 2-CTA cluster on one TPC execute ONE tcgen05.mma with M = 256: each CTA holds its own 128 rows of A and
// HALF of the B tile in shared memory and its own 128 accumulator lanes in TMEM. Only the leader (cluster rank 0) issues
// MMAs; TMA loads of both CTAs signal the leader's barrier; commits are multicast to both CTAs' barriers.
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta window addresses of a cluster CTA carry its rank in bit 24; clearing it addresses the leader's copy
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;
__device__ __forceinline__ void tma_load_2d_2sm(const void* tmap, uint64_t* leader_bar, void* smem, int x, int y) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(leader_bar) & PEER_BIT_MASK), "r"(x), "r"(y)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
        ::"r"(smem_u32(bar)), "r"(cta) : "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(smem_dst)), "n"(NCOLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t addr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far have retired
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// Shared-memory matrix descriptor (sm_100 UMMA), 128-byte swizzle.
//  K-major operand  : tile stored as rows of 128 B (64 bf16 along K); 8-row groups are 1024 B apart (SBO).
//  MN-major operand : tile stored as BLOCK_K rows of 128 B (64 bf16 along M/N); SBO = 1024 B between 8-row
//                     groups along K, LBO = byte distance between 64-element atoms along M/N.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;        // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;        // SWIZZLE_128B
    return d;
}
// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4)                       // D format: f32
         | (1u << 7)                       // A format: bf16
         | (1u << 10)                      // B format: bf16
         | ((a_mn_major ? 1u : 0u) << 15)
         | ((b_mn_major ? 1u : 0u) << 16)
         | ((uint32_t)(N >> 3) << 17)
         | ((uint32_t)(M >> 4) << 24);
}

}  // namespace vlb
