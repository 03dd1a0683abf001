// Optimiser step (global-norm clip + AdamW over one flat parameter buffer), per-parcel Pearson r, parcel->voxel paint,
// and small strided copy/transposes used to stage LoRA B matrices. All HBM-bound, one pass over their operands.
//   AdamW + CosineAnnealingLR + gradient_clip_val=1 : src/litmodule/videollama2_vlb_litmodule.py:345-379,
//                                                     config/experiment/VLB_vllama2_friends_lora.yaml:46
//   PearsonCorrCoef(num_outputs=P) after nan_to_num  : src/utils.py:101-110
//   r**2 painted on atlas labels                     : src/postprocessing/make_acc_brainmaps.py:50-53
#include "common.cuh"

namespace vlb {

__device__ DetSumWs g_sumsq_ws;

__global__ void sumsq_kernel(const float* __restrict__ g, long long n4, float* __restrict__ acc) {
    __shared__ float red[32];
    float s = 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(g)[i];
        s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    s = block_sum(s, red);
    det_grid_add(g_sumsq_ws, s, acc, red);
}

struct AdamArgs {
    float lr, b1, b2, eps, wd, bc1, bc2, max_norm, grad_scale;
};

// p, g, m, v fp32 [n]; pb bf16 working copy [n]. clip = min(1, max_norm / (||g*grad_scale|| + 1e-6)) from *sumsq.
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                             bf16* __restrict__ pb, long long n4, const float* __restrict__ sumsq, AdamArgs a) {
    float clip = 1.f;
    if (a.max_norm > 0.f) {
        const float norm = sqrtf(*sumsq) * a.grad_scale;
        clip = fminf(1.f, a.max_norm / (norm + 1e-6f));
    }
    const float gs = a.grad_scale * clip;
    const float step = a.lr / a.bc1, rbc2 = rsqrtf(a.bc2), decay = 1.f - a.lr * a.wd;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 pv = reinterpret_cast<float4*>(p)[i];
        const float4 gv = reinterpret_cast<const float4*>(g)[i];
        float4 mv = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
        float* pp = &pv.x; const float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gg = gp[j] * gs;
            mp[j] = a.b1 * mp[j] + (1.f - a.b1) * gg;
            vp[j] = a.b2 * vp[j] + (1.f - a.b2) * gg * gg;
            pp[j] = pp[j] * decay - step * mp[j] / (sqrtf(vp[j]) * rbc2 + a.eps);
        }
        reinterpret_cast<float4*>(p)[i] = pv;
        reinterpret_cast<float4*>(m)[i] = mv;
        reinterpret_cast<float4*>(v)[i] = vv;
        uint2 u; u.x = pack_bf16(pv.x, pv.y); u.y = pack_bf16(pv.z, pv.w);
        reinterpret_cast<uint2*>(pb)[i] = u;
    }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ s, bf16* __restrict__ d, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        d[i] = f2bf(s[i]);
}

// dst[r, c] = src[r, c]   (bf16, independent leading dimensions)
__global__ void copy2d_kernel(const bf16* __restrict__ src, int lds, bf16* __restrict__ dst, int ldd, int rows, int cols) {
    const long long total = (long long)rows * cols;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / cols), c = (int)(i % cols);
        dst[(size_t)r * ldd + c] = src[(size_t)r * lds + c];
    }
}
// dst[c, r] = src[r, c]   (bf16 transpose of a skinny matrix: LoRA B^T [r, N] -> B [N, r] inside a wider buffer)
__global__ void transpose2d_kernel(const bf16* __restrict__ src, int lds, bf16* __restrict__ dst, int ldd, int rows, int cols) {
    const long long total = (long long)rows * cols;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i / rows), r = (int)(i % rows);
        dst[(size_t)c * ldd + r] = src[(size_t)r * lds + c];
    }
}

// The same transpose for a whole table of (src, dst) pairs in ONE launch: blockIdx.y picks the row of desc = int64 [n, 6] =
// (src, dst, lds, ldd, rows, cols). The per-step refresh of the staged LoRA B matrices was 7 x 32 launches of ~8 us each.
__global__ void transpose2d_batch_kernel(const long long* __restrict__ desc) {
    const long long* d = desc + (size_t)blockIdx.y * 6;
    const bf16* src = reinterpret_cast<const bf16*>(d[0]);
    bf16* dst = reinterpret_cast<bf16*>(d[1]);
    const int lds = (int)d[2], ldd = (int)d[3], rows = (int)d[4], cols = (int)d[5];
    const long long total = (long long)rows * cols;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i / rows), r = (int)(i % rows);
        dst[(size_t)c * ldd + r] = src[(size_t)r * lds + c];
    }
}

// --------------------------------------------------------------------------------------------- Pearson
template <typename T> __device__ __forceinline__ float ld_clean(const T* p);
template <> __device__ __forceinline__ float ld_clean<float>(const float* p) {
    float v = *p;
    if (isnan(v)) return 0.f;
    if (isinf(v)) return v > 0 ? 3.4028234663852886e38f : -3.4028234663852886e38f;
    return v;
}
template <> __device__ __forceinline__ float ld_clean<bf16>(const bf16* p) {
    float v = bf2f(*p);
    if (isnan(v)) return 0.f;
    if (isinf(v)) return v > 0 ? 3.3895313892515355e38f : -3.3895313892515355e38f;   // bf16 max (torch.nan_to_num)
    return v;
}

// stats [5, P] fp64 += (sum x, sum y, sum xx, sum yy, sum xy) over this CTA's row slab; thread = one column
template <typename T>
__global__ void pearson_update_kernel(const T* __restrict__ x, const T* __restrict__ y, double* __restrict__ stats,
                                      long long N, long long P, int rows_per) {
    const long long col = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (col >= P) return;
    const long long r0 = (long long)blockIdx.y * rows_per;
    const long long r1 = r0 + rows_per < N ? r0 + rows_per : N;
    double sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0;
    for (long long r = r0; r < r1; ++r) {
        const double a = (double)ld_clean<T>(x + r * P + col), b = (double)ld_clean<T>(y + r * P + col);
        sx += a; sy += b; sxx += a * a; syy += b * b; sxy += a * b;
    }
    atomicAdd(stats + col, sx); atomicAdd(stats + P + col, sy); atomicAdd(stats + 2 * P + col, sxx);
    atomicAdd(stats + 3 * P + col, syy); atomicAdd(stats + 4 * P + col, sxy);
}

// Vectorised variant (P % 4 == 0): thread = 4 adjacent columns (one 16-byte fp32 / 8-byte bf16 access per array and row),
// four rows in flight per iteration and ~32 small CTAs per SM so the sweep is bound by HBM, not by load latency
// (ncu r01: the one-column kernel sat at 32 % active warps, fp64 pipe 17 %, long-scoreboard stalls).
template <typename T> struct PearsonVec;
template <> struct PearsonVec<float> {
    static constexpr int V = 4;
    __device__ static __forceinline__ void load(const float* p, float (&v)[4]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static constexpr float kMax = 3.4028234663852886e38f;
};
template <> struct PearsonVec<bf16> {
    static constexpr int V = 4;
    __device__ static __forceinline__ void load(const bf16* p, float (&v)[4]) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        const float2 a = unpack_bf16(t.x), b = unpack_bf16(t.y);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
    static constexpr float kMax = 3.3895313892515355e38f;
};
__device__ __forceinline__ float clean_val(float v, float vmax) {
    if (isnan(v)) return 0.f;
    if (isinf(v)) return v > 0 ? vmax : -vmax;
    return v;
}

template <typename T>
__global__ void __launch_bounds__(128) pearson_update_vec_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                 double* __restrict__ stats, long long N, long long P,
                                                                 int rows_per) {
    constexpr int V = PearsonVec<T>::V;
    const long long col = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * V;
    if (col >= P) return;
    const long long r0 = (long long)blockIdx.y * rows_per;
    const long long r1 = r0 + rows_per < N ? r0 + rows_per : N;
    double sx[V], sy[V], sxx[V], syy[V], sxy[V];
#pragma unroll
    for (int j = 0; j < V; ++j) { sx[j] = 0; sy[j] = 0; sxx[j] = 0; syy[j] = 0; sxy[j] = 0; }
    auto acc = [&](const float (&a)[V], const float (&b)[V]) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const double u = (double)clean_val(a[j], PearsonVec<T>::kMax), w = (double)clean_val(b[j], PearsonVec<T>::kMax);
            sx[j] += u; sy[j] += w; sxx[j] += u * u; syy[j] += w * w; sxy[j] += u * w;
        }
    };
    long long r = r0;
    for (; r + 3 < r1; r += 4) {
        float a[4][V], b[4][V];
#pragma unroll
        for (int k = 0; k < 4; ++k) { PearsonVec<T>::load(x + (r + k) * P + col, a[k]); PearsonVec<T>::load(y + (r + k) * P + col, b[k]); }
#pragma unroll
        for (int k = 0; k < 4; ++k) acc(a[k], b[k]);
    }
    for (; r + 1 < r1; ++r) {
        float a0[V], b0[V];
        PearsonVec<T>::load(x + r * P + col, a0); PearsonVec<T>::load(y + r * P + col, b0);
        acc(a0, b0);
    }
    if (r < r1) {
        float a0[V], b0[V];
        PearsonVec<T>::load(x + r * P + col, a0); PearsonVec<T>::load(y + r * P + col, b0);
        acc(a0, b0);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
        atomicAdd(stats + col + j, sx[j]); atomicAdd(stats + P + col + j, sy[j]); atomicAdd(stats + 2 * P + col + j, sxx[j]);
        atomicAdd(stats + 3 * P + col + j, syy[j]); atomicAdd(stats + 4 * P + col + j, sxy[j]);
    }
}

__global__ void pearson_finalize_kernel(const double* __restrict__ stats, double n, float* __restrict__ r, long long P) {
    const long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= P) return;
    const double sx = stats[c], sy = stats[P + c], sxx = stats[2 * P + c], syy = stats[3 * P + c], sxy = stats[4 * P + c];
    const double cov = sxy - sx * sy / n, vx = sxx - sx * sx / n, vy = syy - sy * sy / n;
    double v = cov / sqrt(vx * vy);
    v = v > 1.0 ? 1.0 : (v < -1.0 ? -1.0 : v);
    r[c] = (float)v;           // NaN when a column is constant, like torchmetrics
}

// vol[i] = label[i] > 0 ? f(r[label[i]-1]) : 0   (NiftiLabelsMasker.inverse_transform of r**2)
__global__ void paint_labels_kernel(const float* __restrict__ r, const int* __restrict__ labels, float* __restrict__ vol,
                                    long long V, int P, int square) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < V; i += (long long)gridDim.x * blockDim.x) {
        const int l = labels[i];
        float v = 0.f;
        if (l > 0 && l <= P) { v = r[l - 1]; if (square) v *= v; }
        vol[i] = v;
    }
}

static inline int grid1d(long long n, int threads) {
    long long b = ceil_div_ll(n, threads);
    long long cap = (long long)num_sms() * 16;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_grad_sumsq(const void* grad_f32, long long n, void* acc_f32, void* stream) {
    VLB_REQUIRE(n % 4 == 0, "grad_sumsq: n=%lld must be a multiple of 4", n);
    sumsq_kernel<<<grid1d(n / 4, 256), 256, 0, (cudaStream_t)stream>>>((const float*)grad_f32, n / 4, (float*)acc_f32);
    return check_launch("grad_sumsq");
}

extern "C" int vlb_adamw_step(void* param_f32, const void* grad_f32, void* m_f32, void* v_f32, void* param_bf16,
                              long long n, const void* sumsq_f32, float lr, float beta1, float beta2, float eps,
                              float weight_decay, int step, float max_norm, float grad_scale, void* stream) {
    VLB_REQUIRE(n % 4 == 0 && step >= 1, "adamw_step: n=%lld step=%d", n, step);
    AdamArgs a;
    a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.wd = weight_decay;
    a.bc1 = 1.f - powf(beta1, (float)step); a.bc2 = 1.f - powf(beta2, (float)step);
    a.max_norm = max_norm; a.grad_scale = grad_scale;
    adamw_kernel<<<grid1d(n / 4, 256), 256, 0, (cudaStream_t)stream>>>((float*)param_f32, (const float*)grad_f32,
                                                                      (float*)m_f32, (float*)v_f32, (bf16*)param_bf16,
                                                                      n / 4, (const float*)sumsq_f32, a);
    return check_launch("adamw_step");
}

extern "C" int vlb_f32_to_bf16(const void* src, void* dst, long long n, void* stream) {
    f32_to_bf16_kernel<<<grid1d(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)src, (bf16*)dst, n);
    return check_launch("f32_to_bf16");
}

extern "C" int vlb_copy2d(const void* src, int lds, void* dst, int ldd, int rows, int cols, int transpose, void* stream) {
    const long long total = (long long)rows * cols;
    if (transpose)
        transpose2d_kernel<<<grid1d(total, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)src, lds, (bf16*)dst, ldd, rows, cols);
    else
        copy2d_kernel<<<grid1d(total, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)src, lds, (bf16*)dst, ldd, rows, cols);
    return check_launch("copy2d");
}

extern "C" int vlb_transpose2d_batch(const void* desc_i64, int n, int blocks_per_item, void* stream) {
    VLB_REQUIRE(n > 0 && n <= 65535 && blocks_per_item > 0, "transpose2d_batch: n=%d blocks=%d", n, blocks_per_item);
    transpose2d_batch_kernel<<<dim3(blocks_per_item, n), 256, 0, (cudaStream_t)stream>>>((const long long*)desc_i64);
    return check_launch("transpose2d_batch");
}

extern "C" int vlb_pearson_update(const void* preds, const void* targets, void* stats_f64, long long N, long long P,
                                  int is_bf16, void* stream) {
    VLB_REQUIRE(N > 0 && P > 0, "pearson_update: N=%lld P=%lld", N, P);
    const int threads = 128;
    const int V = 4;
    const bool vec = (P % V == 0) && (((uintptr_t)preds | (uintptr_t)targets) % 16 == 0);
    const long long cols_per_cta = (long long)threads * (vec ? V : 1);
    const unsigned gx = (unsigned)ceil_div_ll(P, cols_per_cta);
    // enough row slabs to fill the machine (~32 four-warp CTAs per SM), at least 32 rows each
    long long slabs = ceil_div_ll((long long)num_sms() * 32, gx);
    if (slabs > ceil_div_ll(N, 32)) slabs = ceil_div_ll(N, 32);
    if (slabs < 1) slabs = 1;
    const int rows_per = (int)ceil_div_ll(N, slabs);
    dim3 grid(gx, (unsigned)ceil_div_ll(N, rows_per));
    cudaStream_t st = (cudaStream_t)stream;
    if (vec && is_bf16)
        pearson_update_vec_kernel<bf16><<<grid, threads, 0, st>>>((const bf16*)preds, (const bf16*)targets, (double*)stats_f64, N, P, rows_per);
    else if (vec)
        pearson_update_vec_kernel<float><<<grid, threads, 0, st>>>((const float*)preds, (const float*)targets, (double*)stats_f64, N, P, rows_per);
    else if (is_bf16)
        pearson_update_kernel<bf16><<<grid, threads, 0, st>>>((const bf16*)preds, (const bf16*)targets, (double*)stats_f64, N, P, rows_per);
    else
        pearson_update_kernel<float><<<grid, threads, 0, st>>>((const float*)preds, (const float*)targets, (double*)stats_f64, N, P, rows_per);
    return check_launch("pearson_update");
}

extern "C" int vlb_pearson_finalize(const void* stats_f64, long long n_rows, void* r_f32, long long P, void* stream) {
    pearson_finalize_kernel<<<(unsigned)ceil_div_ll(P, 256), 256, 0, (cudaStream_t)stream>>>((const double*)stats_f64,
                                                                                            (double)n_rows, (float*)r_f32, P);
    return check_launch("pearson_finalize");
}

extern "C" int vlb_paint_labels(const void* r_f32, const void* labels_i32, void* vol_f32, long long V, int P, int square,
                                void* stream) {
    paint_labels_kernel<<<grid1d(V, 256), 256, 0, (cudaStream_t)stream>>>((const float*)r_f32, (const int*)labels_i32,
                                                                         (float*)vol_f32, V, P, square);
    return check_launch("paint_labels");
}
