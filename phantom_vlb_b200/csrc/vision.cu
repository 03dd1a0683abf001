// Vision-side data-movement kernels: CLIP patch im2col + embedding assembly, STC depthwise 3x3 / SE / Conv3d gather,
// and the <video> splice of visual rows into the text embedding sequence. All are HBM-bound gathers with 16-byte
// accesses; the dense contractions they feed run on the tcgen05 GEMM.
//   CLIPVisionEmbeddings (HF, reached from VideoLLaMA2/videollama2/model/encoder.py:50)
//   timm RegStage Bottleneck pieces used by STCConnector (VideoLLaMA2/videollama2/model/projector.py:153-184)
//   nn.Conv3d(k=2,s=2,p=1) sampler (projector.py:164-174)
//   prepare_inputs_labels_for_multimodal (VideoLLaMA2/videollama2/model/videollama2_arch.py:161-263)
#include "common.cuh"

namespace vlb {

static inline int grid_for(long long total, int threads) {
    long long b = ceil_div_ll(total, threads);
    long long cap = (long long)num_sms() * 16;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

// pixels fp32 [N,3,IMG,IMG] -> cols bf16 [N*G*G, KP], k = c*P*P + ky*P + kx (conv weight flatten order), zero pad to KP
__global__ void patch_im2col_kernel(const float* __restrict__ pix, bf16* __restrict__ cols, long long N, int IMG, int P,
                                    int G, int KP) {
    const int K = 3 * P * P;
    const long long total = N * G * G * (long long)KP;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % KP);
        const long long pr = i / KP;
        float v = 0.f;
        if (k < K) {
            const int c = k / (P * P), ky = (k / P) % P, kx = k % P;
            const int gx = (int)(pr % G), gy = (int)((pr / G) % G);
            const long long n = pr / (G * G);
            v = pix[((n * 3 + c) * IMG + gy * P + ky) * (long long)IMG + gx * P + kx];
        }
        cols[i] = f2bf(v);
    }
}

// uint8 frames: the CLIPImageProcessor arithmetic the reference runs on the CPU before storing fp32 pixels
// (rescale by 1/255, then (x - mean[c]) / std[c]; extractfeatures.py:344-350 via processor.preprocess) is applied in the
// gather itself, with the same fp32 operation order (round-to-nearest mul, sub, div: bit-identical to the stored fp32
// pixels), so a clip travels as 2.7 MB of bytes instead of 10.8 MB of floats.
struct ClipNorm { float mean[3]; float std[3]; };
__global__ void patch_im2col_u8_kernel(const uint8_t* __restrict__ pix, bf16* __restrict__ cols, long long N, int IMG, int P,
                                       int G, int KP, ClipNorm nm) {
    const int K = 3 * P * P;
    const long long total = N * G * G * (long long)KP;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % KP);
        const long long pr = i / KP;
        float v = 0.f;
        if (k < K) {
            const int c = k / (P * P), ky = (k / P) % P, kx = k % P;
            const int gx = (int)(pr % G), gy = (int)((pr / G) % G);
            const long long n = pr / (G * G);
            const float raw = (float)pix[((n * 3 + c) * IMG + gy * P + ky) * (long long)IMG + gx * P + kx];
            v = __fdiv_rn(__fsub_rn(__fmul_rn(raw, 0.00392156862745098f), nm.mean[c]), nm.std[c]);
        }
        cols[i] = f2bf(v);
    }
}

// hidden0[n, t, :] = LN( (t == 0 ? cls : patch[n, t-1, :]) + pos[t, :] ); one CTA (128 thr) per token; H % 8 == 0, H <= 1024*?
template <int MAXV>
__global__ void __launch_bounds__(128) clip_embed_kernel(const bf16* __restrict__ patch, const bf16* __restrict__ cls,
                                                         const bf16* __restrict__ pos, const bf16* __restrict__ lnw,
                                                         const bf16* __restrict__ lnb, bf16* __restrict__ out, int T,
                                                         int H, float eps) {
    __shared__ float red[32];
    const long long row = blockIdx.x;
    const int t = (int)(row % T);
    const long long n = row / T;
    const bf16* src = (t == 0) ? cls : patch + (n * (T - 1) + (t - 1)) * (long long)H;
    float v[MAXV][8];
    float s = 0.f;
    const int nchunk = H / 8;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * 128;
        if (c < nchunk) {
            float p[8];
            load8(src + c * 8, v[i]);
            load8(pos + (size_t)t * H + c * 8, p);
#pragma unroll
            for (int j = 0; j < 8; ++j) { v[i][j] = bf2f(f2bf(v[i][j] + p[j])); s += v[i][j]; }
        }
    }
    const float mean = block_sum(s, red) / (float)H;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * 128;
        if (c < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
        }
    }
    const float rstd = rsqrtf(block_sum(ss, red) / (float)H + eps);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * 128;
        if (c < nchunk) {
            float w[8], b[8], o[8];
            load8(lnw + c * 8, w); load8(lnb + c * 8, b);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * w[j] + b[j];
            store8(out + row * H + c * 8, o);
        }
    }
}

// drop CLS: in [N, T, H] -> out [N, T-1, H]   (feature_select 'patch', encoder.py:31-39)
__global__ void drop_cls_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, long long N, int T, int H) {
    const int cpr = H / 8;
    const long long total = N * (T - 1) * (long long)cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cpr);
        const long long r = i / cpr;
        const long long n = r / (T - 1);
        const int t = (int)(r % (T - 1));
        *reinterpret_cast<uint4*>(out + r * H + c * 8) =
            *reinterpret_cast<const uint4*>(in + (n * T + t + 1) * (long long)H + c * 8);
    }
}

// depthwise 3x3, stride 1, pad 1, no bias; NHWC x [F, Hh, Ww, C]; w packed [9, C] (tap-major).
// One thread per (frame, output row, 8 channels) walks along x with the 3x3 window held in registers as packed bf16:
// 3 new 16-byte loads per output instead of 9 (the one-pixel-per-thread version was L2-bandwidth bound at 9x the
// algorithmic reads: adjacent threads are adjacent CHANNELS, so no neighbour load ever hit L1). Taps are accumulated in
// the same (dy, dx) order; out-of-image taps enter as zeros.
__global__ void __launch_bounds__(128, 4) dwconv3x3_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                        bf16* __restrict__ y, long long F, int Hh, int Ww, int C) {
    const int cpr = C / 8;
    const long long total = F * Hh * (long long)cpr;
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int c = (int)(i % cpr);
    const long long row = i / cpr;
    const int yy = (int)(row % Hh);
    const long long f = row / Hh;
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    uint4 wq[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) wq[t] = *reinterpret_cast<const uint4*>(w + (size_t)t * C + c * 8);
    const bf16* rp[3];
    bool ok[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const int y2 = yy + d - 1;
        ok[d] = y2 >= 0 && y2 < Hh;
        rp[d] = x + ((f * Hh + (ok[d] ? y2 : yy)) * Ww) * (long long)C + c * 8;
    }
    uint4 left[3], mid[3], right[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        left[d] = zero;
        mid[d] = ok[d] ? *reinterpret_cast<const uint4*>(rp[d]) : zero;
    }
    bf16* yp = y + ((f * Hh + yy) * Ww) * (long long)C + c * 8;
    for (int xx = 0; xx < Ww; ++xx) {
        const bool more = xx + 1 < Ww;
#pragma unroll
        for (int d = 0; d < 3; ++d)
            right[d] = (ok[d] && more) ? *reinterpret_cast<const uint4*>(rp[d] + (long long)(xx + 1) * C) : zero;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            float xv[8], wv[8];
            unpack8(left[d], xv); unpack8(wq[d * 3 + 0], wv);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += xv[j] * wv[j];
            unpack8(mid[d], xv); unpack8(wq[d * 3 + 1], wv);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += xv[j] * wv[j];
            unpack8(right[d], xv); unpack8(wq[d * 3 + 2], wv);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += xv[j] * wv[j];
        }
        store8(yp + (long long)xx * C, acc);
#pragma unroll
        for (int d = 0; d < 3; ++d) { left[d] = mid[d]; mid[d] = right[d]; }
    }
}

// SE squeeze: mean over the HW positions of each frame. x [F, HW, C] -> out [F, C]. Block = 32 channel chunks x 8 position
// slices: a warp reads 512 contiguous bytes per position, the 8 slices run in parallel and are summed in fixed order
// (the one-thread-per-(f, chunk) version walked all HW positions serially with only F*C/8 threads in the grid).
__global__ void __launch_bounds__(256) se_pool_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, long long F,
                                                      int HW, int C) {
    __shared__ float part[8][32][9];
    const int cpr = C / 8;
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long long f = blockIdx.y;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c < cpr) {
        for (int p = threadIdx.y; p < HW; p += 8) {
            float v[8];
            load8(x + (f * HW + p) * (long long)C + c * 8, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += v[j];
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) part[threadIdx.y][threadIdx.x][j] = acc[j];
    __syncthreads();
    if (threadIdx.y == 0 && c < cpr) {
#pragma unroll
        for (int sl = 1; sl < 8; ++sl) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += part[sl][threadIdx.x][j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] /= (float)HW;
        store8(out + f * C + c * 8, acc);
    }
}

// SE excite: y[f, p, :] = x[f, p, :] * gate[f, :]
__global__ void se_scale_kernel(const bf16* __restrict__ x, const bf16* __restrict__ gate, bf16* __restrict__ y,
                                long long F, int HW, int C) {
    const int cpr = C / 8;
    const long long total = F * HW * (long long)cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cpr);
        const long long r = i / cpr;
        const long long f = r / HW;
        float v[8], g[8];
        load8(x + r * C + c * 8, v);
        load8(gate + f * C + c * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= g[j];
        store8(y + r * C + c * 8, v);
    }
}

// Conv3d(k=2,s=2,p=1) gather: x [B, T, Hh, Ww, C] -> cols [B*To*Ho*Wo, 8*C], k = ((kt*2+kh)*2+kw)*C + c, zero padding
__global__ void conv3d_gather_kernel(const bf16* __restrict__ x, bf16* __restrict__ cols, long long B, int T, int Hh,
                                     int Ww, int C, int To, int Ho, int Wo) {
    const int cpr = C / 8;
    const long long total = B * To * Ho * Wo * 8LL * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cpr);
        const int tap = (int)((i / cpr) % 8);
        const long long o = i / (8LL * cpr);
        const int wo = (int)(o % Wo), ho = (int)((o / Wo) % Ho), to = (int)((o / ((long long)Wo * Ho)) % To);
        const long long b = o / ((long long)Wo * Ho * To);
        const int t = to * 2 - 1 + (tap >> 2), h = ho * 2 - 1 + ((tap >> 1) & 1), w = wo * 2 - 1 + (tap & 1);
        uint4 v = make_uint4(0, 0, 0, 0);
        if (t >= 0 && t < T && h >= 0 && h < Hh && w >= 0 && w < Ww)
            v = *reinterpret_cast<const uint4*>(x + (((b * T + t) * Hh + h) * Ww + w) * (long long)C + c * 8);
        *reinterpret_cast<uint4*>(cols + o * 8LL * C + (long long)tap * C + c * 8) = v;
    }
}

// position of the first <video> token (id == video_id) in each row of ids [B, L]; -1 if absent
__global__ void find_video_kernel(const long long* __restrict__ ids, int* __restrict__ pos, int L, long long video_id) {
    __shared__ int best;
    if (threadIdx.x == 0) best = 0x7fffffff;
    __syncthreads();
    const long long* r = ids + (size_t)blockIdx.x * L;
    int m = 0x7fffffff;
    for (int i = threadIdx.x; i < L; i += blockDim.x) if (r[i] == video_id) m = min(m, i);
    if (m != 0x7fffffff) atomicMin(&best, m);
    __syncthreads();
    if (threadIdx.x == 0) pos[blockIdx.x] = (best == 0x7fffffff) ? -1 : best;
}

// out[b, p, :] = p < pos ? E[ids[p]] : (p < pos+Nv ? vis[b, p-pos] : E[ids[p-Nv+1]])      (S = L + Nv - 1)
__global__ void splice_kernel(const long long* __restrict__ ids, const int* __restrict__ pos, const bf16* __restrict__ table,
                              const bf16* __restrict__ vis, bf16* __restrict__ out, long long B, int L, int Nv, int H,
                              int vocab) {
    const int S = L + Nv - 1, cpr = H / 8;
    const long long total = B * S * (long long)cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cpr);
        const long long r = i / cpr;
        const int p = (int)(r % S);
        const long long b = r / S;
        const int vp = pos[b];
        const bf16* src;
        if (vp >= 0 && p >= vp && p < vp + Nv) {
            src = vis + (b * Nv + (p - vp)) * (long long)H;
        } else {
            long long id = ids[b * L + (p < vp || vp < 0 ? p : p - Nv + 1)];
            id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
            src = table + id * (long long)H;
        }
        *reinterpret_cast<uint4*>(out + r * H + c * 8) = *reinterpret_cast<const uint4*>(src + c * 8);
    }
}

// language fp32 ids [B, L] (the datamodule stores ids as float32) -> int64 (.long() truncation)
__global__ void ids_to_long_kernel(const float* __restrict__ f, long long* __restrict__ o, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        o[i] = (long long)f[i];
}

// make_weight_mask (src/litmodule/videollama2_vlb_litmodule.py:178-203), bit-exact layout:
// row = [0]*pad_left ++ repeat_interleave(vis_w, tok_per_frame) ++ [0]*(2+inst) ++ lang_w[:dialog] ++ [0]*(4+pad)
// values are rounded to bf16 (the reference casts to config.dtype) and stored as fp32 for the pooling kernel.
__global__ void weight_mask_kernel(const long long* __restrict__ padvals, const double* __restrict__ visw,
                                   const double* __restrict__ langw, float* __restrict__ out, int Tp, int tok, int NL,
                                   int S) {
    const int b = blockIdx.x;
    const long long pad = padvals[b * 3], inst = padvals[b * 3 + 1], dialog = padvals[b * 3 + 2];
    const long long nvis = (long long)Tp * tok;
    const long long len = nvis + 2 + inst + dialog + 4 + pad;
    const long long left = S - len;
    for (int p = threadIdx.x; p < S; p += blockDim.x) {
        const long long q = p - left;
        float v = 0.f;
        if (q >= 0 && q < nvis) v = bf2f(__double2bfloat16(visw[b * Tp + q / tok]));
        else if (q >= nvis + 2 + inst && q < nvis + 2 + inst + dialog && (q - nvis - 2 - inst) < NL)
            v = bf2f(__double2bfloat16(langw[b * NL + (q - nvis - 2 - inst)]));
        out[(size_t)b * S + p] = v;
    }
}

}  // namespace vlb

using namespace vlb;

extern "C" int vlb_patch_im2col(const void* pixels_f32, void* cols, long long N, int img, int patch, int kpad,
                                void* stream) {
    VLB_REQUIRE(img % patch == 0 && kpad >= 3 * patch * patch && kpad % 8 == 0, "patch_im2col: img=%d patch=%d kpad=%d",
                img, patch, kpad);
    const int G = img / patch;
    patch_im2col_kernel<<<grid_for(N * G * G * (long long)kpad, 256), 256, 0, (cudaStream_t)stream>>>(
        (const float*)pixels_f32, (bf16*)cols, N, img, patch, G, kpad);
    return check_launch("patch_im2col");
}

extern "C" int vlb_patch_im2col_u8(const void* pixels_u8, void* cols, long long N, int img, int patch, int kpad,
                                   const float* mean3, const float* std3, void* stream) {
    VLB_REQUIRE(img % patch == 0 && kpad >= 3 * patch * patch && kpad % 8 == 0, "patch_im2col_u8: img=%d patch=%d kpad=%d",
                img, patch, kpad);
    VLB_REQUIRE(mean3 != nullptr && std3 != nullptr, "patch_im2col_u8: mean/std (3 host floats each) are required");
    ClipNorm nm;
    for (int c = 0; c < 3; ++c) { nm.mean[c] = mean3[c]; nm.std[c] = std3[c]; }
    const int G = img / patch;
    patch_im2col_u8_kernel<<<grid_for(N * G * G * (long long)kpad, 256), 256, 0, (cudaStream_t)stream>>>(
        (const uint8_t*)pixels_u8, (bf16*)cols, N, img, patch, G, kpad, nm);
    return check_launch("patch_im2col_u8");
}

extern "C" int vlb_clip_embed(const void* patch, const void* cls, const void* pos, const void* lnw, const void* lnb,
                              void* out, long long N, int T, int H, float eps, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && H <= 4096, "clip_embed: H=%d", H);
    cudaStream_t st = (cudaStream_t)stream;
    if (H <= 1024)
        clip_embed_kernel<1><<<(unsigned)(N * T), 128, 0, st>>>((const bf16*)patch, (const bf16*)cls, (const bf16*)pos,
                                                               (const bf16*)lnw, (const bf16*)lnb, (bf16*)out, T, H, eps);
    else
        clip_embed_kernel<4><<<(unsigned)(N * T), 128, 0, st>>>((const bf16*)patch, (const bf16*)cls, (const bf16*)pos,
                                                               (const bf16*)lnw, (const bf16*)lnb, (bf16*)out, T, H, eps);
    return check_launch("clip_embed");
}

extern "C" int vlb_drop_cls(const void* in, void* out, long long N, int T, int H, void* stream) {
    VLB_REQUIRE(H % 8 == 0, "drop_cls: H=%d", H);
    drop_cls_kernel<<<grid_for(N * (T - 1) * (long long)(H / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)in, (bf16*)out, N, T, H);
    return check_launch("drop_cls");
}

extern "C" int vlb_dwconv3x3(const void* x, const void* w9c, void* y, long long F, int Hh, int Ww, int C, void* stream) {
    VLB_REQUIRE(C % 8 == 0, "dwconv3x3: C=%d", C);
    VLB_REQUIRE(F > 0 && Hh > 0 && Ww > 0, "dwconv3x3: F=%lld Hh=%d Ww=%d", F, Hh, Ww);
    const long long threads = F * Hh * (long long)(C / 8);
    dwconv3x3_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (const bf16*)w9c, (bf16*)y, F, Hh, Ww, C);
    return check_launch("dwconv3x3");
}

extern "C" int vlb_se_pool(const void* x, void* out, long long F, int HW, int C, void* stream) {
    VLB_REQUIRE(C % 8 == 0, "se_pool: C=%d", C);
    VLB_REQUIRE(F > 0 && F <= 65535 && HW > 0, "se_pool: F=%lld HW=%d", F, HW);
    se_pool_kernel<<<dim3((unsigned)((C / 8 + 31) / 32), (unsigned)F), dim3(32, 8), 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (bf16*)out, F, HW, C);
    return check_launch("se_pool");
}

extern "C" int vlb_se_scale(const void* x, const void* gate, void* y, long long F, int HW, int C, void* stream) {
    VLB_REQUIRE(C % 8 == 0, "se_scale: C=%d", C);
    se_scale_kernel<<<grid_for(F * HW * (long long)(C / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (const bf16*)gate, (bf16*)y, F, HW, C);
    return check_launch("se_scale");
}

extern "C" int vlb_conv3d_gather(const void* x, void* cols, long long B, int T, int Hh, int Ww, int C, void* stream) {
    VLB_REQUIRE(C % 8 == 0, "conv3d_gather: C=%d", C);
    const int To = T / 2 + 1, Ho = Hh / 2 + 1, Wo = Ww / 2 + 1;
    conv3d_gather_kernel<<<grid_for(B * To * Ho * Wo * 8LL * (C / 8), 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, (bf16*)cols, B, T, Hh, Ww, C, To, Ho, Wo);
    return check_launch("conv3d_gather");
}

extern "C" int vlb_splice_embed(const void* ids_i64, const void* table, const void* vis, void* out, void* video_pos_i32,
                                long long B, int L, int Nv, int H, int vocab, long long video_id, void* stream) {
    VLB_REQUIRE(H % 8 == 0 && L > 0 && Nv >= 0, "splice_embed: H=%d L=%d Nv=%d", H, L, Nv);
    cudaStream_t st = (cudaStream_t)stream;
    find_video_kernel<<<(unsigned)B, 256, 0, st>>>((const long long*)ids_i64, (int*)video_pos_i32, L, video_id);
    const int S = L + Nv - 1;
    splice_kernel<<<grid_for(B * S * (long long)(H / 8), 256), 256, 0, st>>>(
        (const long long*)ids_i64, (const int*)video_pos_i32, (const bf16*)table, (const bf16*)vis, (bf16*)out, B, L, Nv,
        H, vocab);
    return check_launch("splice_embed");
}

extern "C" int vlb_ids_to_long(const void* ids_f32, void* ids_i64, long long n, void* stream) {
    ids_to_long_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)ids_f32, (long long*)ids_i64, n);
    return check_launch("ids_to_long");
}

extern "C" int vlb_weight_mask(const void* padvals_i64, const void* vis_w_f64, const void* lang_w_f64, void* out_f32,
                               int B, int Tp, int tok_per_frame, int NL, int S, void* stream) {
    VLB_REQUIRE(B > 0 && S > 0, "weight_mask: B=%d S=%d", B, S);
    weight_mask_kernel<<<B, 256, 0, (cudaStream_t)stream>>>((const long long*)padvals_i64, (const double*)vis_w_f64,
                                                           (const double*)lang_w_f64, (float*)out_f32, Tp, tok_per_frame,
                                                           NL, S);
    return check_launch("weight_mask");
}
