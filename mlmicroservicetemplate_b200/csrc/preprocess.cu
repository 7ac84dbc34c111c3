// K1: fused antialiased-bilinear resize + crop + /255 + (x-mean)/std + HWC->NCHW (+ bf16 cast).
//
// Arithmetic contract (bit-for-bit what torch 2.11 CPU does for
//   F.interpolate(x.float(), size, mode="bilinear", align_corners=False, antialias=True)
// reached through SP/torchvision/transforms/_functional_tensor.py:462-467, followed by
// /255 (:98-99) and sub(mean).div(std) (:928)):
//   * triangle-filter weights per output index i, computed with the promotions of
//     ATen's `_compute_indices_min_size_weights_aa`:
//       scale = float(in)/float(out); support = scale>=1 ? scale : 1; invscale = scale>=1 ? float(1.0/scale) : 1
//       center = float(double(scale)*(i+0.5));  xmin = max(int64(double(center-support)+0.5), 0)
//       xsize  = clamp(min(int64(double(center+support)+0.5), in) - xmin, 0, 2*ceil(support)+1)
//       w_j    = tri(float((double(float(j+xmin)-center)+0.5)*double(invscale))),  w_j /= sum_j w_j   (fp32)
//   * horizontal pass first into fp32, then vertical; each pass accumulates taps in order as
//       t = x0*w0;  then groups of four taps with separate multiply and add;  the (xsize-1)%4
//       trailing taps with fused multiply-add
//     (the shape the compiled ATen inner loop has — established by probing, see oracle/preprocess_np.py).
//   * a pass whose input and output size are equal degenerates to weights (1, 0) and is exact.
//
// Work decomposition: one CTA per (image, band of BAND output rows).  The CTA streams the source
// rows its band needs through a double-buffered shared-memory row buffer with 16-byte cp.async
// copies (coalesced; head/tail bytes of unaligned rows are copied bytewise), thread `ox` does
// the horizontal pass for output column ox (3 channels) and feeds BAND x 3 register accumulators
// of the vertical pass, so the fp32 intermediate never leaves the SM.
#include <cstdio>
#include <cuda_bf16.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int OUT = B2P_OUT_HW;      // 224
constexpr int BAND = 8;              // output rows per CTA
constexpr int THREADS = 256;         // 224 compute lanes + 32 that only help copying

__device__ __forceinline__ float tri_filter(float x) {
    x = fabsf(x);
    return x < 1.0f ? static_cast<float>(1.0 - static_cast<double>(x)) : 0.0f;
}

struct AxisGeom {
    float scale, support, invscale;
    int max_taps;
};

__device__ __forceinline__ AxisGeom axis_geom(int in_size, int out_size) {
    AxisGeom g;
    g.scale = __fdiv_rn(static_cast<float>(in_size), static_cast<float>(out_size));
    g.support = g.scale >= 1.0f ? g.scale : 1.0f;
    g.invscale = g.scale >= 1.0f ? static_cast<float>(1.0 / static_cast<double>(g.scale)) : 1.0f;
    g.max_taps = static_cast<int>(ceilf(g.support)) * 2 + 1;
    return g;
}

// Index range of output index i (in resized-image coordinates).
__device__ __forceinline__ void axis_range(const AxisGeom& g, int i, int in_size, float& center,
                                           int& xmin, int& xsize) {
    center = static_cast<float>(static_cast<double>(g.scale) * (static_cast<double>(i) + 0.5));
    const long long lo = static_cast<long long>(static_cast<double>(__fsub_rn(center, g.support)) + 0.5);
    const long long hi = static_cast<long long>(static_cast<double>(__fadd_rn(center, g.support)) + 0.5);
    xmin = static_cast<int>(lo > 0 ? lo : 0);
    int xs = static_cast<int>(hi < in_size ? hi : in_size) - xmin;
    xs = xs < 0 ? 0 : xs;
    xsize = xs > g.max_taps ? g.max_taps : xs;
}

__device__ __forceinline__ float axis_raw_weight(const AxisGeom& g, float center, int xmin, int j) {
    const float a = __fsub_rn(static_cast<float>(j + xmin), center);
    const float t = static_cast<float>((static_cast<double>(a) + 0.5) * static_cast<double>(g.invscale));
    return tri_filter(t);
}

__device__ __forceinline__ void cp_async_16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Copies `nbytes` from global `g` into shared `buf + (g & 15)` so that 16-byte vectors line up.
__device__ __forceinline__ void stage_row(uint8_t* buf, const uint8_t* g, int nbytes) {
    const int shift = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 15u);
    uint8_t* dst = buf + shift;
    int head = (16 - shift) & 15;
    head = head < nbytes ? head : nbytes;
    const int nvec = (nbytes - head) >> 4;
    const int tail0 = head + (nvec << 4);
    for (int i = threadIdx.x; i < nvec; i += THREADS) cp_async_16(dst + head + (i << 4), g + head + (i << 4));
    for (int i = threadIdx.x; i < head; i += THREADS) dst[i] = __ldg(g + i);
    for (int i = tail0 + threadIdx.x; i < nbytes; i += THREADS) dst[i] = __ldg(g + i);
}

template <bool OUT_F32>
__global__ void __launch_bounds__(THREADS)
preprocess_kernel(const uint8_t* __restrict__ slab, const b2p_image* __restrict__ imgs, void* __restrict__ out,
                  int wh_taps, int wv_taps, int rowbuf_bytes) {
    extern __shared__ __align__(16) uint8_t smem[];
    float* wh = reinterpret_cast<float*>(smem);                       // [wh_taps][OUT]
    float* wv = wh + wh_taps * OUT;                                   // [BAND][wv_taps]
    int* vmeta = reinterpret_cast<int*>(wv + BAND * wv_taps);         // ymin[BAND], ysize[BAND]
    uint8_t* rowbuf = reinterpret_cast<uint8_t*>(vmeta + 2 * BAND);   // 2 x rowbuf_bytes (16B aligned)

    const b2p_image im = imgs[blockIdx.y];
    const int band0 = blockIdx.x * BAND;
    const int tid = threadIdx.x;
    const uint8_t* src = slab + im.offset;
    const int row_bytes = im.w * 3;

    // ---- horizontal weights for this thread's output column, transposed into smem
    int xmin = 0, xsize = 0;
    if (tid < OUT) {
        const AxisGeom g = axis_geom(im.w, im.res_w);
        float center;
        axis_range(g, im.left + tid, im.w, center, xmin, xsize);
        float total = 0.f;
        for (int j = 0; j < xsize; ++j) {
            const float w = axis_raw_weight(g, center, xmin, j);
            wh[j * OUT + tid] = w;
            total = __fadd_rn(total, w);
        }
        if (total != 0.f)
            for (int j = 0; j < xsize; ++j) wh[j * OUT + tid] = __fdiv_rn(wh[j * OUT + tid], total);
    }
    // ---- vertical weights for the band's rows
    if (tid < BAND) {
        const AxisGeom g = axis_geom(im.h, im.res_h);
        float center;
        int ymin, ysize;
        axis_range(g, im.top + band0 + tid, im.h, center, ymin, ysize);
        float total = 0.f;
        for (int j = 0; j < ysize; ++j) {
            const float w = axis_raw_weight(g, center, ymin, j);
            wv[tid * wv_taps + j] = w;
            total = __fadd_rn(total, w);
        }
        if (total != 0.f)
            for (int j = 0; j < ysize; ++j) wv[tid * wv_taps + j] = __fdiv_rn(wv[tid * wv_taps + j], total);
        vmeta[tid] = ymin;
        vmeta[BAND + tid] = ysize;
    }
    __syncthreads();
    const int iy0 = vmeta[0];
    const int iy1 = vmeta[BAND - 1] + vmeta[2 * BAND - 1];   // ranges are monotone in the output row

    float acc[BAND][3];
#pragma unroll
    for (int r = 0; r < BAND; ++r) acc[r][0] = acc[r][1] = acc[r][2] = 0.f;

    const int hmain = xsize > 0 ? ((xsize - 1) >> 2) << 2 : 0;   // taps 1..hmain: mul then add

    if (iy0 < iy1) stage_row(rowbuf, src + static_cast<size_t>(iy0) * row_bytes, row_bytes);
    cp_async_commit();

    for (int iy = iy0; iy < iy1; ++iy) {
        cp_async_wait_all();
        __syncthreads();
        const int cur = (iy - iy0) & 1;
        if (iy + 1 < iy1)
            stage_row(rowbuf + (cur ^ 1) * rowbuf_bytes, src + static_cast<size_t>(iy + 1) * row_bytes,
                      row_bytes);
        cp_async_commit();

        if (tid < OUT) {
            const uint8_t* g = src + static_cast<size_t>(iy) * row_bytes;
            const uint8_t* s = rowbuf + cur * rowbuf_bytes + (reinterpret_cast<uintptr_t>(g) & 15u) + xmin * 3;
            // horizontal pass, 3 channels
            float w = wh[tid];
            float h0 = __fmul_rn(static_cast<float>(s[0]), w);
            float h1 = __fmul_rn(static_cast<float>(s[1]), w);
            float h2 = __fmul_rn(static_cast<float>(s[2]), w);
            int j = 1;
            for (; j <= hmain; ++j) {
                w = wh[j * OUT + tid];
                h0 = __fadd_rn(h0, __fmul_rn(static_cast<float>(s[3 * j + 0]), w));
                h1 = __fadd_rn(h1, __fmul_rn(static_cast<float>(s[3 * j + 1]), w));
                h2 = __fadd_rn(h2, __fmul_rn(static_cast<float>(s[3 * j + 2]), w));
            }
            for (; j < xsize; ++j) {
                w = wh[j * OUT + tid];
                h0 = __fmaf_rn(static_cast<float>(s[3 * j + 0]), w, h0);
                h1 = __fmaf_rn(static_cast<float>(s[3 * j + 1]), w, h1);
                h2 = __fmaf_rn(static_cast<float>(s[3 * j + 2]), w, h2);
            }
            // vertical pass: this source row is tap t of every band row whose range contains it
#pragma unroll
            for (int r = 0; r < BAND; ++r) {
                const int t = iy - vmeta[r];
                const int ys = vmeta[BAND + r];
                if (t >= 0 && t < ys) {
                    const float vw = wv[r * wv_taps + t];
                    const int vmain = ((ys - 1) >> 2) << 2;
                    if (t == 0) {
                        acc[r][0] = __fmul_rn(h0, vw);
                        acc[r][1] = __fmul_rn(h1, vw);
                        acc[r][2] = __fmul_rn(h2, vw);
                    } else if (t <= vmain) {
                        acc[r][0] = __fadd_rn(acc[r][0], __fmul_rn(h0, vw));
                        acc[r][1] = __fadd_rn(acc[r][1], __fmul_rn(h1, vw));
                        acc[r][2] = __fadd_rn(acc[r][2], __fmul_rn(h2, vw));
                    } else {
                        acc[r][0] = __fmaf_rn(h0, vw, acc[r][0]);
                        acc[r][1] = __fmaf_rn(h1, vw, acc[r][1]);
                        acc[r][2] = __fmaf_rn(h2, vw, acc[r][2]);
                    }
                }
            }
        }
    }

    if (tid < OUT) {
        const float mean[3] = {0.485f, 0.456f, 0.406f};
        const float stdv[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
        for (int r = 0; r < BAND; ++r) {
            const int oy = band0 + r;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float v = __fdiv_rn(__fsub_rn(__fdiv_rn(acc[r][c], 255.0f), mean[c]), stdv[c]);
                const size_t o = ((static_cast<size_t>(blockIdx.y) * 3 + c) * OUT + oy) * OUT + tid;
                if (OUT_F32) reinterpret_cast<float*>(out)[o] = v;
                else reinterpret_cast<__nv_bfloat16*>(out)[o] = __float2bfloat16_rn(v);
            }
        }
    }
}

// Crop-only fast path: every image of the batch already has the resized size (res == source), so
// the resize passes are the identity and each output value depends on one source byte.  The
// normalisation is a 3 x 256 table (built once per handle with the same IEEE divisions as the general
// kernel, hence bit-identical).  One thread = 8 consecutive pixels of one row: 24 source bytes fetched
// as seven aligned 32-bit words + funnel shifts (any byte alignment), three 16-byte planar stores.
__global__ void norm_lut_kernel(float* __restrict__ lut) {
    const int v = threadIdx.x;                       // 256 threads
    const float mean[3] = {0.485f, 0.456f, 0.406f};
    const float stdv[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
    for (int c = 0; c < 3; ++c)
        lut[c * 256 + v] = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(v), 255.0f), mean[c]), stdv[c]);
}

template <bool OUT_F32>
__global__ void __launch_bounds__(256)
preprocess_identity_kernel(const uint8_t* __restrict__ slab, const b2p_image* __restrict__ imgs, void* __restrict__ out,
                           const float* __restrict__ lut_g, int n) {
    __shared__ float lut[3 * 256];
    for (int i = threadIdx.x; i < 3 * 256; i += 256) lut[i] = __ldg(lut_g + i);
    __syncthreads();
    const long t = blockIdx.x * 256L + threadIdx.x;
    constexpr int GROUPS = OUT / 8;                  // 28 groups of 8 pixels per row
    if (t >= static_cast<long>(n) * OUT * GROUPS) return;
    const int g = static_cast<int>(t % GROUPS);
    const int row = static_cast<int>((t / GROUPS) % OUT);
    const int img = static_cast<int>(t / (GROUPS * OUT));
    const b2p_image im = imgs[img];
    const uint8_t* src = slab + im.offset + (static_cast<size_t>(im.top + row) * im.w + im.left + 8 * g) * 3;
    const uint32_t sh = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(src) & 3u) * 8u;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(src) & ~static_cast<uintptr_t>(3));
    uint32_t raw[7];
#pragma unroll
    for (int k = 0; k < 6; ++k) raw[k] = __ldg(w + k);
    raw[6] = sh ? __ldg(w + 6) : 0u;                 // only touched when the 24 bytes straddle a 7th word
    uint32_t b[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) b[k] = __funnelshift_r(raw[k], raw[k + 1], sh);
    // b[] now holds the 24 bytes r0 g0 b0 r1 g1 b1 ... little-endian
    float v[3][8];
#pragma unroll
    for (int px = 0; px < 8; ++px)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int byte = px * 3 + c;
            v[c][px] = lut[c * 256 + ((b[byte >> 2] >> ((byte & 3) * 8)) & 0xFFu)];
        }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const size_t o = ((static_cast<size_t>(img) * 3 + c) * OUT + row) * OUT + 8 * g;
        if (OUT_F32) {
            float4* o4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + o);
            o4[0] = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
            o4[1] = make_float4(v[c][4], v[c][5], v[c][6], v[c][7]);
        } else {
            uint4 q;
            q.x = pack_bf16x2(v[c][0], v[c][1]); q.y = pack_bf16x2(v[c][2], v[c][3]);
            q.z = pack_bf16x2(v[c][4], v[c][5]); q.w = pack_bf16x2(v[c][6], v[c][7]);
            *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(out) + o) = q;
        }
    }
}

inline int host_max_taps(int in_size, int out_size) {
    const float scale = static_cast<float>(in_size) / static_cast<float>(out_size);
    const float support = scale >= 1.0f ? scale : 1.0f;
    return static_cast<int>(ceilf(support)) * 2 + 1;
}

}  // namespace

cudaError_t launch_norm_lut(float* lut, cudaStream_t s) {
    norm_lut_kernel<<<1, 256, 0, s>>>(lut);
    return cudaGetLastError();
}

cudaError_t launch_preprocess(const uint8_t* slab, const b2p_image* imgs_dev, const b2p_image* imgs_host,
                              int n, void* out, int out_f32, const float* norm_lut, cudaStream_t s, int* launches) {
    if (n <= 0) return cudaSuccess;
    int wh_taps = 1, wv_taps = 1, max_w = 1;
    for (int i = 0; i < n; ++i) {
        const b2p_image& im = imgs_host[i];
        const int th = host_max_taps(im.w, im.res_w), tv = host_max_taps(im.h, im.res_h);
        wh_taps = th > wh_taps ? th : wh_taps;
        wv_taps = tv > wv_taps ? tv : wv_taps;
        max_w = im.w > max_w ? im.w : max_w;
    }
    bool all_identity = true;
    for (int i = 0; i < n; ++i)
        all_identity = all_identity && imgs_host[i].h == imgs_host[i].res_h && imgs_host[i].w == imgs_host[i].res_w;
    if (all_identity && norm_lut != nullptr) {
        const long threads = static_cast<long>(n) * OUT * (OUT / 8);
        const unsigned grid = static_cast<unsigned>((threads + 255) / 256);
        if (out_f32) preprocess_identity_kernel<true><<<grid, 256, 0, s>>>(slab, imgs_dev, out, norm_lut, n);
        else preprocess_identity_kernel<false><<<grid, 256, 0, s>>>(slab, imgs_dev, out, norm_lut, n);
        if (launches) *launches += 1;
        return cudaGetLastError();
    }
    const int rowbuf_bytes = ((max_w * 3 + 15) / 16) * 16 + 32;
    const size_t smem = static_cast<size_t>(wh_taps) * OUT * 4 + static_cast<size_t>(BAND) * wv_taps * 4 +
                        2 * BAND * 4 + 2 * static_cast<size_t>(rowbuf_bytes);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    dim3 grid(OUT / BAND, n);
    cudaError_t e;
    if (out_f32) {
        e = cudaFuncSetAttribute(preprocess_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        preprocess_kernel<true><<<grid, THREADS, smem, s>>>(slab, imgs_dev, out, wh_taps, wv_taps, rowbuf_bytes);
    } else {
        e = cudaFuncSetAttribute(preprocess_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        preprocess_kernel<false><<<grid, THREADS, smem, s>>>(slab, imgs_dev, out, wh_taps, wv_taps, rowbuf_bytes);
    }
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace b2p
