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
#include <cstdlib>
#include <cuda_bf16.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int OUT = B2P_OUT_HW;      // 224
constexpr int BAND = 8;              // output rows per CTA
constexpr int THREADS = 256;         // widest CTA: 224 compute lanes + 32 that only help copying (narrower CTAs: 128 / 64)
constexpr int RING = 2;              // raw source rows in flight per CTA (deeper rings were measured: no gain,
                                     // the kernel is issue-bound, and they cost occupancy on wide images)

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
template <int N>
__device__ __forceinline__ void cp_async_wait_pending() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Copies `nbytes` from global `g` into shared `buf + (g & 15)` so that 16-byte vectors line up.
template <int NT>
__device__ __forceinline__ void stage_row(uint8_t* buf, const uint8_t* g, int nbytes) {
    const int shift = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 15u);
    uint8_t* dst = buf + shift;
    int head = (16 - shift) & 15;
    head = head < nbytes ? head : nbytes;
    const int nvec = (nbytes - head) >> 4;
    const int tail0 = head + (nvec << 4);
    for (int i = threadIdx.x; i < nvec; i += NT) cp_async_16(dst + head + (i << 4), g + head + (i << 4));
    for (int i = threadIdx.x; i < head; i += NT) dst[i] = __ldg(g + i);
    for (int i = tail0 + threadIdx.x; i < nbytes; i += NT) dst[i] = __ldg(g + i);
}

// Packed fp32x2 arithmetic (sm_100): one instruction rounds two independent IEEE fp32 results exactly as
// the scalar instruction would, so the per-channel bit-exact contract is untouched while the
// multiply/add count per tap halves.  A pixel is kept as two pairs: (R, G) and (B, unused).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// a + b with ONE rounding, written as fma(a, one, b) with `one` = (1.0f, 1.0f) coming from a kernel
// argument.  A plain add.rn.f32x2 is not usable here: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2
// into a single FFMA2 (it never does that for the scalar .rn forms, and --fmad=false does not stop it),
// which would turn torch's un-fused taps into fused ones.  A product feeding an FMA's *multiplicand*
// cannot be contracted, and a*1 is exact, so this is the same IEEE result as the add.
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b, f32x2 one) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(one), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ void lds_pixel(uint32_t addr, f32x2& rg, f32x2& bx) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(rg), "=l"(bx) : "r"(addr));
}

template <bool OUT_F32, int NT>
__global__ void __launch_bounds__(NT)
preprocess_kernel(const uint8_t* __restrict__ slab, const b2p_image* __restrict__ imgs, void* __restrict__ out,
                  int wh_taps, int wv_taps, int rowbuf_bytes, int frow_bytes, float one_f, int cols, int skip_identity) {
    extern __shared__ __align__(16) uint8_t smem[];
    const f32x2 one = pack2(one_f, one_f);
    float* wh = reinterpret_cast<float*>(smem);                       // [wh_taps][cols]
    float* wv = wh + wh_taps * cols;                                   // [BAND][wv_taps]
    int* vmeta = reinterpret_cast<int*>(wv + BAND * wv_taps);         // ymin[BAND], ysize[BAND]
    uint8_t* rowbuf = reinterpret_cast<uint8_t*>(vmeta + 2 * BAND);   // RING x rowbuf_bytes raw uint8 rows (16B aligned)
    float4* frow = reinterpret_cast<float4*>(rowbuf + RING * rowbuf_bytes);   // the current row as fp32 (R,G,B,0) pixels

    const b2p_image im = imgs[blockIdx.y];
    if (skip_identity && im.h == im.res_h && im.w == im.res_w) return;   // the crop-only kernel takes this image
    const int band0 = blockIdx.x * BAND;
    const int tid = threadIdx.x;
    // A CTA covers `cols` output columns starting at col0 (224 for ordinary widths; fewer for very wide sources, so
    // that the staged window of a source row always fits shared memory whatever the image width).
    const int col0 = blockIdx.z * cols;
    const bool active = tid < cols;
    const int ocol = col0 + tid;                                  // this thread's output column (within the crop)
    const uint8_t* src = slab + im.offset;
    const int row_bytes = im.w * 3;
    (void)frow_bytes;
    __shared__ int s_win[2];

    // ---- horizontal weights for this thread's output column, transposed into smem
    int xmin = 0, xsize = 0;
    if (active) {
        const AxisGeom g = axis_geom(im.w, im.res_w);
        float center;
        axis_range(g, im.left + ocol, im.w, center, xmin, xsize);
        if (tid == 0) s_win[0] = xmin;                            // ranges are monotone in the output column
        if (tid == cols - 1) s_win[1] = xmin + xsize;
        float total = 0.f;
        for (int j = 0; j < xsize; ++j) {
            const float w = axis_raw_weight(g, center, xmin, j);
            wh[j * cols + tid] = w;
            total = __fadd_rn(total, w);
        }
        if (total != 0.f)
            for (int j = 0; j < xsize; ++j) wh[j * cols + tid] = __fdiv_rn(wh[j * cols + tid], total);
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
    const int win_lo = s_win[0];                              // source pixels [win_lo, win_hi) feed this CTA's columns
    const int win_px = s_win[1] - win_lo;
    const int win_bytes = win_px * 3;
    src += static_cast<size_t>(win_lo) * 3;

    f32x2 acc_rg[BAND], acc_b[BAND];
#pragma unroll
    for (int r = 0; r < BAND; ++r) acc_rg[r] = acc_b[r] = 0ull;

    const int hmain = xsize > 0 ? ((xsize - 1) >> 2) << 2 : 0;   // taps 1..hmain: mul then add
    const uint32_t frow_base = smem_u32(frow) + static_cast<uint32_t>(xmin - win_lo) * 16u;

    // raw rows stream through a ring of RING buffers: a row is requested RING iterations before it is
    // consumed, so the ~1 us global-memory latency is overlapped instead of paid once per row
#pragma unroll
    for (int k = 0; k < RING; ++k) {
        if (iy0 + k < iy1) stage_row<NT>(rowbuf + k * rowbuf_bytes, src + static_cast<size_t>(iy0 + k) * row_bytes, win_bytes);
        cp_async_commit();
    }

    for (int iy = iy0; iy < iy1; ++iy) {
        cp_async_wait_pending<RING - 1>();
        __syncthreads();                     // raw row iy landed; everyone is done reading frow (row iy-1)
        const int cur = (iy - iy0) % RING;
        {   // uint8 -> fp32 once per source sample (exact), all 256 threads
            const uint8_t* g = src + static_cast<size_t>(iy) * row_bytes;
            const uint8_t* raw = rowbuf + cur * rowbuf_bytes + (reinterpret_cast<uintptr_t>(g) & 15u);
            for (int px = tid; px < win_px; px += NT)
                frow[px] = make_float4(static_cast<float>(raw[3 * px]), static_cast<float>(raw[3 * px + 1]),
                                       static_cast<float>(raw[3 * px + 2]), 0.f);
        }
        __syncthreads();                     // frow complete, raw buffer `cur` free again
        if (iy + RING < iy1)
            stage_row<NT>(rowbuf + cur * rowbuf_bytes, src + static_cast<size_t>(iy + RING) * row_bytes, win_bytes);
        cp_async_commit();

        if (active) {
            // horizontal pass: t = x0*w0; taps 1..hmain un-fused; the rest fused (torch's loop shape)
            float w = wh[tid];
            f32x2 w2 = pack2(w, w), prg, pb;
            lds_pixel(frow_base, prg, pb);
            f32x2 hrg = mul2(prg, w2), hb = mul2(pb, w2);
            int j = 1;
            for (; j <= hmain; j += 4) {        // hmain is a multiple of 4: the un-fused taps come in fours
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    w = wh[(j + u) * cols + tid];
                    w2 = pack2(w, w);
                    lds_pixel(frow_base + static_cast<uint32_t>(j + u) * 16u, prg, pb);
                    hrg = add2(mul2(prg, w2), hrg, one);
                    hb = add2(mul2(pb, w2), hb, one);
                }
            }
            for (; j < xsize; ++j) {
                w = wh[j * cols + tid];
                w2 = pack2(w, w);
                lds_pixel(frow_base + static_cast<uint32_t>(j) * 16u, prg, pb);
                hrg = fma2(prg, w2, hrg);
                hb = fma2(pb, w2, hb);
            }
            // vertical pass: this source row is tap t of every band row whose range contains it
#pragma unroll
            for (int r = 0; r < BAND; ++r) {
                const int t = iy - vmeta[r];
                const int ys = vmeta[BAND + r];
                if (t >= 0 && t < ys) {
                    const float vw = wv[r * wv_taps + t];
                    const f32x2 vw2 = pack2(vw, vw);
                    const int vmain = ((ys - 1) >> 2) << 2;
                    if (t == 0) {
                        acc_rg[r] = mul2(hrg, vw2);
                        acc_b[r] = mul2(hb, vw2);
                    } else if (t <= vmain) {
                        acc_rg[r] = add2(mul2(hrg, vw2), acc_rg[r], one);
                        acc_b[r] = add2(mul2(hb, vw2), acc_b[r], one);
                    } else {
                        acc_rg[r] = fma2(hrg, vw2, acc_rg[r]);
                        acc_b[r] = fma2(hb, vw2, acc_b[r]);
                    }
                }
            }
        }
    }

    if (active) {
        const float mean[3] = {0.485f, 0.456f, 0.406f};
        const float stdv[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
        for (int r = 0; r < BAND; ++r) {
            const int oy = band0 + r;
            float a[3], unused;
            unpack2(acc_rg[r], a[0], a[1]);
            unpack2(acc_b[r], a[2], unused);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float v = __fdiv_rn(__fsub_rn(__fdiv_rn(a[c], 255.0f), mean[c]), stdv[c]);
                const size_t o = ((static_cast<size_t>(blockIdx.y) * 3 + c) * OUT + oy) * OUT + ocol;
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
    if (im.h != im.res_h || im.w != im.res_w) return;   // mixed batch: the general kernel takes this image
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

// Crop-only images straight into the stem's input layout: the 2x2 space-to-depth tensor [n,112,112,16] bf16 that
// OP_STEM_PACK would otherwise build from the NCHW tensor (channel = (di*2+dj)*3 + c, channels 12..15 zero; see
// misc_kernels.cu: stem_pack_kernel).  Same table, same rounding, so the values are bit-identical to K1 followed by
// the pack; what disappears is one write and one read of the 77 MB NCHW tensor per 256 images and one launch.
// One thread = four s2d pixels of one row: 2 source rows x 24 bytes (seven aligned 32-bit words + funnel shifts, any
// byte alignment - the load pattern of preprocess_identity_kernel) in, 128 contiguous bytes out.
__global__ void __launch_bounds__(256)
preprocess_identity_s2d_kernel(const uint8_t* __restrict__ slab, const b2p_image* __restrict__ imgs,
                               uint4* __restrict__ out, const float* __restrict__ lut_g, int n) {
    __shared__ float lut[3 * 256];
    __shared__ uint4 stage[256 * 8];                 // 32 KB: one 4 KB transposition buffer per warp
    pdl_launch_dependents();
    for (int i = threadIdx.x; i < 3 * 256; i += 256) lut[i] = __ldg(lut_g + i);     // written once at create
    pdl_wait();                                      // the slot written below may still be read by the previous step
    __syncthreads();
    constexpr int HALF = OUT / 2;                    // 112 s2d pixels per row
    constexpr int GROUPS = HALF / 4;                 // 28 groups of 4
    const long t = blockIdx.x * 256L + threadIdx.x;
    const long t_max = static_cast<long>(n) * HALF * GROUPS - 1;
    const long tc = t < t_max ? t : t_max;           // threads past the end recompute the last group (never stored)
    const int g = static_cast<int>(tc % GROUPS);
    const int oi = static_cast<int>((tc / GROUPS) % HALF);
    const int img = static_cast<int>(tc / (GROUPS * HALF));
    const b2p_image im = imgs[img];
    uint32_t q[4][6];                                // [s2d pixel][channel pair]
#pragma unroll
    for (int di = 0; di < 2; ++di) {
        const uint8_t* src = slab + im.offset + (static_cast<size_t>(im.top + 2 * oi + di) * im.w + im.left + 8 * g) * 3;
        const uint32_t sh = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(src) & 3u) * 8u;
        const uint32_t* w = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(src) & ~static_cast<uintptr_t>(3));
        uint32_t raw[7];
#pragma unroll
        for (int k = 0; k < 6; ++k) raw[k] = __ldg(w + k);
        raw[6] = sh ? __ldg(w + 6) : 0u;             // only touched when the 24 bytes straddle a 7th word
        uint32_t b[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) b[k] = __funnelshift_r(raw[k], raw[k + 1], sh);
        // b[] = 24 bytes r0 g0 b0 r1 g1 b1 ...; s2d pixel p takes source pixels 2p (dj = 0) and 2p+1 (dj = 1)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            float v[6];
#pragma unroll
            for (int e = 0; e < 6; ++e) {
                const int byte = p * 6 + e;
                v[e] = lut[(e % 3) * 256 + ((b[byte >> 2] >> ((byte & 3) * 8)) & 0xFFu)];
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) q[p][di * 3 + k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
        }
    }
    // Thread t owns output bytes [128 t, 128 t + 128) of the tensor (t enumerates (img, row, group) in memory order),
    // so a warp owns 4 KB.  Written directly, every store instruction would scatter 16-byte pieces 128 bytes apart;
    // instead the warp transposes through shared memory (XOR-swizzled, conflict-free) and stores 512 contiguous
    // bytes per instruction.
    const int lane = threadIdx.x & 31;
    uint4* wbuf = stage + (threadIdx.x & ~31) * 8;        // this warp's 32 x 8 uint4
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        uint4 a, c;
        a.x = q[p][0]; a.y = q[p][1]; a.z = q[p][2]; a.w = q[p][3];
        c.x = q[p][4]; c.y = q[p][5]; c.z = 0u;      c.w = 0u;
        wbuf[lane * 8 + ((2 * p) ^ (lane & 7))] = a;
        wbuf[lane * 8 + ((2 * p + 1) ^ (lane & 7))] = c;
    }
    __syncwarp();
    const long total16 = static_cast<long>(n) * HALF * GROUPS * 8;          // uint4 elements in the tensor
    const long warp_base = (t - lane) * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int e = j * 32 + lane;                      // element of the warp's 4 KB, in memory order
        const int src_t = e >> 3, src_c = e & 7;
        const uint4 v = wbuf[src_t * 8 + (src_c ^ (src_t & 7))];
        if (warp_base + e < total16) out[warp_base + e] = v;
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

bool preprocess_all_identity(const b2p_image* imgs_host, int n) {
    for (int i = 0; i < n; ++i)
        if (imgs_host[i].h != imgs_host[i].res_h || imgs_host[i].w != imgs_host[i].res_w) return false;
    return true;
}

cudaError_t launch_preprocess_s2d(const uint8_t* slab, const b2p_image* imgs_dev, int n, void* out_s2d,
                                  const float* norm_lut, cudaStream_t s) {
    const long threads = static_cast<long>(n) * (OUT / 2) * (OUT / 8);
    return launch_pdl(preprocess_identity_s2d_kernel, dim3(static_cast<unsigned>((threads + 255) / 256)), dim3(256), 0, s, slab,
                      imgs_dev, static_cast<uint4*>(out_s2d), norm_lut, n);
}

// Host mirror of axis_range(): first source index and one-past-last of the taps of output index i.
static void host_tap_window(int in_size, int out_size, int i, int* lo, int* hi) {
    const float scale = static_cast<float>(in_size) / static_cast<float>(out_size);
    const float support = scale >= 1.0f ? scale : 1.0f;
    const int max_taps = static_cast<int>(ceilf(support)) * 2 + 1;
    const float center = static_cast<float>(static_cast<double>(scale) * (static_cast<double>(i) + 0.5));
    const long long a = static_cast<long long>(static_cast<double>(center - support) + 0.5);
    const long long b = static_cast<long long>(static_cast<double>(center + support) + 0.5);
    const int xmin = static_cast<int>(a > 0 ? a : 0);
    int xs = static_cast<int>(b < in_size ? b : in_size) - xmin;
    xs = xs < 0 ? 0 : (xs > max_taps ? max_taps : xs);
    *lo = xmin;
    *hi = xmin + xs;
}

cudaError_t launch_preprocess(const uint8_t* slab, const b2p_image* imgs_dev, const b2p_image* imgs_host,
                              int n, void* out, int out_f32, const float* norm_lut, cudaStream_t s, int* launches) {
    if (n <= 0) return cudaSuccess;
    int n_identity = 0;
    for (int i = 0; i < n; ++i)
        n_identity += imgs_host[i].h == imgs_host[i].res_h && imgs_host[i].w == imgs_host[i].res_w;
    const bool lut_path = norm_lut != nullptr && n_identity > 0;
    if (lut_path) {
        // crop-only images (all of them, or the crop-only part of a mixed batch): table look-up kernel
        const long threads = static_cast<long>(n) * OUT * (OUT / 8);
        const unsigned grid = static_cast<unsigned>((threads + 255) / 256);
        if (out_f32) preprocess_identity_kernel<true><<<grid, 256, 0, s>>>(slab, imgs_dev, out, norm_lut, n);
        else preprocess_identity_kernel<false><<<grid, 256, 0, s>>>(slab, imgs_dev, out, norm_lut, n);
        if (launches) *launches += 1;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess || n_identity == n) return e;
    }
    // General kernel for the images that need a real resize.  Shared memory holds the tap tables plus the window of
    // a source row that feeds a CTA's output columns; very wide sources are handled by giving a CTA fewer columns
    // (224 / 112 / 56 / 32 / 16 / 8), so no image width is refused.
    int wh_taps = 1, wv_taps = 1;
    for (int i = 0; i < n; ++i) {
        const b2p_image& im = imgs_host[i];
        if (lut_path && im.h == im.res_h && im.w == im.res_w) continue;
        const int th = host_max_taps(im.w, im.res_w), tv = host_max_taps(im.h, im.res_h);
        wh_taps = th > wh_taps ? th : wh_taps;
        wv_taps = tv > wv_taps ? tv : wv_taps;
    }
    // Columns per CTA: 224 (one 256-thread CTA per band) whenever the staged window fits; very wide sources get
    // 112 / 56 / ... columns on 128- / 64-thread CTAs.  B2P_K1_COLS forces a narrower value (A/B runs: narrower chunks
    // raise occupancy but were measured 0-15 % slower, profiles/r02u_k1_ablation.txt).
    static const int forced_cols = [] { const char* e = getenv("B2P_K1_COLS"); return e ? atoi(e) : 0; }();
    const int col_choices[6] = {224, 112, 56, 32, 16, 8};
    int cols = 0, rowbuf_bytes = 0, frow_bytes = 0;
    size_t smem = 0;
    for (int c : col_choices) {
        if (forced_cols && c > forced_cols) continue;
        int max_win = 1;
        for (int i = 0; i < n; ++i) {
            const b2p_image& im = imgs_host[i];
            if (lut_path && im.h == im.res_h && im.w == im.res_w) continue;
            for (int c0 = 0; c0 < OUT; c0 += c) {
                int lo, hi, lo2, hi2;
                host_tap_window(im.w, im.res_w, im.left + c0, &lo, &hi);
                host_tap_window(im.w, im.res_w, im.left + c0 + c - 1, &lo2, &hi2);
                max_win = hi2 - lo > max_win ? hi2 - lo : max_win;
            }
        }
        rowbuf_bytes = ((max_win * 3 + 15) / 16) * 16 + 32;
        frow_bytes = max_win * 16 + 16;
        smem = static_cast<size_t>(wh_taps) * c * 4 + static_cast<size_t>(BAND) * wv_taps * 4 + 2 * BAND * 4 +
               RING * static_cast<size_t>(rowbuf_bytes) + frow_bytes;
        if (smem <= 200 * 1024) { cols = c; break; }      // widest chunk that fits (narrower ones measured 0-15 % slower)
    }
    if (cols == 0) return cudaErrorInvalidValue;     // not reachable for images below ~1.4 M pixels per row
    dim3 grid(OUT / BAND, n, OUT / cols);
    const int nt = cols > 112 ? 256 : (cols > 56 ? 128 : 64);
    cudaError_t e = cudaSuccess;
#define B2P_K1_LAUNCH(F32, NT)                                                                                          \
    do {                                                                                                                \
        e = cudaFuncSetAttribute(preprocess_kernel<F32, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,               \
                                 static_cast<int>(smem));                                                               \
        if (e != cudaSuccess) return e;                                                                                 \
        preprocess_kernel<F32, NT><<<grid, NT, smem, s>>>(slab, imgs_dev, out, wh_taps, wv_taps, rowbuf_bytes,          \
                                                          frow_bytes, 1.0f, cols, lut_path ? 1 : 0);                    \
    } while (0)
    if (out_f32) {
        if (nt == 256) B2P_K1_LAUNCH(true, 256); else if (nt == 128) B2P_K1_LAUNCH(true, 128); else B2P_K1_LAUNCH(true, 64);
    } else {
        if (nt == 256) B2P_K1_LAUNCH(false, 256); else if (nt == 128) B2P_K1_LAUNCH(false, 128); else B2P_K1_LAUNCH(false, 64);
    }
#undef B2P_K1_LAUNCH
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace b2p
