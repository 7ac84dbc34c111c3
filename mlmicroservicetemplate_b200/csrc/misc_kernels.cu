// Memory-bound helpers around the GEMM core: stem space-to-depth pack, layout change, max pool,
// global average pool, and the K3 softmax / top-k epilogue.  All are plain coalesced 16-byte
// vector kernels; none of them has data reuse worth staging.
#include <cfloat>
#include <cuda_bf16.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

// ----------------------------------------------------------------------------- stem pack
// NCHW bf16 [n,3,H,W] -> NHWC [n,H/2,W/2,16]: channel = (di*2+dj)*3 + c for the 2x2 block
// (di,dj), channels 12..15 zero.  Turns the 7x7/s2 stem (SP/torchvision/models/resnet.py:197)
// into a 4x4/s1 conv with 16-channel pixels that TMA im2col can fetch (32 B per pixel).
__global__ void stem_pack_kernel(const uint16_t* __restrict__ x, uint4* __restrict__ out, int n, int h,
                                 int w) {
    pdl_launch_dependents();
    pdl_wait();
    const int oh = h >> 1, ow = w >> 1;
    const long total = static_cast<long>(n) * oh * ow;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int oj = static_cast<int>(i % ow);
        const long t = i / ow;
        const int oi = static_cast<int>(t % oh);
        const int img = static_cast<int>(t / oh);
        uint32_t v[2][3];   // [di][c] = {x[c][2oi+di][2oj], x[c][2oi+di][2oj+1]} packed
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int di = 0; di < 2; ++di) {
                const size_t off = ((static_cast<size_t>(img) * 3 + c) * h + (2 * oi + di)) * w + 2 * oj;
                v[di][c] = *reinterpret_cast<const uint32_t*>(x + off);
            }
        }
        // channel order: (di=0,dj=0,c0..2) (di=0,dj=1,c0..2) (di=1,dj=0,..) (di=1,dj=1,..)
        uint16_t ch[16];
#pragma unroll
        for (int di = 0; di < 2; ++di)
#pragma unroll
            for (int dj = 0; dj < 2; ++dj)
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    ch[(di * 2 + dj) * 3 + c] =
                        static_cast<uint16_t>(dj ? (v[di][c] >> 16) : (v[di][c] & 0xFFFFu));
        ch[12] = ch[13] = ch[14] = ch[15] = 0;
        uint4 a, b;
        a.x = ch[0] | (uint32_t(ch[1]) << 16);   a.y = ch[2] | (uint32_t(ch[3]) << 16);
        a.z = ch[4] | (uint32_t(ch[5]) << 16);   a.w = ch[6] | (uint32_t(ch[7]) << 16);
        b.x = ch[8] | (uint32_t(ch[9]) << 16);   b.y = ch[10] | (uint32_t(ch[11]) << 16);
        b.z = 0;                                  b.w = 0;
        out[2 * i] = a;
        out[2 * i + 1] = b;
    }
}

// ----------------------------------------------------------------------------- NCHW -> NHWC
__global__ void nchw_to_nhwc_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ out, int n,
                                    int c, int hw) {
    __shared__ uint16_t tile[32][33];
    const int img = blockIdx.z;
    const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
    const uint16_t* xi = x + static_cast<size_t>(img) * c * hw;
    uint16_t* oi = out + static_cast<size_t>(img) * c * hw;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int cc = c0 + j, pp = p0 + threadIdx.x;
        if (cc < c && pp < hw) tile[j][threadIdx.x] = xi[static_cast<size_t>(cc) * hw + pp];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int pp = p0 + j, cc = c0 + threadIdx.x;
        if (cc < c && pp < hw) oi[static_cast<size_t>(pp) * c + cc] = tile[threadIdx.x][j];
    }
}

// ----------------------------------------------------------------------------- max pool
// NHWC bf16, 8 channels (16 bytes) per thread.  Out-of-range taps are skipped, which is what
// nn.MaxPool2d's implicit -inf padding does (SP/torchvision/models/resnet.py:200).
__global__ void maxpool_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int n, int in_h,
                               int in_w, int c8, int out_h, int out_w, int k, int stride, int pad) {
    pdl_launch_dependents();
    pdl_wait();
    const long total = static_cast<long>(n) * out_h * out_w * c8;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int cg = static_cast<int>(i % c8);
        long t = i / c8;
        const int ow = static_cast<int>(t % out_w);
        t /= out_w;
        const int oh = static_cast<int>(t % out_h);
        const int img = static_cast<int>(t / out_h);
        float m[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) m[e] = -FLT_MAX;
        for (int r = 0; r < k; ++r) {
            const int ih = oh * stride - pad + r;
            if (ih < 0 || ih >= in_h) continue;
            for (int s = 0; s < k; ++s) {
                const int iw = ow * stride - pad + s;
                if (iw < 0 || iw >= in_w) continue;
                const uint4 v = __ldg(x + ((static_cast<size_t>(img) * in_h + ih) * in_w + iw) * c8 + cg);
                m[0] = fmaxf(m[0], bf16_lo(v.x)); m[1] = fmaxf(m[1], bf16_hi(v.x));
                m[2] = fmaxf(m[2], bf16_lo(v.y)); m[3] = fmaxf(m[3], bf16_hi(v.y));
                m[4] = fmaxf(m[4], bf16_lo(v.z)); m[5] = fmaxf(m[5], bf16_hi(v.z));
                m[6] = fmaxf(m[6], bf16_lo(v.w)); m[7] = fmaxf(m[7], bf16_hi(v.w));
            }
        }
        uint4 o;
        o.x = pack_bf16x2(m[0], m[1]); o.y = pack_bf16x2(m[2], m[3]);
        o.z = pack_bf16x2(m[4], m[5]); o.w = pack_bf16x2(m[6], m[7]);
        out[i] = o;
    }
}

// ----------------------------------------------------------------------------- global avg pool
// [n, hw, c] bf16 -> [n, c] bf16, fp32 accumulation (AdaptiveAvgPool2d((1,1)), resnet.py:205).
__global__ void avgpool_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int n, int hw,
                               int c8) {
    pdl_launch_dependents();
    pdl_wait();
    const long total = static_cast<long>(n) * c8;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int cg = static_cast<int>(i % c8);
        const int img = static_cast<int>(i / c8);
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const uint4* p = x + static_cast<size_t>(img) * hw * c8 + cg;
        // seven independent loads in flight per thread (49 = 7 x 7 for ResNet), accumulated in row order
        int j = 0;
        for (; j + 7 <= hw; j += 7) {
            uint4 v[7];
#pragma unroll
            for (int u = 0; u < 7; ++u) v[u] = __ldg(p + static_cast<size_t>(j + u) * c8);
#pragma unroll
            for (int u = 0; u < 7; ++u) {
                a[0] += bf16_lo(v[u].x); a[1] += bf16_hi(v[u].x); a[2] += bf16_lo(v[u].y); a[3] += bf16_hi(v[u].y);
                a[4] += bf16_lo(v[u].z); a[5] += bf16_hi(v[u].z); a[6] += bf16_lo(v[u].w); a[7] += bf16_hi(v[u].w);
            }
        }
        for (; j < hw; ++j) {
            const uint4 v = __ldg(p + static_cast<size_t>(j) * c8);
            a[0] += bf16_lo(v.x); a[1] += bf16_hi(v.x); a[2] += bf16_lo(v.y); a[3] += bf16_hi(v.y);
            a[4] += bf16_lo(v.z); a[5] += bf16_hi(v.z); a[6] += bf16_lo(v.w); a[7] += bf16_hi(v.w);
        }
        const float d = static_cast<float>(hw);
        uint4 o;
        o.x = pack_bf16x2(a[0] / d, a[1] / d); o.y = pack_bf16x2(a[2] / d, a[3] / d);
        o.z = pack_bf16x2(a[4] / d, a[5] / d); o.w = pack_bf16x2(a[6] / d, a[7] / d);
        out[i] = o;
    }
}

// ----------------------------------------------------------------------------- K3 softmax + top-k
// One warp per image.  Each lane keeps its strided share of the row in registers (classes <=
// 32 * MAX_PER_LANE), the max / sum / argmax reductions are butterfly shuffles.  Ties resolve to
// the lowest class index, matching torch.topk / argmax on the oracle side.
constexpr int MAX_PER_LANE = 40;   // up to 1280 classes

__global__ void softmax_topk_kernel(const float* __restrict__ logits, int n, int classes, int k,
                                    float* __restrict__ probs, int32_t* __restrict__ idx) {
    pdl_launch_dependents();
    pdl_wait();                                         // the logits come from the kernel before this one
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= n) return;
    const float* row = logits + static_cast<size_t>(warp) * classes;
    float v[MAX_PER_LANE];
    float mx = -FLT_MAX;
#pragma unroll
    for (int j = 0; j < MAX_PER_LANE; ++j) {
        const int c = j * 32 + lane;
        v[j] = c < classes ? row[c] : -FLT_MAX;
        mx = fmaxf(mx, v[j]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < MAX_PER_LANE; ++j) {
        const int c = j * 32 + lane;
        const float e = c < classes ? expf(v[j] - mx) : 0.f;
        v[j] = c < classes ? e : -1.f;      // exp() >= 0, so -1 marks "absent / already taken"
        sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    for (int t = 0; t < k; ++t) {
        float best = -1.f;
        int best_c = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < MAX_PER_LANE; ++j) {
            const int c = j * 32 + lane;
            if (v[j] > best) { best = v[j]; best_c = c; }   // ascending c within a lane: first wins
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oc = __shfl_xor_sync(0xffffffffu, best_c, o);
            if (ob > best || (ob == best && oc < best_c)) { best = ob; best_c = oc; }
        }
        if (lane == 0) {
            probs[static_cast<size_t>(warp) * k + t] = best / sum;
            idx[static_cast<size_t>(warp) * k + t] = best_c;
        }
#pragma unroll
        for (int j = 0; j < MAX_PER_LANE; ++j)
            if (j * 32 + lane == best_c) v[j] = -1.f;
    }
}

inline int grid_for(long total, int block, int cap = 148 * 16) {
    long g = (total + block - 1) / block;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace

cudaError_t launch_stem_pack(const void* x, void* out, int n, int h, int w, cudaStream_t s) {
    const long total = static_cast<long>(n) * (h / 2) * (w / 2);
    return launch_pdl(stem_pack_kernel, dim3(grid_for(total, 256)), dim3(256), 0, s, static_cast<const uint16_t*>(x),
                      static_cast<uint4*>(out), n, h, w);
}

cudaError_t launch_nchw_to_nhwc(const void* x, void* out, int n, int c, int h, int w, cudaStream_t s) {
    const int hw = h * w;
    dim3 grid((hw + 31) / 32, (c + 31) / 32, n), block(32, 8);
    nchw_to_nhwc_kernel<<<grid, block, 0, s>>>(static_cast<const uint16_t*>(x),
                                               static_cast<uint16_t*>(out), n, c, hw);
    return cudaGetLastError();
}

cudaError_t launch_maxpool(const void* x, void* out, int n, int in_h, int in_w, int c, int out_h,
                           int out_w, int k, int stride, int pad, cudaStream_t s) {
    const long total = static_cast<long>(n) * out_h * out_w * (c / 8);
    return launch_pdl(maxpool_kernel, dim3(grid_for(total, 256)), dim3(256), 0, s, static_cast<const uint4*>(x),
                      static_cast<uint4*>(out), n, in_h, in_w, c / 8, out_h, out_w, k, stride, pad);
}

cudaError_t launch_avgpool(const void* x, void* out, int n, int hw, int c, cudaStream_t s) {
    const long total = static_cast<long>(n) * (c / 8);
    return launch_pdl(avgpool_kernel, dim3(grid_for(total, 128)), dim3(128), 0, s, static_cast<const uint4*>(x),
                      static_cast<uint4*>(out), n, hw, c / 8);
}

cudaError_t launch_softmax_topk(const float* logits, int n, int classes, int k, float* probs,
                                int32_t* idx, cudaStream_t s) {
    if (classes > 32 * MAX_PER_LANE) return cudaErrorInvalidValue;
    const int warps_per_block = 4;
    const int grid = (n + warps_per_block - 1) / warps_per_block;
    return launch_pdl(softmax_topk_kernel, dim3(grid), dim3(warps_per_block * 32), 0, s, logits, n, classes, k, probs, idx);
}

}  // namespace b2p
