// K2b: the ViT-specific pieces of the classifier forward (SP/torchvision/models/vision_transformer.py):
//   patchify   conv_proj's 16x16/s16 patches as GEMM rows (:213-215, :268-287)
//   tokens     [class token | patch embeddings] + positional embedding (:289-301, :154-155)
//   layernorm  nn.LayerNorm(eps=1e-6) (:96, :152), optionally only for the class-token row (:301)
//   attention  nn.MultiheadAttention's scaled-dot-product core (:103, :111) on tcgen05:
//              S = Q K^T and O = P V as UMMA with TMEM accumulators, softmax between them by the
//              epilogue warps, one CTA per (image, head, 128-row query tile).
// The linear layers around them (packed QKV, out-proj, MLP fc1+GELU, fc2) run on the generic
// conv_gemm kernel as 1x1 convolutions over a 197x1 "image".
#include <cfloat>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

// ------------------------------------------------------------------------------ patchify
// NCHW bf16 [n,3,224,224] -> [n*196, 768] bf16, column = c*256 + ky*16 + kx (the flattened conv_proj
// weight order).  One thread moves one 16-pixel patch row (32 bytes).
__global__ void patchify_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int n) {
    pdl_launch_dependents();
    pdl_wait();
    const long total = static_cast<long>(n) * 3 * 224 * 14;       // (img, c, y, px)
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int px = static_cast<int>(i % 14);
        long t = i / 14;
        const int y = static_cast<int>(t % 224);
        t /= 224;
        const int c = static_cast<int>(t % 3);
        const int img = static_cast<int>(t / 3);
        const uint4 a = __ldg(x + 2 * i), b = __ldg(x + 2 * i + 1);      // input is exactly (img,c,y,px)-ordered
        const int py = y >> 4, ky = y & 15;
        const size_t row = static_cast<size_t>(img) * 196 + py * 14 + px;
        uint4* o = out + (row * 768 + c * 256 + ky * 16) / 8;
        o[0] = a;
        o[1] = b;
    }
}

// ------------------------------------------------------------------------------ tokens
// out[img][0] = cls + pos[0];  out[img][t] = proj[img][t-1] + pos[t]   (bf16 out, fp32 math)
__global__ void tokens_kernel(const uint4* __restrict__ proj, const float* __restrict__ cls, const float* __restrict__ pos,
                              uint4* __restrict__ out, int n, int seq, int dim) {
    pdl_launch_dependents();
    pdl_wait();
    const int d8 = dim / 8;
    const long total = static_cast<long>(n) * seq * d8;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int g = static_cast<int>(i % d8);
        const long t2 = i / d8;
        const int t = static_cast<int>(t2 % seq);
        const int img = static_cast<int>(t2 / seq);
        float v[8];
        if (t == 0) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = cls[g * 8 + e];
        } else {
            const uint4 p = __ldg(proj + (static_cast<size_t>(img) * (seq - 1) + (t - 1)) * d8 + g);
            v[0] = bf16_lo(p.x); v[1] = bf16_hi(p.x); v[2] = bf16_lo(p.y); v[3] = bf16_hi(p.y);
            v[4] = bf16_lo(p.z); v[5] = bf16_hi(p.z); v[6] = bf16_lo(p.w); v[7] = bf16_hi(p.w);
        }
        const float* pp = pos + static_cast<size_t>(t) * dim + g * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] += __ldg(pp + e);
        uint4 o;
        o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
        o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
        out[i] = o;
    }
}

// ------------------------------------------------------------------------------ layernorm
// One warp per row of `dim` (<= 1024, multiple of 256) bf16 elements; statistics in fp32.
// row_stride/rows_out let the final LN touch only the class-token row of each image.
template <int DIM>
__global__ void layernorm_kernel(const uint4* __restrict__ x, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, uint4* __restrict__ out, long rows, long in_row_stride8,
                                 float eps) {
    pdl_launch_dependents();
    pdl_wait();
    constexpr int PER_LANE = DIM / 256;            // uint4 (8 elements) per lane
    const long row = (blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const uint4* xr = x + row * in_row_stride8;
    float v[PER_LANE * 8];
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < PER_LANE; ++j) {
        const uint4 p = __ldg(xr + j * 32 + lane);
        v[8 * j + 0] = bf16_lo(p.x); v[8 * j + 1] = bf16_hi(p.x); v[8 * j + 2] = bf16_lo(p.y); v[8 * j + 3] = bf16_hi(p.y);
        v[8 * j + 4] = bf16_lo(p.z); v[8 * j + 5] = bf16_hi(p.z); v[8 * j + 6] = bf16_lo(p.w); v[8 * j + 7] = bf16_hi(p.w);
#pragma unroll
        for (int e = 0; e < 8; ++e) sum += v[8 * j + e];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum / DIM;
    float sq = 0.f;
#pragma unroll
    for (int j = 0; j < PER_LANE * 8; ++j) { const float d = v[j] - mean; sq += d * d; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq / DIM + eps);
    uint4* orow = out + row * (DIM / 8);
#pragma unroll
    for (int j = 0; j < PER_LANE; ++j) {
        const int c0 = (j * 32 + lane) * 8;
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c0 + 4));
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c0)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c0 + 4));
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        float y[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) y[e] = (v[8 * j + e] - mean) * rstd * gg[e] + bb[e];
        uint4 o;
        o.x = pack_bf16x2(y[0], y[1]); o.y = pack_bf16x2(y[2], y[3]);
        o.z = pack_bf16x2(y[4], y[5]); o.w = pack_bf16x2(y[6], y[7]);
        orow[j * 32 + lane] = o;
    }
}

// ------------------------------------------------------------------------------ attention
constexpr int HD = 64;                 // head dimension
constexpr int KV_ROWS = 208;           // keys padded to a multiple of 16 (UMMA N / K granularity), >= 197
constexpr int Q_BYTES = 128 * 128;     // 128 queries x 64 dims bf16, 128B swizzle
constexpr int KV_BYTES = 256 * 128;    // room for 208 key (value) rows of 128 bytes
constexpr int P_CHUNK_BYTES = 128 * 128;
// P (4 chunks of 16 KB) is written only after S = QK^T has completed, when Q and K are dead: chunks 0-1
// overlay K, chunk 2 overlays Q, chunk 3 gets its own 16 KB.

// tcgen05 instruction descriptor, bf16 x bf16 -> fp32, A K-major, B MN-major (bit 16).
__host__ __device__ constexpr uint32_t umma_idesc_bf16_b_mn(uint32_t m, uint32_t n) {
    return umma_idesc_bf16(m, n) | (1u << 16);
}

// One CTA per SM walks the (image, head, query tile) items.  Two items are in flight at any time, in two slots
// (slot = item parity): each slot has its own Q/K/V/P shared-memory stage, its own 256 TMEM columns (S, later
// overwritten by O) and its own group of four softmax warps, so the softmax of one item runs under the loads, the
// MMAs and the epilogue of the other, and the per-CTA costs (TMEM allocation, barrier set-up, cold first load) are
// paid once per SM instead of once per item (the first version, one item per CTA, was 15 % slower at batch 64).
//   warp 0   TMA producer:  Q / K / V of item i into stage i % 2 as soon as PV(i - 2) has released it
//   warp 1   MMA issuer:    S(i) = Q K^T, then O(i - 1) = P V as soon as the softmax of i - 1 has published P
//   warps 2-5 / 6-9  softmax + O epilogue of the even / odd items
constexpr int A2_THREADS = 320;
constexpr int A2_STAGE_BYTES = Q_BYTES + 2 * KV_BYTES + P_CHUNK_BYTES;      // 96 KB: Q | K | V | P chunk 3
constexpr int A2_SMEM = 1024 + 2 * A2_STAGE_BYTES + 256;

__global__ void __launch_bounds__(A2_THREADS, 1)
attention_persistent_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                            uint16_t* __restrict__ out, int seq, int dim, int heads, int n_items, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * A2_STAGE_BYTES);
    uint64_t* full_bar = bars;            // [2] loads of the stage have landed
    uint64_t* s_bar = bars + 2;           // [2] S = QK^T complete in TMEM
    uint64_t* p_bar = bars + 4;           // [2] P written to shared memory by the four softmax warps
    uint64_t* o_bar = bars + 6;           // [2] O = PV complete: the stage's shared memory is free again
    uint64_t* t_bar = bars + 8;           // [2] the epilogue has read O: the slot's TMEM columns are free again
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_q);
            tma_prefetch_desc(&tmap_kv);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int b = 0; b < 2; ++b) {
                mbar_init(&full_bar[b], 1);
                mbar_init(&s_bar[b], 1);
                mbar_init(&p_bar[b], 4);
                mbar_init(&o_bar[b], 1);
                mbar_init(&t_bar[b], 4);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    pdl_launch_dependents();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    // item -> (image, head, query tile); consecutive items share K and V (the two query tiles of a head)
    auto item_coords = [&](int item, int& img, int& head, int& mt) {
        mt = item & 1;
        const int ih = item >> 1;
        img = ih / heads;
        head = ih - img * heads;
    };
    const int first = static_cast<int>(blockIdx.x), step = static_cast<int>(gridDim.x);

    if (warp_idx == 0) {
        if (elect_one()) {
            int it = 0;
            for (int item = first; item < n_items; item += step, ++it) {
                const int b = it & 1;
                if (it >= 2) mbar_wait(&o_bar[b], static_cast<uint32_t>(((it - 2) >> 1) & 1));   // PV(it - 2) done with this stage
                int img, head, mt;
                item_coords(item, img, head, mt);
                const int row0 = img * seq;
                uint8_t* st = smem + b * A2_STAGE_BYTES;
                mbar_arrive_expect_tx(&full_bar[b], Q_BYTES + 2 * KV_ROWS * 128);
                tma_load_2d(&tmap_q, &full_bar[b], st, head * HD, row0 + mt * 128);
                tma_load_2d(&tmap_kv, &full_bar[b], st + Q_BYTES, dim + head * HD, row0);
                tma_load_2d(&tmap_kv, &full_bar[b], st + Q_BYTES + KV_BYTES, 2 * dim + head * HD, row0);
            }
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(128, KV_ROWS);
            constexpr uint32_t idesc_o = umma_idesc_bf16_b_mn(128, HD);
            auto issue_pv = [&](int it) {
                const int b = it & 1;
                mbar_wait(&p_bar[b], static_cast<uint32_t>((it >> 1) & 1));
                tc_fence_after();
                const uint32_t st = smem_u32(smem + b * A2_STAGE_BYTES);
                const uint32_t pk = st + Q_BYTES, pq = st, p3 = st + Q_BYTES + 2 * KV_BYTES, v_addr = st + Q_BYTES + KV_BYTES;
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 256);
#pragma unroll
                for (int ks = 0; ks < KV_ROWS / 16; ++ks) {
                    const int c = ks >> 2;
                    const uint32_t pa = c == 0 ? pk : (c == 1 ? pk + P_CHUNK_BYTES : (c == 2 ? pq : p3));
                    const uint64_t p_desc = umma_desc_kmajor(pa, 128) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t v_desc = umma_desc_kmajor(v_addr + ks * 2048, 128);
                    umma_bf16(d_tmem, p_desc, v_desc, idesc_o, static_cast<uint32_t>(ks != 0));
                }
                umma_commit(&o_bar[b]);
            };
            int it = 0;
            for (int item = first; item < n_items; item += step, ++it) {
                const int b = it & 1;
                mbar_wait(&full_bar[b], static_cast<uint32_t>((it >> 1) & 1));
                if (it >= 2) mbar_wait(&t_bar[b], static_cast<uint32_t>(((it - 2) >> 1) & 1));   // O(it - 2) has been read
                tc_fence_after();
                const uint32_t st = smem_u32(smem + b * A2_STAGE_BYTES);
                const uint64_t q_desc = umma_desc_kmajor(st, 128);
                const uint64_t k_desc = umma_desc_kmajor(st + Q_BYTES, 128);
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 256);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)
                    umma_bf16(d_tmem, q_desc + static_cast<uint64_t>(2 * k), k_desc + static_cast<uint64_t>(2 * k), idesc_s,
                              static_cast<uint32_t>(k != 0));
                umma_commit(&s_bar[b]);
                if (it >= 1) issue_pv(it - 1);
            }
            if (it >= 1) issue_pv(it - 1);
        }
    } else {
        // ---- softmax + O epilogue: group g = warps 2+4g .. 5+4g takes the items of parity g
        const int g = (warp_idx - 2) >> 2;
        const int quad = warp_idx & 3;                          // TMEM lane quadrant this warp may read
        const int lane = static_cast<int>(lane_id());
        const int r = quad * 32 + lane;                         // query row inside the tile
        const uint32_t st = smem_u32(smem + g * A2_STAGE_BYTES);
        const uint32_t pk = st + Q_BYTES, pq = st, p3 = st + Q_BYTES + 2 * KV_BYTES;
        auto p_chunk_addr = [&](int c) -> uint32_t { return c == 0 ? pk : (c == 1 ? pk + P_CHUNK_BYTES : (c == 2 ? pq : p3)); };
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(g * 256);
        const uint32_t p_row = static_cast<uint32_t>(r * 128);
        int it = g;
        for (int item = first + g * step; item < n_items; item += 2 * step, it += 2) {
            const uint32_t ph = static_cast<uint32_t>((it >> 1) & 1);
            int img, head, mt;
            item_coords(item, img, head, mt);
            mbar_wait(&s_bar[g], ph);
            tc_fence_after();
            // The softmax is instruction-bound (ncu: 16 k warp instructions per item before this was trimmed), so every
            // element costs as little as possible: the key mask is only applied in the one 32-column chunk that
            // straddles `seq`, exp((s - max) * scale) is one FFMA and one MUFU.EX2, and the row sum adds the fp32
            // values (the bf16 rounding of P is unbiased; summing the rounded values instead cost three more
            // instructions per element for a 1e-4 relative change of the normaliser).
            const int full_chunks = seq >> 5;                   // chunks whose 32 keys are all valid
            // pass 1: row maximum over the valid keys
            float mx = -FLT_MAX;
#pragma unroll 1
            for (int c = 0; c < 7; ++c) {
                if (c * 32 >= seq) break;
                uint32_t v[32];
                tmem_ld_32x32(t_addr + c * 32, v);
                tmem_ld_wait();
                if (c < full_chunks) {
                    float m0 = __uint_as_float(v[0]), m1 = __uint_as_float(v[1]), m2 = __uint_as_float(v[2]), m3 = __uint_as_float(v[3]);
#pragma unroll
                    for (int j = 4; j < 32; j += 4) {
                        m0 = fmaxf(m0, __uint_as_float(v[j]));
                        m1 = fmaxf(m1, __uint_as_float(v[j + 1]));
                        m2 = fmaxf(m2, __uint_as_float(v[j + 2]));
                        m3 = fmaxf(m3, __uint_as_float(v[j + 3]));
                    }
                    mx = fmaxf(mx, fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c * 32 + j < seq) mx = fmaxf(mx, __uint_as_float(v[j]));
                }
            }
            // pass 2: p = 2^(s * scale - max * scale) as unnormalised bf16 into the K-major swizzled P tiles
            const float neg_mxs = -mx * scale_log2e;
            float sum0 = 0.f, sum1 = 0.f;
#pragma unroll 1
            for (int c = 0; c < 7; ++c) {
                float p[32];
                if (c * 32 < seq) {
                    uint32_t v[32];
                    tmem_ld_32x32(t_addr + c * 32, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        float e;
                        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(__uint_as_float(v[j]), scale_log2e, neg_mxs)));
                        p[j] = e;
                    }
                    if (c >= full_chunks) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c * 32 + j >= seq) p[j] = 0.f;
                    }
#pragma unroll
                    for (int j = 0; j < 32; j += 2) { sum0 += p[j]; sum1 += p[j + 1]; }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) p[j] = 0.f;
                }
                const int chunk = c >> 1;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint4 o;
                    o.x = pack_bf16x2(p[8 * j + 0], p[8 * j + 1]); o.y = pack_bf16x2(p[8 * j + 2], p[8 * j + 3]);
                    o.z = pack_bf16x2(p[8 * j + 4], p[8 * j + 5]); o.w = pack_bf16x2(p[8 * j + 6], p[8 * j + 7]);
                    const int piece = ((c & 1) * 4 + j) ^ (r & 7);
                    st_shared_v4(p_chunk_addr(chunk) + p_row + static_cast<uint32_t>(piece * 16), o);
                }
            }
            const float sum = sum0 + sum1;
            fence_proxy_async_smem();          // P must be visible to the tensor core's async proxy
            tc_fence_before();                 // every S read is done before O overwrites the columns
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_bar[g]);
            // ---- O epilogue
            mbar_wait(&o_bar[g], ph);
            tc_fence_after();
            const float inv = 1.0f / sum;
            const int token = mt * 128 + r;
            uint4 packed[8];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t v[32];
                tmem_ld_32x32(t_addr + h * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    packed[h * 4 + j].x = pack_bf16x2(__uint_as_float(v[8 * j + 0]) * inv, __uint_as_float(v[8 * j + 1]) * inv);
                    packed[h * 4 + j].y = pack_bf16x2(__uint_as_float(v[8 * j + 2]) * inv, __uint_as_float(v[8 * j + 3]) * inv);
                    packed[h * 4 + j].z = pack_bf16x2(__uint_as_float(v[8 * j + 4]) * inv, __uint_as_float(v[8 * j + 5]) * inv);
                    packed[h * 4 + j].w = pack_bf16x2(__uint_as_float(v[8 * j + 6]) * inv, __uint_as_float(v[8 * j + 7]) * inv);
                }
            }
            tc_fence_before();                 // O is in registers: the slot's TMEM columns may take S(it + 2)
            __syncwarp();
            if (lane == 0) mbar_arrive(&t_bar[g]);
            if (token < seq) {
                uint4* o = reinterpret_cast<uint4*>(out + (static_cast<size_t>(img * seq + token) * dim + head * HD));
#pragma unroll
                for (int j = 0; j < 8; ++j) st_v4(o + j, packed[j]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_vit_encode = nullptr;

bool vit_encode_2d(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows) {
    if (!g_vit_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess || !fn)
            return false;
        g_vit_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return g_vit_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

inline int grid_for(long total, int block, int cap = 148 * 16) {
    long g = (total + block - 1) / block;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace

cudaError_t launch_patchify(const void* x, void* out, int n, cudaStream_t s) {
    const long total = static_cast<long>(n) * 3 * 224 * 14;
    return launch_pdl(patchify_kernel, dim3(grid_for(total, 256)), dim3(256), 0, s, static_cast<const uint4*>(x),
                      static_cast<uint4*>(out), n);
}

cudaError_t launch_tokens(const void* proj, const float* cls, const float* pos, void* out, int n, int seq, int dim,
                          cudaStream_t s) {
    const long total = static_cast<long>(n) * seq * (dim / 8);
    return launch_pdl(tokens_kernel, dim3(grid_for(total, 256)), dim3(256), 0, s, static_cast<const uint4*>(proj), cls, pos,
                      static_cast<uint4*>(out), n, seq, dim);
}

cudaError_t launch_layernorm(const void* x, const float* gamma, const float* beta, void* out, long rows, int dim,
                             long in_row_stride_elems, float eps, cudaStream_t s) {
    if (dim != 768) return cudaErrorInvalidValue;
    const int warps_per_block = 8;
    const long blocks = (rows + warps_per_block - 1) / warps_per_block;
    return launch_pdl(layernorm_kernel<768>, dim3(static_cast<unsigned>(blocks)), dim3(warps_per_block * 32), 0, s,
                      static_cast<const uint4*>(x), gamma, beta, static_cast<uint4*>(out), rows, in_row_stride_elems / 8, eps);
}

bool attention_plan_init(AttentionPlan& p, const void* qkv, int batch_cap, int seq, int dim, int heads, char* err,
                         size_t err_len) {
    if (dim != heads * HD || seq > KV_ROWS || seq <= 128) {
        snprintf(err, err_len, "attention supports head_dim 64 and sequence lengths 129..%d (two query tiles)", KV_ROWS);
        return false;
    }
    p.seq = seq;
    p.dim = dim;
    p.heads = heads;
    const uint64_t rows = static_cast<uint64_t>(batch_cap) * seq;
    if (!vit_encode_2d(&p.tmap_q, qkv, 3ull * dim, rows, HD, 128) ||
        !vit_encode_2d(&p.tmap_kv, qkv, 3ull * dim, rows, HD, KV_ROWS)) {
        snprintf(err, err_len, "attention tensor map encoding failed");
        return false;
    }
    return true;
}

cudaError_t attention_plan_launch(const AttentionPlan& p, void* out, int n, cudaStream_t s) {
    static std::atomic<bool> attr_set[64];   // zero-initialised; replicas of different GPUs launch from different threads
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(attention_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, A2_SMEM);
        if (e != cudaSuccess) return e;
        attr_set[dev].store(true, std::memory_order_release);
    }
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(HD));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_items = n * p.heads * 2;            // two 128-query tiles per (image, head)
    const int grid = n_items < sms ? n_items : sms;
    return launch_pdl(attention_persistent_kernel, dim3(grid), dim3(A2_THREADS), A2_SMEM, s, p.tmap_q, p.tmap_kv,
                      static_cast<uint16_t*>(out), p.seq, p.dim, p.heads, n_items, scale_log2e);
}

}  // namespace b2p
