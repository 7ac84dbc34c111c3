// K2 stem: the 7x7/s2/p3 convolution (+ folded BN + ReLU) of SP/torchvision/models/resnet.py:197-199,
// optionally fused with the 3x3/s2/p1 max pool that follows it (resnet.py:200), as one persistent
// tcgen05 kernel over the 2x2 space-to-depth image [n,112,112,16] (see program.py: the stem becomes a
// 4x4/s1 conv with 16-channel pixels, padding 2 before / 1 after).
//
// Why a dedicated kernel: with 32-byte pixels the generic TMA-im2col path issues sixteen 4 KB loads
// per 128-pixel tile and is latency bound (measured 580 us per 256 images, 14 % of a forward).
// Here a tile is one full output row (112 pixels; the fused variant: the three conv rows one pooled
// row needs).  The input patch for the whole tile - (rows+3) x 120 pixels of 32 bytes - is fetched ONCE by
// one tiled-TMA load into the 32-byte-swizzled K-major layout.  Every filter tap (a, b) is then just a
// different START ADDRESS into that patch (shifted by whole pixels): the tensor core derives the swizzle
// from the shared-memory address bits, so row-shifted descriptors read exactly what TMA wrote.  Sixteen
// K=16 MMAs per conv row accumulate the 4x4 window in TMEM; the 32 KB of weights (no-swizzle layout:
// 8 x 16 B core matrices, SBO 128 B, LBO 1 KB) sit in shared memory for the lifetime of the CTA.
//
// Fused pooling: pooled row j needs conv rows 2j-1, 2j, 2j+1.  A CTA walks SEGMENTS of consecutive pooled rows of
// one image and keeps every conv row in one of eight 64-column TMEM slots (a ring indexed by the CTA's running
// conv-row count), so row 2j+1 is computed once and read by pooled rows j and j+1: two new conv rows per tile
// instead of three (the kernel is bound by its N = 64 MMAs - 48 cycles each, the shared-memory read rate - so the
// recomputed row was a third of its time).  Only the first tile of a segment computes three.  The epilogue takes
// the vertical max in registers, applies bias+ReLU (both commute with max), rounds to bf16, exchanges columns
// through shared memory for the horizontal 3-tap max and TMA-stores the pooled row [56][64].  The 411 MB/256-image
// stem activation never reaches HBM.
#include <atomic>
#include <cstdio>
#include <cuda_bf16.h>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int S2D_W = 112;              // s2d image is 112 x 112 (224 x 224 input)
constexpr int OUT_C = 64;
constexpr int PATCH_W = 120;            // 2 pad + 112 + 1 pad, rounded up to a multiple of 8
constexpr int TAPS = 16;
constexpr int W_IMAGE_BYTES = TAPS * 2 * OUT_C * 16;   // [tap][k half][out ch][8 x bf16] = 32 KB
constexpr int STEM_THREADS = 320;      // TMA, MMA, 2 epilogue groups x 4 warps (group g drains TMEM stage g)
constexpr int STEM_STAGES = 4;

template <bool POOL>
struct StemCfg {
    static constexpr int ROWS = POOL ? 3 : 1;                   // conv rows per tile
    static constexpr int PATCH_ROWS = ROWS + 3;
    static constexpr int PLANE_BYTES = PATCH_ROWS * PATCH_W * 16;
    static constexpr int STAGE_BYTES = 2 * PLANE_BYTES;
    static constexpr int SLOTS = POOL ? 8 : 2;                  // 64-column TMEM slots, one conv row each
    static constexpr int TMEM_COLS = SLOTS * OUT_C;             // 512 / 128
    static constexpr int RING = POOL ? 3 : 2;                   // tiles between the MMA warp and the epilogue
    static constexpr int COLBUF_BYTES = POOL ? 128 * 128 : 0;   // conv columns x 64 ch bf16, swizzled (per group)
    static constexpr int OUTBUF_BYTES = 128 * 128;              // staging for the TMA store (per group)
    static constexpr int OUT_ROWS = POOL ? 56 : 112;            // rows of the store box
    static constexpr int SMEM_BYTES = 1024 + W_IMAGE_BYTES + STEM_STAGES * STAGE_BYTES + 512 /*overrun slack*/ +
                                      2 * (COLBUF_BYTES + OUTBUF_BYTES) + 256 + 1024 /*second alignment*/;
};

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* m, uint64_t* bar, void* dst, int32_t c0, int32_t c1,
                                            int32_t c2, int32_t c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// No-swizzle K-major UMMA descriptor: 8-row core matrices of 16-byte rows.
__device__ __forceinline__ uint64_t umma_desc_noswizzle(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;   // distance between the two K core matrices
    d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;   // distance between 8-row groups
    d |= 1ull << 46;                                    // descriptor version (Blackwell); layout type 0 = no swizzle
    return d;
}
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

template <bool POOL>
__global__ void __launch_bounds__(STEM_THREADS, 1)
stem_kernel(const __grid_constant__ CUtensorMap tmap_in, const __grid_constant__ CUtensorMap tmap_out,
            const uint8_t* __restrict__ w_image, const float* __restrict__ bias, int n_imgs, int relu, int seg_len) {
    using C = StemCfg<POOL>;
    constexpr int ROWS = C::ROWS;
    constexpr int RING = C::RING;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* smem_w = smem;                                             // 32 KB weight image
    uint8_t* smem_patch = smem_w + W_IMAGE_BYTES;                       // STAGES x 2 planes
    uint8_t* smem_col = smem_patch + STEM_STAGES * C::STAGE_BYTES + 512;
    smem_col = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_col) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_out = smem_col + 2 * C::COLBUF_BYTES;                 // 1024B aligned (COLBUF is 16 KB or 0)
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_out + 2 * C::OUTBUF_BYTES);
    uint64_t* empty_bar = full_bar + STEM_STAGES;
    uint64_t* tmem_full_bar = empty_bar + STEM_STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + RING;
    uint64_t* w_bar = tmem_empty_bar + RING;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int tiles_per_img = POOL ? 56 : 112;
    // a CTA's k-th tile: segment (k / seg_len) of its round-robin share, row (k % seg_len) of that segment
    const int segs_per_img = tiles_per_img / seg_len;
    const int total_segs = n_imgs * segs_per_img;
    auto tile_of = [&](int k, int& img, int& t, bool& first) -> bool {
        const int si = k / seg_len, within = k - si * seg_len;
        const int seg = static_cast<int>(blockIdx.x) + si * static_cast<int>(gridDim.x);
        if (seg >= total_segs) return false;
        img = seg / segs_per_img;
        t = (seg - img * segs_per_img) * seg_len + within;
        first = within == 0;
        return true;
    };

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_in);
            tma_prefetch_desc(&tmap_out);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < STEM_STAGES; ++i) {
                mbar_init(&full_bar[i], 1);
                mbar_init(&empty_bar[i], 1);
            }
            for (int i = 0; i < RING; ++i) {
                mbar_init(&tmem_full_bar[i], 1);
                mbar_init(&tmem_empty_bar[i], 4);
            }
            mbar_init(w_bar, 1);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, C::TMEM_COLS);
        tmem_relinquish();
    }
    pdl_launch_dependents();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp_idx == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            mbar_arrive_expect_tx(w_bar, W_IMAGE_BYTES);
            bulk_load_1d(smem_w, w_image, W_IMAGE_BYTES, w_bar);
            int stage = 0;
            uint32_t phase = 0;
            int img, t;
            bool first;
            for (int k = 0; tile_of(k, img, t, first); ++k) {
                // first NEW conv row of the tile, then the first s2d row its 4x4 window touches (pad 2 before)
                const int conv_row0 = POOL ? (first ? 2 * t - 1 : 2 * t) : t;
                const int in_row0 = conv_row0 - 2;
                mbar_wait(&empty_bar[stage], phase ^ 1u);
                mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                uint8_t* dst = smem_patch + stage * C::STAGE_BYTES;
                tma_load_4d(&tmap_in, &full_bar[stage], dst, 0, -2, in_row0, img);   // whole patch, 32B-swizzled pixels
                if (++stage == STEM_STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp_idx == 1) {
        // ======================================================================= MMA issuer
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, OUT_C);
            mbar_wait(w_bar, 0);
            const uint32_t w_addr = smem_u32(smem_w);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t crow = 0;                                 // conv rows this CTA has computed so far
            // The slots written by tile k held rows that tiles <= k - dist read: with two new rows per tile (three at
            // a segment start, segments >= 2 tiles) any three consecutive tiles keep at most 8 rows alive, so the
            // epilogue may lag three tiles; one-tile segments (3 rows each) and the un-pooled variant (2 slots): two.
            const int dist = (POOL && seg_len > 1) ? 3 : 2;
            int img, t;
            bool first;
            for (int k = 0; tile_of(k, img, t, first); ++k) {
                const int wk = k - dist;
                if (wk >= 0) mbar_wait(&tmem_empty_bar[wk % RING], static_cast<uint32_t>((wk / RING) & 1));
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint32_t patch = smem_u32(smem_patch + stage * C::STAGE_BYTES);
                const int nrows = POOL ? (first ? 3 : 2) : 1;
#pragma unroll
                for (int ri = 0; ri < ROWS; ++ri) {
                    if (ri >= nrows) break;
                    const uint32_t d_tmem = tmem_base + ((crow + static_cast<uint32_t>(ri)) & (C::SLOTS - 1)) * OUT_C;
#pragma unroll
                    for (int a = 0; a < 4; ++a) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            // tap (a, b) = the same patch, start address shifted by whole 32-byte pixels: the
                            // tensor core applies the 32B swizzle from the address bits, as TMA did when writing
                            const uint32_t a_addr = patch + static_cast<uint32_t>(((ri + a) * PATCH_W + b) * 32);
                            const uint64_t a_desc = umma_desc_kmajor(a_addr, 32);
                            const uint64_t b_desc = umma_desc_noswizzle(w_addr + (a * 4 + b) * 2048, 1024, 128);
                            umma_bf16(d_tmem, a_desc, b_desc, idesc, static_cast<uint32_t>((a | b) != 0));
                        }
                    }
                }
                crow += static_cast<uint32_t>(nrows);
                umma_commit(&empty_bar[stage]);
                umma_commit(&tmem_full_bar[k % RING]);
                if (++stage == STEM_STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        // ========================================================================= epilogue
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int q = quad * 32 + lane;                       // conv output column handled by this thread
        // two groups of four warps: group g handles every second tile of this CTA, i.e. TMEM stage g, with
        // its own column / staging buffers and named barrier, so one group's pooling overlaps the other's
        const int group = (warp_idx - 2) >> 2;
        const bool leader = ((warp_idx - 2) & 3) == 0 && (lane == 0);
        const uint32_t bar_id = 1u + static_cast<uint32_t>(group);
        smem_col += group * C::COLBUF_BYTES;
        smem_out += group * C::OUTBUF_BYTES;
        const uint32_t swz = static_cast<uint32_t>(q & 7);
        uint32_t crow = 0;
        int img, t;
        bool first;
        for (int k = 0; tile_of(k, img, t, first); ++k) {
            const int nrows = POOL ? (first ? 3 : 2) : 1;
            // TMEM slots of the tile's conv rows: the new ones, preceded by the previous tile's last row
            const uint32_t row0 = POOL ? (first ? crow : crow - 1u) : crow;
            crow += static_cast<uint32_t>(nrows);
            if ((k & 1) != group) continue;
            const int my_ring = k % RING;
            mbar_wait(&tmem_full_bar[my_ring], static_cast<uint32_t>((k / RING) & 1));
            tc_fence_after();
            const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
            uint4 packed[8];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float f[32];
                // issue the TMEM loads of all conv rows first, wait once (the loads overlap each other)
                uint32_t v[ROWS][32];
#pragma unroll
                for (int ri = 0; ri < ROWS; ++ri)
                    tmem_ld_32x32(t_lane + ((row0 + static_cast<uint32_t>(ri)) & (C::SLOTS - 1)) * OUT_C + static_cast<uint32_t>(h * 32), v[ri]);
                tmem_ld_wait();
#pragma unroll
                for (int ri = 0; ri < ROWS; ++ri) {
                    // conv row 2t-1 does not exist for t == 0: the pool window ignores it (implicit -inf pad)
                    const bool valid = !POOL || ri > 0 || t > 0;
                    if (ri == 0) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = valid ? __uint_as_float(v[0][j]) : -3.0e38f;
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], __uint_as_float(v[ri][j]));
                    }
                }
                const float4* b4 = reinterpret_cast<const float4*>(bias + h * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 b = __ldg(b4 + j);
                    f[4 * j + 0] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
                }
                if (relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    packed[h * 4 + j].x = pack_bf16x2(f[8 * j + 0], f[8 * j + 1]);
                    packed[h * 4 + j].y = pack_bf16x2(f[8 * j + 2], f[8 * j + 3]);
                    packed[h * 4 + j].z = pack_bf16x2(f[8 * j + 4], f[8 * j + 5]);
                    packed[h * 4 + j].w = pack_bf16x2(f[8 * j + 6], f[8 * j + 7]);
                }
            }
            // accumulators are in registers now: give the TMEM stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[my_ring]);

            // the previous tile's TMA store must have finished reading the staging buffer
            if (leader) tma_store_wait_read<0>();
            named_bar_sync(bar_id, 128);
            if (POOL) {
                // conv columns -> shared (swizzled 16B chunks), then the horizontal 3-tap max
                const uint32_t c_base = smem_u32(smem_col) + static_cast<uint32_t>(q * 128);
#pragma unroll
                for (int j = 0; j < 8; ++j) st_shared_v4(c_base + ((static_cast<uint32_t>(j) ^ swz) << 4), packed[j]);
                named_bar_sync(bar_id, 128);
                const uint32_t col0 = smem_u32(smem_col), out0 = smem_u32(smem_out);
                for (int item = q; item < 56 * 8; item += 128) {
                    const int i = item >> 3, c = item & 7;          // pooled column, 8-channel chunk
                    const int x1 = 2 * i, x2 = 2 * i + 1, x0 = 2 * i - 1;
                    uint4 m = ld_shared_v4(col0 + x1 * 128 + ((static_cast<uint32_t>(c) ^ (x1 & 7)) << 4));
                    const uint4 n2 = ld_shared_v4(col0 + x2 * 128 + ((static_cast<uint32_t>(c) ^ (x2 & 7)) << 4));
                    m.x = bf16x2_max(m.x, n2.x); m.y = bf16x2_max(m.y, n2.y);
                    m.z = bf16x2_max(m.z, n2.z); m.w = bf16x2_max(m.w, n2.w);
                    if (x0 >= 0) {
                        const uint4 n0 = ld_shared_v4(col0 + x0 * 128 + ((static_cast<uint32_t>(c) ^ (x0 & 7)) << 4));
                        m.x = bf16x2_max(m.x, n0.x); m.y = bf16x2_max(m.y, n0.y);
                        m.z = bf16x2_max(m.z, n0.z); m.w = bf16x2_max(m.w, n0.w);
                    }
                    st_shared_v4(out0 + i * 128 + ((static_cast<uint32_t>(c) ^ (i & 7)) << 4), m);
                }
            } else {
                const uint32_t o_base = smem_u32(smem_out) + static_cast<uint32_t>(q * 128);
#pragma unroll
                for (int j = 0; j < 8; ++j) st_shared_v4(o_base + ((static_cast<uint32_t>(j) ^ swz) << 4), packed[j]);
            }
            fence_proxy_async_smem();
            named_bar_sync(bar_id, 128);
            if (leader) {
                const int out_row0 = (img * tiles_per_img + t) * C::OUT_ROWS;
                tma_store_2d(&tmap_out, smem_out, 0, out_row0);
                tma_store_commit();
            }
        }
        if (leader) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_stem_encode = nullptr;

bool stem_encode_init() {
    if (g_stem_encode) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
        return false;
    g_stem_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return true;
}

}  // namespace

void stem_pack_weight_image(const uint16_t* w_packed /*[64][4][4][16] bf16*/, uint16_t* image /*16384 elems*/) {
    for (int tap = 0; tap < TAPS; ++tap)
        for (int h = 0; h < 2; ++h)
            for (int n = 0; n < OUT_C; ++n)
                for (int e = 0; e < 8; ++e)
                    image[((tap * 2 + h) * OUT_C + n) * 8 + e] = w_packed[(n * TAPS + tap) * 16 + h * 8 + e];
}

bool stem_plan_init(StemPlan& p, const void* src_s2d, void* dst, int batch_cap, int pool, char* err, size_t err_len) {
    if (!stem_encode_init()) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available");
        return false;
    }
    p.pool = pool;
    const int patch_rows = (pool ? 3 : 1) + 3;
    cuuint64_t dims[4] = {16, S2D_W, S2D_W, static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {32, 32ull * S2D_W, 32ull * S2D_W * S2D_W};
    cuuint32_t box[4] = {16, PATCH_W, static_cast<cuuint32_t>(patch_rows), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = g_stem_encode(&p.tmap_in, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src_s2d), dims, strides,
                               box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(err, err_len, "stem input tensor map failed (%d)", static_cast<int>(r));
        return false;
    }
    const int out_hw = pool ? 56 : 112;
    cuuint64_t odims[2] = {OUT_C, static_cast<cuuint64_t>(batch_cap) * out_hw * out_hw};
    cuuint64_t ostrides[1] = {OUT_C * 2};
    cuuint32_t obox[2] = {OUT_C, static_cast<cuuint32_t>(out_hw)};
    cuuint32_t oestr[2] = {1, 1};
    r = g_stem_encode(&p.tmap_out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dst, odims, ostrides, obox, oestr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(err, err_len, "stem output tensor map failed (%d)", static_cast<int>(r));
        return false;
    }
    return true;
}

cudaError_t stem_plan_launch(const StemPlan& p, int n, int relu, int num_sms, cudaStream_t stream) {
    // Segment length (pooled rows a CTA walks inside one image): longer segments recompute fewer conv rows
    // (2L + 1 per L pooled rows) but quantise worse over the SMs; small batches keep one row per CTA.
    int seg_len = 1;
    if (p.pool && n > 0) {
        double best = 0.0;
        for (int len : {1, 2, 4, 7, 14, 28}) {
            const int segs = n * (56 / len);
            const int rounds = (segs + num_sms - 1) / num_sms;
            const double cost = static_cast<double>(rounds) * (2 * len + 1);     // conv rows on the busiest SM
            if (best == 0.0 || cost < best) { best = cost; seg_len = len; }
        }
    }
    const int total = n * ((p.pool ? 56 : 112) / seg_len);
    const int grid = total < num_sms ? total : num_sms;
    static std::atomic<bool> attr_set[64][2];   // zero-initialised; replicas of different GPUs launch from different threads
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (p.pool) {
        if (!attr_set[dev][1].load(std::memory_order_acquire)) {
            cudaError_t e = cudaFuncSetAttribute(stem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 StemCfg<true>::SMEM_BYTES);
            if (e != cudaSuccess) return e;
            attr_set[dev][1].store(true, std::memory_order_release);
        }
        return launch_pdl(stem_kernel<true>, dim3(grid), dim3(STEM_THREADS), StemCfg<true>::SMEM_BYTES, stream, p.tmap_in,
                          p.tmap_out, p.w_image, p.bias, n, relu, seg_len);
    } else {
        if (!attr_set[dev][0].load(std::memory_order_acquire)) {
            cudaError_t e = cudaFuncSetAttribute(stem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 StemCfg<false>::SMEM_BYTES);
            if (e != cudaSuccess) return e;
            attr_set[dev][0].store(true, std::memory_order_release);
        }
        return launch_pdl(stem_kernel<false>, dim3(grid), dim3(STEM_THREADS), StemCfg<false>::SMEM_BYTES, stream, p.tmap_in,
                          p.tmap_out, p.w_image, p.bias, n, relu, 1);
    }
    return cudaGetLastError();
}

}  // namespace b2p
