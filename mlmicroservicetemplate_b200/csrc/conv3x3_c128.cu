// K2 specialisation: 3x3 / stride 1 / pad 1 convolution on 28x28x128 -> 128 channels (+ folded BN + ReLU): the conv2 of
// ResNet-50's 28x28 bottlenecks (SP/torchvision/models/resnet.py:143-163) and the plain 3x3 layers of ResNet-18/34's
// layer2 (resnet.py:89-105).
//
// Why not the generic implicit GEMM (conv_gemm.cu): a 128x128 tile on ONE CTA reads 8 KB of operands from shared memory
// per 64-cycle MMA - all of the SM's 128 B/clk - so the TMA writes that refill the operand ring (144 KB of im2col rows
// + 288 KB of weights per tile) come straight out of the tensor pipe's time (benchmarks/micro/mma_rate.cu,
// profiles/r02zb_l2_smem_bounds.txt).  Here
//   * a CTA PAIR (cta_group::2) works on two tiles at once: each CTA supplies its own tile's 128 rows and HALF of the
//     weights (64 output channels), so an MMA reads 6 KB per SM;
//   * that half - 18 tiles [64 out x 64 in], 144 KB - stays in shared memory for the whole kernel: the weights are read
//     from the L2 once per CTA instead of once per tile;
//   * a tile = four output rows of one image = 4 x 28 pixels; its input patch = 6 rows x 32 pixels (30 used: the pitch
//     is padded to 32) x 128 channels arrives as two plain tiled TMA boxes ("planes", one per 64-channel half, in the
//     128-byte-swizzled K-major layout; out-of-image rows / columns are zero-filled = the conv padding) instead of nine
//     im2col gathers; for tap (r, s) the A operand is a plane with the descriptor start shifted by (r*32 + s) rows of
//     128 bytes (the tensor core derives the swizzle from the address bits, as TMA did when writing).  Accumulator row
//     m is patch pixel m: rows 32*i .. 32*i+27 are output row i, the rest fall on the pitch padding;
//   * epilogue: four warps per CTA; a thread owns one pixel: bias, ReLU, bf16, and its 256 contiguous output bytes go
//     straight to global memory (no staging tile: shared memory is full).
#include <atomic>
#include <cstdio>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int HW = 28;
constexpr int CH = 128;                      // input = output channels
constexpr int PITCH = 32;                    // patch row pitch in pixels
constexpr int TILE_ROWS = 4;                 // output rows per tile
constexpr int PATCH_ROWS = TILE_ROWS + 2;
constexpr int PLANE_BYTES = PATCH_ROWS * PITCH * 128;     // one 64-channel half of a tile's patch: 24 KB
constexpr int PLANES = 3;                                 // two in use by the current tile, one being prefetched
constexpr int W_TILES = 18;                               // (channel half, tap)
constexpr int W_TILE_BYTES = (CH / 2) * 128;              // this CTA's [64 out x 64 in] bf16 = 8 KB
constexpr int C3B_THREADS = 192;                          // TMA, MMA, 4 epilogue warps
constexpr int C3B_SMEM = 1024 + PLANES * PLANE_BYTES + W_TILES * W_TILE_BYTES + 512;
static_assert(C3B_SMEM <= 227 * 1024, "shared memory budget");

// Tiled 4-D load issued by either CTA of a pair: the data lands in the issuer's shared memory, the bytes are credited to
// the leader's barrier.
__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* m, uint32_t bar_cluster_addr, void* dst, int32_t c0,
                                                 int32_t c1, int32_t c2, int32_t c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

#ifdef C3B_TRACE
__device__ long long g_c3b_trace[256][8];
#define C3B_T0(v) const long long v = clock64()
#define C3B_ACC(slot, v) g_acc[slot] += clock64() - (v)
#else
#define C3B_T0(v)
#define C3B_ACC(slot, v)
#endif

__global__ void __launch_bounds__(C3B_THREADS, 1)
conv3x3_c128_kernel(const __grid_constant__ CUtensorMap tmap_in, const __grid_constant__ CUtensorMap tmap_w,
                    const float* __restrict__ bias, uint8_t* __restrict__ out, int n_imgs, int relu, int reverse) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_planes = smem;                                        // PLANES x 24 KB
    uint8_t* smem_w = smem_planes + PLANES * PLANE_BYTES;               // 18 x 8 KB, resident (also absorbs the 256 bytes
                                                                        // a shifted descriptor reads past the last plane)
    uint64_t* w_full = reinterpret_cast<uint64_t*>(smem_w + W_TILES * W_TILE_BYTES);   // [18], used once
    uint64_t* p_full = w_full + W_TILES;
    uint64_t* p_empty = p_full + PLANES;
    uint64_t* tmem_full_bar = p_empty + PLANES;                          // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;                        // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const uint32_t rank = cluster_ctarank();
    constexpr int tiles_per_img = HW / TILE_ROWS;
    const int total_tiles = n_imgs * tiles_per_img;
    // the pair's contiguous run of tile pairs [j_begin, j_end); CTA `rank` takes tile 2j + rank
    const int n_pt = (total_tiles + 1) / 2;
    const int n_pairs = static_cast<int>(gridDim.x >> 1), pr = static_cast<int>(blockIdx.x >> 1);
    const int per = n_pt / n_pairs, extra = n_pt % n_pairs;
    const int j_begin = pr * per + (pr < extra ? pr : extra);
    const int j_end = j_begin + per + (pr < extra ? 1 : 0);
    auto tile_of = [&](int j, int& img, int& t) -> bool {       // -> false for the phantom tile of an odd total
        int idx = 2 * j + static_cast<int>(rank);
        const bool real = idx < total_tiles;
        if (!real) idx = total_tiles - 1;
        const int tt = reverse ? total_tiles - 1 - idx : idx;
        img = tt / tiles_per_img;
        t = tt - img * tiles_per_img;
        return real;
    };

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_in);
            tma_prefetch_desc(&tmap_w);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < W_TILES; ++i) mbar_init(&w_full[i], 1);
            for (int i = 0; i < PLANES; ++i) { mbar_init(&p_full[i], 1); mbar_init(&p_empty[i], 1); }
            for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full_bar[i], 1); mbar_init(&tmem_empty_bar[i], 8); }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_pair(tmem_slot, 256);
        tmem_relinquish_pair();
    }
    pdl_launch_dependents();
    tc_fence_before();
    cluster_sync_all();                       // the peer signals this CTA's barriers: both must be initialised
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp_idx == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            const bool arms = rank == 0;
            const uint32_t w_full0 = mapa_shared(smem_u32(&w_full[0]), 0);
            const uint32_t p_full0 = mapa_shared(smem_u32(&p_full[0]), 0);
            // the weights do not depend on the previous kernel: fetch them before waiting for it
            for (int u = 0; u < W_TILES; ++u) {                // u = channel half * 9 + tap
                const int kc = u / 9, tap = u - 9 * kc;
                if (arms) mbar_arrive_expect_tx(&w_full[u], 2 * W_TILE_BYTES);
                tma_load_2d_pair_hint(&tmap_w, w_full0 + static_cast<uint32_t>(u * 8), smem_w + u * W_TILE_BYTES,
                                      tap * CH + 64 * kc, static_cast<int>(rank) * (CH / 2), L2_EVICT_LAST);
            }
            pdl_wait();
            int pl = 0;
            uint32_t pl_phase = 0;
            for (int j = j_begin; j < j_end; ++j) {
                int img, t;
                tile_of(j, img, t);
                for (int kc = 0; kc < 2; ++kc) {
                    mbar_wait(&p_empty[pl], pl_phase ^ 1u);
                    if (arms) mbar_arrive_expect_tx(&p_full[pl], 2 * PLANE_BYTES);
                    // rows t*4 - 1 .. t*4 + 4, columns -1 .. 30, channels 64 kc ..
                    tma_load_4d_pair(&tmap_in, p_full0 + static_cast<uint32_t>(pl * 8), smem_planes + pl * PLANE_BYTES,
                                     64 * kc, -1, TILE_ROWS * t - 1, img);
                    if (++pl == PLANES) { pl = 0; pl_phase ^= 1u; }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ======================================================================= MMA issuer (the leader, for both CTAs)
        if (rank == 0 && elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(256, CH);
            int pl = 0, acc = 0;
            uint32_t pl_phase = 0, acc_phase = 0;
#ifdef C3B_TRACE
            long long g_acc[4] = {0, 0, 0, 0};
            C3B_T0(t_all);
#endif
            for (int j = j_begin; j < j_end; ++j) {
                { C3B_T0(tw); mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u); C3B_ACC(0, tw); }
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * CH);
                for (int kc = 0; kc < 2; ++kc) {
                    { C3B_T0(tw); mbar_wait(&p_full[pl], pl_phase); C3B_ACC(1, tw); }
                    tc_fence_after();
                    const uint32_t plane = smem_u32(smem_planes + pl * PLANE_BYTES);
#pragma unroll
                    for (int tap = 0; tap < 9; ++tap) {
                        const int r = tap / 3, s = tap - 3 * r;
                        const int u = kc * 9 + tap;
                        { C3B_T0(tw); mbar_wait(&w_full[u], 0); C3B_ACC(2, tw); }   // completes once; instant afterwards
                        tc_fence_after();
                        const uint64_t a = umma_desc_kmajor(plane + static_cast<uint32_t>((r * PITCH + s) * 128), 128);
                        const uint64_t b = umma_desc_kmajor(smem_u32(smem_w + u * W_TILE_BYTES), 128);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_pair(d_tmem, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc,
                                           static_cast<uint32_t>((kc | tap | k) != 0));
                    }
                    umma_commit_pair(&p_empty[pl]);
                    if (++pl == PLANES) { pl = 0; pl_phase ^= 1u; }
                }
                umma_commit_pair(&tmem_full_bar[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            }
#ifdef C3B_TRACE
            g_acc[3] = clock64() - t_all;
            for (int i = 0; i < 4; ++i) g_c3b_trace[blockIdx.x][i] = g_acc[i];
#endif
        }
    } else {
        // ========================================================================= epilogue: thread = one pixel of the tile
        pdl_wait();                                            // the output buffer may still be read by a predecessor
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;                        // accumulator row = patch-linear pixel
        const int orow = m >> 5, ocol = m & 31;
        const uint32_t tmem_empty0 = mapa_shared(smem_u32(&tmem_empty_bar[0]), 0);   // the leader's
        int acc = 0;
        uint32_t acc_phase = 0;
#ifdef C3B_TRACE
        long long g_acc[2] = {0, 0};
        C3B_T0(t_all);
#endif
        for (int j = j_begin; j < j_end; ++j) {
            int img, t;
            const bool real = tile_of(j, img, t);
            uint8_t* dst = out + (static_cast<size_t>((img * HW + TILE_ROWS * t + orow) * HW + ocol)) * (CH * 2);
            const bool store = real && ocol < HW;
            { C3B_T0(tw); mbar_wait(&tmem_full_bar[acc], acc_phase); C3B_ACC(0, tw); }
            tc_fence_after();
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) +
                                        static_cast<uint32_t>(acc * CH + half * 64);
                uint32_t v[64];
                tmem_ld_32x32(t_addr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                tmem_ld_32x32(t_addr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                const float4* b4 = reinterpret_cast<const float4*>(bias + half * 64);
                tmem_ld_wait();
                if (half == 1) {                               // last read of this accumulator: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tmem_empty0 + static_cast<uint32_t>(acc * 8));
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 ba = __ldg(b4 + 2 * q), bb = __ldg(b4 + 2 * q + 1);
                    uint4 o;
                    o.x = bf16x2_from_f32x2(f32x2_add(f32x2_pack(v[8 * q + 0], v[8 * q + 1]), f32x2_pack(ba.x, ba.y)));
                    o.y = bf16x2_from_f32x2(f32x2_add(f32x2_pack(v[8 * q + 2], v[8 * q + 3]), f32x2_pack(ba.z, ba.w)));
                    o.z = bf16x2_from_f32x2(f32x2_add(f32x2_pack(v[8 * q + 4], v[8 * q + 5]), f32x2_pack(bb.x, bb.y)));
                    o.w = bf16x2_from_f32x2(f32x2_add(f32x2_pack(v[8 * q + 6], v[8 * q + 7]), f32x2_pack(bb.z, bb.w)));
                    if (relu) { o.x = bf16x2_relu(o.x); o.y = bf16x2_relu(o.y); o.z = bf16x2_relu(o.z); o.w = bf16x2_relu(o.w); }
                    if (store) st_v4(dst + half * 128 + q * 16, o);
                }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
        }
#ifdef C3B_TRACE
        if (threadIdx.x == 64) {
            g_c3b_trace[blockIdx.x][4] = g_acc[0];
            g_c3b_trace[blockIdx.x][5] = clock64() - t_all;
            g_c3b_trace[blockIdx.x][6] = j_end - j_begin;
        }
#endif
    }

    // a pair leaves together: the leader's MMAs read the peer's shared memory and signal its barriers
    tc_fence_before();
    cluster_sync_all();
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 256);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_c3b_encode = nullptr;

bool c3b_encode_init() {
    if (g_c3b_encode) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
        return false;
    g_c3b_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return true;
}

}  // namespace

bool conv3x3_c128_supported(const b2p_op& o) {
    return o.kind == B2P_OP_CONV && o.kh == 3 && o.kw == 3 && o.stride == 1 && o.pad_lo == 1 && o.pad_hi == 1 &&
           o.in_h == HW && o.in_w == HW && o.in_c == CH && o.out_c == CH && o.src2_c <= 0 && o.res == B2P_SLOT_NONE &&
           (o.relu == 0 || o.relu == 1);
}

bool conv3x3_c128_plan_set(Conv3x3bPlan& p, const void* src, void* dst, int batch_cap, char* err, size_t err_len) {
    if (!c3b_encode_init()) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available");
        return false;
    }
    cuuint64_t dims[4] = {CH, HW, HW, static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {CH * 2, static_cast<cuuint64_t>(CH) * 2 * HW, static_cast<cuuint64_t>(CH) * 2 * HW * HW};
    cuuint32_t box[4] = {64, PITCH, PATCH_ROWS, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_c3b_encode(&p.tmap_in, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, err_len, "3x3 (28x28x128) patch tensor map failed");
        return false;
    }
    {   // weights [128 out][9 taps * 128 in] bf16, one [64 out x 64 in] box per (tap, channel half) and CTA
        cuuint64_t wd[2] = {9 * CH, CH};
        cuuint64_t ws[1] = {9 * CH * 2};
        cuuint32_t wb[2] = {64, CH / 2};
        cuuint32_t we[2] = {1, 1};
        if (g_c3b_encode(&p.tmap_w, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.weight), wd, ws, wb, we,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            snprintf(err, err_len, "3x3 (28x28x128) weight tensor map failed");
            return false;
        }
    }
    p.out = dst;
    return true;
}

cudaError_t conv3x3_c128_launch(const Conv3x3bPlan& p, int n, int relu, int reverse, int num_sms, cudaStream_t s) {
    const int total = n * (HW / TILE_ROWS);
    if (total <= 0) return cudaSuccess;
    const int n_pt = (total + 1) / 2;
    const int pairs = n_pt < num_sms / 2 ? n_pt : num_sms / 2;
    if (pairs < 1) return cudaErrorInvalidConfiguration;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(conv3x3_c128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C3B_SMEM);
        if (e != cudaSuccess) return e;
        attr_set[dev].store(true, std::memory_order_release);
    }
    // CTA pairs: a 2-wide cluster (the two SMs of a TPC) on top of the programmatic-serialization attribute
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(C3B_THREADS);
    cfg.dynamicSmemBytes = C3B_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, conv3x3_c128_kernel, p.tmap_in, p.tmap_w, p.bias, static_cast<uint8_t*>(p.out), n, relu,
                              reverse);
}

}  // namespace b2p
