// K2 specialisation: 3x3 / stride 1 / pad 1 convolution on 56x56x64 -> 64 channels (+ folded BN + optional
// residual + ReLU): ResNet-50's layer1 conv2 (SP/torchvision/models/resnet.py:143-163) and every conv of
// ResNet-18/34's layer1 (resnet.py:89-105).
//
// The generic TMA-im2col path fetches every input pixel nine times (once per filter tap).  With only 64
// output channels there is too little math per byte to hide that: measured 94 us per 256 images, 75 % L2
// hit rate, 45 % L2 throughput, the MMA warp waiting on loads (profiles/r01_ncu_full_conv_gemm_summary.csv).
// Here - exactly as in the stem kernel - a tile's input patch is fetched ONCE and every tap is a
// different start address into it:
//   * a tile = two output rows of one image = 2 x 56 pixels; its patch = 4 input rows x 64 pixels (58 used;
//     the pitch is padded to 64) x 128 B, written by ONE tiled-TMA load in the 128-byte-swizzled K-major
//     layout; out-of-range rows/columns arrive as zeros (= the conv padding);
//   * for tap (r, s) the A operand is the same patch with the descriptor start shifted by (r*64 + s) rows of
//     128 B (the tensor core's swizzle is address-based, see desc_sw128_rowshift), four K=16 steps per tap;
//     accumulator row m is patch pixel m: rows 0..55 are output row 2t, rows 64..119 output row 2t+1,
//     the rest fall on the pitch padding and are dropped;
//   * the nine 64x64 weight tiles (72 KB) stay in shared memory for the lifetime of the CTA;
//   * epilogue: two groups of four warps (one per TMEM stage): bias, residual (TMA-loaded), ReLU, bf16,
//     swizzled staging, two 56-row TMA stores.
#include <atomic>
#include <cstdio>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int HW = 56;
constexpr int CH = 64;                       // input = output channels
constexpr int PITCH = 64;                    // patch row pitch in pixels
constexpr int PATCH_ROWS = 4;
constexpr int STAGE_BYTES = PATCH_ROWS * PITCH * 128;     // one patch: 32 KB
constexpr int W_IMAGE_BYTES = 9 * CH * 128;               // nine 64 x 64 bf16 weight tiles = 73,728
constexpr int TILE_BYTES = 128 * 128;                     // staging / residual tile: 128 rows x 64 ch bf16
constexpr int C3_THREADS = 320;                           // TMA, MMA, 2 epilogue groups x 4 warps
template <bool HAS_RES>
struct C3Cfg {
    static constexpr int STAGES = HAS_RES ? 2 : 3;             // patch ring depth (the residual tiles cost one stage)
    static constexpr int TILES = HAS_RES ? 4 : 2;               // staging (x2) + residual (x2) tiles
    static constexpr int SMEM = 1024 + W_IMAGE_BYTES + STAGES * STAGE_BYTES + 512 + 1024 + TILES * TILE_BYTES + 256;
};

__device__ __forceinline__ void tma_load_4d_tiled(const CUtensorMap* m, uint64_t* bar, void* dst, int32_t c0, int32_t c1,
                                                  int32_t c2, int32_t c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// 128-byte-swizzled K-major descriptor whose start address is any whole number of 128-byte rows into the
// patch (not 1024-byte aligned).  Measured on B200: the tensor core derives the XOR pattern from the
// shared-memory ADDRESS bits, exactly as TMA did when it wrote the patch, so a row-shifted start address
// needs nothing else - setting the descriptor's "matrix base offset" field to (address >> 7) & 7 gives
// WRONG results (tests/test_gpu_layers.py caught it), leaving it 0 is bit-correct.
__device__ __forceinline__ uint64_t desc_sw128_rowshift(uint32_t smem_addr) { return umma_desc_kmajor(smem_addr, 128); }

#ifdef C3_TRACE
__device__ long long g_c3_trace[256][8];
#define C3_T0(v) const long long v = clock64()
#define C3_ACC(slot, v) g_acc[slot] += clock64() - (v)
#else
#define C3_T0(v)
#define C3_ACC(slot, v)
#endif

template <bool HAS_RES>
__global__ void __launch_bounds__(C3_THREADS, 1)
conv3x3_c64_kernel(const __grid_constant__ CUtensorMap tmap_in, const __grid_constant__ CUtensorMap tmap_out,
                   const __grid_constant__ CUtensorMap tmap_res, const __grid_constant__ CUtensorMap tmap_w,
                   const float* __restrict__ bias, int n_imgs, int relu, int reverse) {
    constexpr int STAGES = C3Cfg<HAS_RES>::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_w = smem;
    uint8_t* smem_patch = smem_w + W_IMAGE_BYTES;
    uint8_t* smem_tiles = smem_patch + STAGES * STAGE_BYTES + 512;
    smem_tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_tiles) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_stage_out = smem_tiles;                    // 2 x TILE_BYTES (one per epilogue group)
    uint8_t* smem_resid = smem_tiles + 2 * TILE_BYTES;       // 2 x TILE_BYTES (one per TMEM stage)
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_tiles + C3Cfg<HAS_RES>::TILES * TILE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;            // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;            // [2]
    uint64_t* res_full_bar = tmem_empty_bar + 2;             // [2]
    uint64_t* res_empty_bar = res_full_bar + 2;              // [2]
    uint64_t* w_bar = res_empty_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int tiles_per_img = HW / 2;
    const int total_tiles = n_imgs * tiles_per_img;

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_in);
            tma_prefetch_desc(&tmap_out);
            if (HAS_RES) tma_prefetch_desc(&tmap_res);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
            for (int i = 0; i < 2; ++i) {
                mbar_init(&tmem_full_bar[i], 1);
                mbar_init(&tmem_empty_bar[i], 4);
                mbar_init(&res_full_bar[i], 1);
                mbar_init(&res_empty_bar[i], 4);
            }
            mbar_init(w_bar, 1);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 128);
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
            for (int tap = 0; tap < 9; ++tap) tma_load_2d(&tmap_w, w_bar, smem_w + tap * 8192, tap * CH, 0);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
                const int tt = reverse ? total_tiles - 1 - tile : tile;
                const int img = tt / tiles_per_img;
                const int t = tt - img * tiles_per_img;
                mbar_wait(&empty_bar[stage], phase ^ 1u);
                mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
                uint8_t* dst = smem_patch + stage * STAGE_BYTES;
                tma_load_4d_tiled(&tmap_in, &full_bar[stage], dst, 0, -1, 2 * t - 1, img);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                if (HAS_RES) {
                    // residual rows of this tile: two 56-pixel rows into rows [0,56) and [64,120) of the tile buffer
                    const int a = it & 1;
                    const uint32_t rphase = static_cast<uint32_t>((it >> 1) & 1);
                    mbar_wait(&res_empty_bar[a], rphase ^ 1u);
                    mbar_arrive_expect_tx(&res_full_bar[a], 2 * HW * 128);
                    const int row0 = (img * HW + 2 * t) * HW;
                    tma_load_2d(&tmap_res, &res_full_bar[a], smem_resid + a * TILE_BYTES, 0, row0);
                    tma_load_2d(&tmap_res, &res_full_bar[a], smem_resid + a * TILE_BYTES + 64 * 128, 0, row0 + HW);
                }
            }
        }
    } else if (warp_idx == 1) {
        // ======================================================================= MMA issuer
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, CH);
            mbar_wait(w_bar, 0);
            const uint32_t w_addr = smem_u32(smem_w);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
#ifdef C3_TRACE
            long long g_acc[4] = {0, 0, 0, 0};
            C3_T0(t_all);
#endif
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                { C3_T0(tw); mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u); C3_ACC(0, tw); }
                { C3_T0(tw); mbar_wait(&full_bar[stage], phase); C3_ACC(1, tw); }
                tc_fence_after();
                const uint32_t patch = smem_u32(smem_patch + stage * STAGE_BYTES);
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * CH);
#pragma unroll
                for (int r = 0; r < 3; ++r) {
#pragma unroll
                    for (int s = 0; s < 3; ++s) {
#pragma unroll
                        const uint64_t a_tap = desc_sw128_rowshift(patch + static_cast<uint32_t>((r * PITCH + s) * 128));
                        const uint64_t b_tap = umma_desc_kmajor(w_addr + (r * 3 + s) * 8192, 128);
#pragma unroll
                        for (int kc = 0; kc < 4; ++kc)
                            umma_bf16(d_tmem, a_tap + static_cast<uint64_t>(2 * kc), b_tap + static_cast<uint64_t>(2 * kc), idesc,
                                      static_cast<uint32_t>((r | s | kc) != 0));
                    }
                }
                umma_commit(&empty_bar[stage]);
                umma_commit(&tmem_full_bar[acc]);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
#ifdef C3_TRACE
                ++g_acc[2];
#endif
            }
#ifdef C3_TRACE
            g_acc[3] = clock64() - t_all;
            for (int i = 0; i < 4; ++i) g_c3_trace[blockIdx.x][i] = g_acc[i];
#endif
        }
    } else {
        // ========================================================================= epilogue (group g <-> TMEM stage g)
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;                        // accumulator row = patch-linear pixel
        const int group = (warp_idx - 2) >> 2;
        const bool leader = ((warp_idx - 2) & 3) == 0 && (lane == 0);
        const uint32_t bar_id = 1u + static_cast<uint32_t>(group);
        const uint32_t swz = static_cast<uint32_t>(m & 7);
        uint8_t* my_out = smem_stage_out + group * TILE_BYTES;
        const uint32_t res_row = smem_u32(smem_resid + group * TILE_BYTES) + static_cast<uint32_t>(m * 128);
        const int acc = group;
        uint32_t acc_phase = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
            if ((it & 1) != group) continue;
            const int tt = reverse ? total_tiles - 1 - tile : tile;
            const int img = tt / tiles_per_img;
            const int t = tt - img * tiles_per_img;
#ifdef C3_TRACE
            const long long tw_full = clock64();
#endif
            mbar_wait(&tmem_full_bar[acc], acc_phase);
#ifdef C3_TRACE
            if (threadIdx.x == 64 || threadIdx.x == 192) g_c3_trace[blockIdx.x][4 + group] += clock64() - tw_full;
#endif
            tc_fence_after();
            if (HAS_RES) mbar_wait(&res_full_bar[acc], acc_phase);
            const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(acc * CH);
            uint4 packed[8];
            {
                // both halves requested from TMEM before anything waits; packed fp32 pairs for the bias / residual adds,
                // ReLU on the rounded bf16 pair (same roundings as the scalar form)
                uint32_t v[64];
                tmem_ld_32x32(t_addr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                tmem_ld_32x32(t_addr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                const float4* b4 = reinterpret_cast<const float4*>(bias);
                uint4 r[8];
                if (HAS_RES) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) r[j] = ld_shared_v4(res_row + ((static_cast<uint32_t>(j) ^ swz) << 4));
                }
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 ba = __ldg(b4 + 2 * j), bb = __ldg(b4 + 2 * j + 1);
                    uint64_t a0 = f32x2_add(f32x2_pack(v[8 * j + 0], v[8 * j + 1]), f32x2_pack(ba.x, ba.y));
                    uint64_t a1 = f32x2_add(f32x2_pack(v[8 * j + 2], v[8 * j + 3]), f32x2_pack(ba.z, ba.w));
                    uint64_t a2 = f32x2_add(f32x2_pack(v[8 * j + 4], v[8 * j + 5]), f32x2_pack(bb.x, bb.y));
                    uint64_t a3 = f32x2_add(f32x2_pack(v[8 * j + 6], v[8 * j + 7]), f32x2_pack(bb.z, bb.w));
                    if (HAS_RES) {
                        a0 = f32x2_add(a0, f32x2_from_bf16x2(r[j].x));
                        a1 = f32x2_add(a1, f32x2_from_bf16x2(r[j].y));
                        a2 = f32x2_add(a2, f32x2_from_bf16x2(r[j].z));
                        a3 = f32x2_add(a3, f32x2_from_bf16x2(r[j].w));
                    }
                    uint4 o;
                    o.x = bf16x2_from_f32x2(a0); o.y = bf16x2_from_f32x2(a1);
                    o.z = bf16x2_from_f32x2(a2); o.w = bf16x2_from_f32x2(a3);
                    if (relu) { o.x = bf16x2_relu(o.x); o.y = bf16x2_relu(o.y); o.z = bf16x2_relu(o.z); o.w = bf16x2_relu(o.w); }
                    packed[j] = o;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&tmem_empty_bar[acc]);
                if (HAS_RES) mbar_arrive(&res_empty_bar[acc]);
            }
            acc_phase ^= 1u;

            if (leader) tma_store_wait_read<0>();             // this group's previous stores have read the staging tile
            named_bar_sync(bar_id, 128);
            const uint32_t o_row = smem_u32(my_out) + static_cast<uint32_t>(m * 128);
#pragma unroll
            for (int j = 0; j < 8; ++j) st_shared_v4(o_row + ((static_cast<uint32_t>(j) ^ swz) << 4), packed[j]);
            fence_proxy_async_smem();
            named_bar_sync(bar_id, 128);
            if (leader) {
                const int row0 = (img * HW + 2 * t) * HW;
                tma_store_2d(&tmap_out, my_out, 0, row0);
                tma_store_2d(&tmap_out, my_out + 64 * 128, 0, row0 + HW);
                tma_store_commit();
            }
        }
        if (leader) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_c3_encode = nullptr;

bool c3_encode_init() {
    if (g_c3_encode) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
        return false;
    g_c3_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return true;
}

bool c3_encode_rows(CUtensorMap* m, const void* base, uint64_t rows) {
    cuuint64_t dims[2] = {CH, rows};
    cuuint64_t strides[1] = {CH * 2};
    cuuint32_t box[2] = {CH, HW};
    cuuint32_t estr[2] = {1, 1};
    return g_c3_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

bool conv3x3_c64_supported(const b2p_op& o) {
    return o.kind == B2P_OP_CONV && o.kh == 3 && o.kw == 3 && o.stride == 1 && o.pad_lo == 1 && o.pad_hi == 1 &&
           o.in_h == HW && o.in_w == HW && o.in_c == CH && o.out_c == CH && o.src2_c <= 0 && (o.relu == 0 || o.relu == 1);
}

bool conv3x3_c64_plan_set(Conv3x3Plan& p, const void* src, void* dst, const void* res, int batch_cap, char* err,
                          size_t err_len) {
    if (!c3_encode_init()) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available");
        return false;
    }
    cuuint64_t dims[4] = {CH, HW, HW, static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {CH * 2, static_cast<cuuint64_t>(CH) * 2 * HW, static_cast<cuuint64_t>(CH) * 2 * HW * HW};
    cuuint32_t box[4] = {CH, PITCH, PATCH_ROWS, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_c3_encode(&p.tmap_in, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, err_len, "3x3 patch tensor map failed");
        return false;
    }
    {   // weights [64 out][9 taps * 64 in] bf16, one 64 x 64 box per tap
        cuuint64_t wd[2] = {9 * CH, CH};
        cuuint64_t ws[1] = {9 * CH * 2};
        cuuint32_t wb[2] = {CH, CH};
        cuuint32_t we[2] = {1, 1};
        if (g_c3_encode(&p.tmap_w, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.weight), wd, ws, wb, we,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            snprintf(err, err_len, "3x3 weight tensor map failed");
            return false;
        }
    }
    const uint64_t rows = static_cast<uint64_t>(batch_cap) * HW * HW;
    p.has_res = res != nullptr;
    if (!c3_encode_rows(&p.tmap_out, dst, rows) || !c3_encode_rows(&p.tmap_res, res ? res : dst, rows)) {
        snprintf(err, err_len, "3x3 output tensor map failed");
        return false;
    }
    return true;
}

cudaError_t conv3x3_c64_launch(const Conv3x3Plan& p, int n, int relu, int reverse, int num_sms, cudaStream_t s) {
    const int total = n * (HW / 2);
    const int grid = total < num_sms ? total : num_sms;
    static std::atomic<bool> attr_set[64][2];   // zero-initialised; replicas of different GPUs launch from different threads
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (p.has_res) {
        if (!attr_set[dev][1].load(std::memory_order_acquire)) {
            cudaError_t e = cudaFuncSetAttribute(conv3x3_c64_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 C3Cfg<true>::SMEM);
            if (e != cudaSuccess) return e;
            attr_set[dev][1].store(true, std::memory_order_release);
        }
        return launch_pdl(conv3x3_c64_kernel<true>, dim3(grid), dim3(C3_THREADS), C3Cfg<true>::SMEM, s, p.tmap_in, p.tmap_out, p.tmap_res,
                          p.tmap_w, p.bias, n, relu, reverse);
    }
    if (!attr_set[dev][0].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(conv3x3_c64_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             C3Cfg<false>::SMEM);
        if (e != cudaSuccess) return e;
        attr_set[dev][0].store(true, std::memory_order_release);
    }
    return launch_pdl(conv3x3_c64_kernel<false>, dim3(grid), dim3(C3_THREADS), C3Cfg<false>::SMEM, s, p.tmap_in, p.tmap_out, p.tmap_res,
                      p.tmap_w, p.bias, n, relu, reverse);
}

}  // namespace b2p
