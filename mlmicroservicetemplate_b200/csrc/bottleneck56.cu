// K2 fusion: a whole identity bottleneck block of the 56x56 stage as ONE persistent tcgen05 kernel
// (SP/torchvision/models/resnet.py:143-163):
//
//   t1  = relu(W1 x  + b1)        1x1, 256 -> 64
//   t2  = relu(W2 * t1 + b2)      3x3 / pad 1, 64 -> 64
//   out = relu(W3 t2 + b3 + x)    1x1, 64 -> 256, identity shortcut
//
// Layer by layer the two 64-channel intermediates make an HBM round trip each (4 x 103 MB per 256 images and block)
// and the 3x3 runs as a kernel of its own.  Here only the block input is read and the block output written.
//
// Work unit: one CTA takes two output rows of one image.
//   * the x patch - 4 input rows (2 output rows + halo) x 64 pixels (58 used) x 256 ch - streams through a ring of
//     [2 rows x 64 px x 64 ch] TMA boxes (128B swizzle, out-of-image pixels arrive as zeros), each with its 64-column
//     slice of W1; conv1 runs on all four rows (the halo rows are recomputed: 2x the conv1 math, no exchange between
//     CTAs).  The boxes of the tile after next are requested into the L2 ahead of time (TMA prefetch);
//   * epilogues A / C turn the conv1 accumulator into the bf16 t1 PATCH in shared memory (zero outside the image:
//     that is conv2's padding) in exactly the K-major swizzled layout a TMA load would have produced;
//   * conv2 = 9 taps x 4 K-steps of MMAs whose A descriptor is the patch start shifted by (r * 64 + s) pixels
//     (the row-shift trick of conv3x3_c64.cu), W2 (72 KB) resident; epilogue A writes t2 as the next A operand;
//   * conv3 (N = 256, W3 resident) accumulates into 256 TMEM columns; epilogue B adds bias and the residual (read from
//     global memory: the same pixels went through the x ring a moment ago, so it is an L2 hit), applies ReLU and
//     TMA-stores two 56-pixel rows per 64-channel chunk.
// TMEM (512 columns): conv1 2 x 64 | conv2 64 | - | conv3 256.
//
// Why not a CTA pair (cta_group::2) with the weights split across two SMs, which would leave room for deeper rings:
// measured, a cta_group::2 MMA with N = 64 takes as long as one with N = 256 (the kernel ran its tensor pipe
// back to back at 24 % of the peak op rate: profiles/r02k_bottleneck_pair_ncu.txt), so the 68 narrow MMAs of
// conv1 + conv2 cost four times their single-CTA price.
//
// Warps (512 threads): 0 = TMA producer (weights once, then x / W1 boxes), 1 = TMEM owner + MMA issuer, 2-3 idle,
// 4-7 = epilogue A (t1 patch rows 0-1, then t2), 12-15 = epilogue C (t1 patch rows 2-3), 8-11 = epilogue B (conv3
// accumulator + residual -> output).  The MMA order per tile i is conv2(i), conv1(i+1), conv3(i), so the tensor pipe
// works on the next tile's conv1 while epilogue A produces t2(i).
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int HW = 56;
constexpr int CIN = 256;
constexpr int MID = 64;                     // bottleneck width of the 56x56 stage
constexpr int OUTC = 256;
constexpr int KC1 = CIN / 64;               // conv1 K chunks
constexpr int PITCH = 64;                   // patch row pitch in pixels
constexpr int TILE16K = 128 * 128;          // 128 rows x 64 ch bf16
constexpr int W1_CHUNK = 64 * 128;          // 64 output rows x 64 K
constexpr int PATCH_BYTES = 4 * PITCH * 128 + 1024;   // 4 rows + the rows a shifted descriptor reads past the end
constexpr int X_STAGES = 2;
constexpr int X_STAGE_BYTES = TILE16K + W1_CHUNK;     // x box + its W1 slice
constexpr int W2_BYTES = 9 * 64 * 128;
constexpr int W3_BYTES = OUTC * 128;
constexpr int BN_THREADS = 512;
constexpr int BN_SMEM = 1024 + W2_BYTES + W3_BYTES + PATCH_BYTES + TILE16K + X_STAGES * X_STAGE_BYTES + TILE16K + 512;
static_assert(BN_SMEM <= 227 * 1024, "shared memory budget");
constexpr uint32_t COL_ACC1 = 0, COL_ACC2 = 128, COL_ACC3 = 256;

struct BnArgs {
    int n_imgs;
    const float* b1;
    const float* b2;
    const float* b3;
    const uint16_t* x;      // the block input again, for the residual (bf16 NHWC)
    int trace;              // B2P_BNECK_TRACE=1: CTA 0 prints clock stamps of its first tiles (development aid)
};

__device__ __forceinline__ void tma_load_4d_hint(const CUtensorMap* m, uint64_t* bar, void* dst, int32_t c0, int32_t c1,
                                                 int32_t c2, int32_t c3, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
        : "memory");
}
// Asks the L2 for a box ahead of time (no shared-memory destination, no barrier): the x ring is only two boxes deep,
// so a box that still has to come from HBM when it is requested (~2 us under load) would bound a CTA at ~20 GB/s.
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap* m, int32_t c0, int32_t c1, int32_t c2, int32_t c3) {
    asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(
                     reinterpret_cast<uint64_t>(m)),
                 "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
// Waits with a watchdog: a protocol error becomes a trap (launch failure) after ~2 s instead of a hung GPU.
__device__ __forceinline__ void mbar_wait_wd(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 4000000000ll) {
            printf("bottleneck56: barrier %u (block %d thread %d) timed out, parity %u\n", addr, blockIdx.x, threadIdx.x, parity);
            __trap();
        }
    }
}

__global__ void __launch_bounds__(BN_THREADS, 1)
bottleneck56_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                    const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_w3,
                    const __grid_constant__ CUtensorMap tmap_out, const BnArgs args) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* s_w2 = smem;
    uint8_t* s_w3 = s_w2 + W2_BYTES;
    uint8_t* s_patch = s_w3 + W3_BYTES;
    uint8_t* s_t2 = s_patch + PATCH_BYTES;
    uint8_t* s_x = s_t2 + TILE16K;                      // X_STAGES x (x box | W1 slice)
    uint8_t* s_out = s_x + X_STAGES * X_STAGE_BYTES;    // output staging tile
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_out + TILE16K);
    uint64_t* x_full = bars;                    // [X_STAGES]
    uint64_t* x_empty = x_full + X_STAGES;      // [X_STAGES]  commit
    uint64_t* acc1_full = x_empty + X_STAGES;   // commit
    uint64_t* acc1_empty = acc1_full + 1;       // 8 warp arrivals (epilogues A and C)
    uint64_t* t1_ready = acc1_empty + 1;        // 8 warp arrivals
    uint64_t* t1_free = t1_ready + 1;           // commit
    uint64_t* acc2_full = t1_free + 1;          // commit
    uint64_t* acc2_empty = acc2_full + 1;       // 4
    uint64_t* t2_ready = acc2_empty + 1;        // 4
    uint64_t* t2_free = t2_ready + 1;           // commit
    uint64_t* acc3_full = t2_free + 1;          // commit
    uint64_t* acc3_empty = acc3_full + 1;       // 4
    uint64_t* w_bar = acc3_empty + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    __shared__ long long s_trace[8][24];        // [tile][event] clock stamps of CTA 0 (args.trace)
    const bool tracing = args.trace && blockIdx.x == 0;
#define BN_TRACE(it_, ev_) do { if (tracing && (it_) >= 8 && (it_) < 16) s_trace[(it_) - 8][ev_] = clock64(); } while (0)
    const int tiles = args.n_imgs * (HW / 2);
    const int first = static_cast<int>(blockIdx.x), step = static_cast<int>(gridDim.x);

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w1); tma_prefetch_desc(&tmap_w2);
            tma_prefetch_desc(&tmap_w3); tma_prefetch_desc(&tmap_out);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < X_STAGES; ++i) { mbar_init(&x_full[i], 1); mbar_init(&x_empty[i], 1); }
            mbar_init(acc1_full, 1); mbar_init(acc1_empty, 8);
            mbar_init(t1_ready, 8);  mbar_init(t1_free, 1);
            mbar_init(acc2_full, 1); mbar_init(acc2_empty, 4);
            mbar_init(t2_ready, 4);  mbar_init(t2_free, 1);
            mbar_init(acc3_full, 1); mbar_init(acc3_empty, 4);
            mbar_init(w_bar, 1);
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

    // tile -> (image, first output row)
    auto tile_pos = [&](int T, int& img, int& row0) {
        img = T / (HW / 2);
        row0 = (T - img * (HW / 2)) * 2;
    };

    if (warp_idx == 0) {
        // ================================================================ TMA producer: W2, W3 once, then x + W1 boxes
        if (elect_one()) {
            mbar_arrive_expect_tx(w_bar, W2_BYTES + W3_BYTES);
            for (int tap = 0; tap < 9; ++tap)
                tma_load_2d_hint(&tmap_w2, w_bar, s_w2 + tap * 8192, tap * 64, 0, L2_EVICT_LAST);
            for (int h = 0; h < 2; ++h)
                tma_load_2d_hint(&tmap_w3, w_bar, s_w3 + h * TILE16K, 0, h * 128, L2_EVICT_LAST);
            auto prefetch_tile = [&](int T) {
                if (T >= tiles) return;
                int img, row0;
                tile_pos(T, img, row0);
                for (int kc = 0; kc < KC1; ++kc) {
                    tma_prefetch_4d(&tmap_x, kc * 64, -1, row0 - 1, img);
                    tma_prefetch_4d(&tmap_x, kc * 64, -1, row0 + 1, img);
                }
            };
            prefetch_tile(first);
            prefetch_tile(first + step);
            int stage = 0;
            uint32_t phase = 0;
            for (int T = first; T < tiles; T += step) {
                int img, row0;
                tile_pos(T, img, row0);
                prefetch_tile(T + 2 * step);
                for (int j = 0; j < 2; ++j) {
                    for (int kc = 0; kc < KC1; ++kc) {
                        mbar_wait_wd(&x_empty[stage], phase ^ 1u);
                        mbar_arrive_expect_tx(&x_full[stage], X_STAGE_BYTES);
                        uint8_t* dst = s_x + stage * X_STAGE_BYTES;
                        // patch rows 2j, 2j+1 = image rows row0 - 1 + 2j ..; columns start at -1 (the left halo)
                        tma_load_4d_hint(&tmap_x, &x_full[stage], dst, kc * 64, -1, row0 - 1 + 2 * j, img, L2_EVICT_NORMAL);
                        tma_load_2d_hint(&tmap_w1, &x_full[stage], dst + TILE16K, kc * 64, 0, L2_EVICT_LAST);
                        if (++stage == X_STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ================================================================ MMA issuer
        if (elect_one()) {
            constexpr uint32_t idesc64 = umma_idesc_bf16(128, MID);
            constexpr uint32_t idesc256 = umma_idesc_bf16(128, OUTC);
            mbar_wait_wd(w_bar, 0);
            tc_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            auto conv1 = [&](int it) {
                mbar_wait_wd(acc1_empty, (static_cast<uint32_t>(it) & 1u) ^ 1u);
                tc_fence_after();
                for (int j = 0; j < 2; ++j) {
                    const uint32_t d = tmem_base + COL_ACC1 + static_cast<uint32_t>(j * MID);
                    for (int kc = 0; kc < KC1; ++kc) {
                        mbar_wait_wd(&x_full[stage], phase);
                        tc_fence_after();
                        const uint32_t base = smem_u32(s_x + stage * X_STAGE_BYTES);
                        const uint64_t a = umma_desc_kmajor(base, 128);
                        const uint64_t b = umma_desc_kmajor(base + TILE16K, 128);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc64,
                                      static_cast<uint32_t>((kc | k) != 0));
                        umma_commit(&x_empty[stage]);
                        if (++stage == X_STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
                umma_commit(acc1_full);
            };
            int it = 0;
            if (first < tiles) conv1(0);
            for (int T = first; T < tiles; T += step, ++it) {
                const uint32_t par = static_cast<uint32_t>(it) & 1u;
                // ---- conv2(it): nine taps, each the t1 patch shifted by (r * 64 + s) pixels
                mbar_wait_wd(t1_ready, par);
                mbar_wait_wd(acc2_empty, par ^ 1u);
                tc_fence_after();
                BN_TRACE(it, 0);
                {
                    const uint32_t patch = smem_u32(s_patch), w2 = smem_u32(s_w2);
                    const uint32_t d = tmem_base + COL_ACC2;
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
#pragma unroll
                        for (int s = 0; s < 3; ++s) {
                            const uint64_t a = umma_desc_kmajor(patch + static_cast<uint32_t>((r * PITCH + s) * 128), 128);
                            const uint64_t b = umma_desc_kmajor(w2 + static_cast<uint32_t>((r * 3 + s) * 8192), 128);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_bf16(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc64,
                                          static_cast<uint32_t>((r | s | k) != 0));
                        }
                    }
                }
                umma_commit(t1_free);
                umma_commit(acc2_full);
                BN_TRACE(it, 1);
                // ---- conv1 of the NEXT tile keeps the tensor pipe busy while epilogue A produces t2(it)
                if (T + step < tiles) conv1(it + 1);
                BN_TRACE(it, 2);
                // ---- conv3(it)
                mbar_wait_wd(t2_ready, par);
                BN_TRACE(it, 3);
                mbar_wait_wd(acc3_empty, par ^ 1u);
                tc_fence_after();
                BN_TRACE(it, 4);
                {
                    const uint32_t d = tmem_base + COL_ACC3;
                    const uint64_t a = umma_desc_kmajor(smem_u32(s_t2), 128);
                    const uint64_t b = umma_desc_kmajor(smem_u32(s_w3), 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc256,
                                  static_cast<uint32_t>(k != 0));
                }
                umma_commit(t2_free);
                umma_commit(acc3_full);
                BN_TRACE(it, 5);
            }
        }
    } else if ((warp_idx >= 4 && warp_idx < 8) || warp_idx >= 12) {
        // ================================================================ epilogues A / C: acc1 -> t1 patch, acc2 -> t2
        // conv1 -> this epilogue -> conv2 is the serial chain of a tile, so the two 128-row halves of the patch are
        // converted by two warp groups at once: warps 4-7 take patch rows 0-1 (and then t2), warps 12-15 rows 2-3.
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;                              // accumulator row
        const int j = warp_idx >= 12 ? 1 : 0;                        // which half of the patch
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        int it = 0;
        for (int T = first; T < tiles; T += step, ++it) {
            const uint32_t par = static_cast<uint32_t>(it) & 1u;
            int img, row0;
            tile_pos(T, img, row0);
            // ---- t1 patch
            mbar_wait_wd(acc1_full, par);
            tc_fence_after();
            if (threadIdx.x == 128) BN_TRACE(it, 6);
            mbar_wait_wd(t1_free, par ^ 1u);                         // conv2 of the previous tile has read the patch
            if (threadIdx.x == 128) BN_TRACE(it, 7);
            {
                const int pr = j * 128 + m;                          // patch-linear pixel
                const int irow = row0 - 1 + (pr >> 6), icol = (pr & 63) - 1;
                const bool inside = irow >= 0 && irow < HW && icol >= 0 && icol < HW;
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC1 + static_cast<uint32_t>(j * MID + h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b1 + h * 32);
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const float4 ba = __ldg(b4 + 2 * g), bb = __ldg(b4 + 2 * g + 1);
                        uint4 o;
                        o.x = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 0]) + ba.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 1]) + ba.y, 0.f));
                        o.y = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 2]) + ba.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 3]) + ba.w, 0.f));
                        o.z = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 4]) + bb.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 5]) + bb.y, 0.f));
                        o.w = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 6]) + bb.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 7]) + bb.w, 0.f));
                        if (!inside) o = make_uint4(0u, 0u, 0u, 0u);  // conv2's zero padding (relu(b1) != 0 there)
                        packed[h * 4 + g] = o;
                    }
                }
                const uint32_t row_addr = smem_u32(s_patch) + static_cast<uint32_t>(pr * 128);
                const uint32_t swz = static_cast<uint32_t>(pr & 7);
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(t1_ready);
                mbar_arrive(acc1_empty);
            }
            if (threadIdx.x == 128) BN_TRACE(it, 8);
            if (j != 0) continue;
            // ---- t2 tile (warps 4-7)
            mbar_wait_wd(acc2_full, par);
            tc_fence_after();
            if (threadIdx.x == 128) BN_TRACE(it, 9);
            {
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC2 + static_cast<uint32_t>(h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b2 + h * 32);
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const float4 ba = __ldg(b4 + 2 * g), bb = __ldg(b4 + 2 * g + 1);
                        packed[h * 4 + g].x = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 0]) + ba.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 1]) + ba.y, 0.f));
                        packed[h * 4 + g].y = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 2]) + ba.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 3]) + ba.w, 0.f));
                        packed[h * 4 + g].z = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 4]) + bb.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 5]) + bb.y, 0.f));
                        packed[h * 4 + g].w = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 6]) + bb.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 7]) + bb.w, 0.f));
                    }
                }
                mbar_wait_wd(t2_free, par ^ 1u);                     // conv3 of the previous tile has read t2
                const uint32_t row_addr = smem_u32(s_t2) + static_cast<uint32_t>(m * 128);
                const uint32_t swz = static_cast<uint32_t>(m & 7);
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(t2_ready);
                mbar_arrive(acc2_empty);
            }
            if (threadIdx.x == 128) BN_TRACE(it, 10);
        }
    } else if (warp_idx >= 8 && warp_idx < 12) {
        // ================================================================ epilogue B: acc3 + residual -> out
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;                              // accumulator row: pixel (m >> 6, m & 63) of the tile
        const bool leader = warp_idx == 8 && lane == 0;
        const bool real = (m & 63) < HW;                             // rows on the pitch padding carry nothing
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const uint32_t swz = static_cast<uint32_t>(m & 7);
        const uint32_t row_addr = smem_u32(s_out) + static_cast<uint32_t>(m * 128);
        int it = 0;
        for (int T = first; T < tiles; T += step, ++it) {
            int img, row0;
            tile_pos(T, img, row0);
            const int grow = (img * HW + row0) * HW;                 // row of the [n*56*56, 256] view
            // this thread's residual pixel; the same bytes went through the x ring a moment ago (L2 hit)
            const uint4* res = reinterpret_cast<const uint4*>(
                args.x + (static_cast<size_t>(grow) + static_cast<size_t>((m >> 6) * HW + (m & 63))) * CIN);
            mbar_wait_wd(acc3_full, static_cast<uint32_t>(it) & 1u);
            tc_fence_after();
            if (leader) BN_TRACE(it, 11);
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint4 r[8];
#pragma unroll
                for (int g = 0; g < 8; ++g) r[g] = real ? ld_nc_v4(res + c * 8 + g) : make_uint4(0u, 0u, 0u, 0u);
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC3 + static_cast<uint32_t>(c * 64 + h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b3 + c * 64 + h * 32);
                    float f[32];
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        const float4 b = __ldg(b4 + g);
                        f[4 * g + 0] = __uint_as_float(v[4 * g + 0]) + b.x;
                        f[4 * g + 1] = __uint_as_float(v[4 * g + 1]) + b.y;
                        f[4 * g + 2] = __uint_as_float(v[4 * g + 2]) + b.z;
                        f[4 * g + 3] = __uint_as_float(v[4 * g + 3]) + b.w;
                    }
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint4 rr = r[h * 4 + g];
                        f[8 * g + 0] += bf16_lo(rr.x); f[8 * g + 1] += bf16_hi(rr.x);
                        f[8 * g + 2] += bf16_lo(rr.y); f[8 * g + 3] += bf16_hi(rr.y);
                        f[8 * g + 4] += bf16_lo(rr.z); f[8 * g + 5] += bf16_hi(rr.z);
                        f[8 * g + 6] += bf16_lo(rr.w); f[8 * g + 7] += bf16_hi(rr.w);
                    }
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        packed[h * 4 + g].x = pack_bf16x2(fmaxf(f[8 * g + 0], 0.f), fmaxf(f[8 * g + 1], 0.f));
                        packed[h * 4 + g].y = pack_bf16x2(fmaxf(f[8 * g + 2], 0.f), fmaxf(f[8 * g + 3], 0.f));
                        packed[h * 4 + g].z = pack_bf16x2(fmaxf(f[8 * g + 4], 0.f), fmaxf(f[8 * g + 5], 0.f));
                        packed[h * 4 + g].w = pack_bf16x2(fmaxf(f[8 * g + 6], 0.f), fmaxf(f[8 * g + 7], 0.f));
                    }
                }
                if (c == 3) {
                    // last chunk of the tile: this warp's TMEM reads are done, hand the accumulator back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc3_empty);
                }
                // the single staging tile: the previous chunk's store must have read it
                if (leader) tma_store_wait_read<0>();
                named_bar_sync(1, 128);
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
                fence_proxy_async_smem();
                named_bar_sync(1, 128);
                if (leader) {
                    tma_store_2d(&tmap_out, s_out, c * 64, grow);
                    tma_store_2d(&tmap_out, s_out + 64 * 128, c * 64, grow + HW);
                    tma_store_commit();
                    BN_TRACE(it, 12 + c);
                }
            }
        }
        if (leader) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncthreads();
    if (tracing && threadIdx.x == 0) {
        const long long t0 = s_trace[0][0];
        for (int t = 0; t < 8; ++t) {
            printf("tile %d:", t + 8);
            for (int e = 0; e < 16; ++e) printf(" %lld", s_trace[t][e] - t0);
            printf("\n");
        }
    }
#undef BN_TRACE
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_bn_encode = nullptr;

bool bn_encode_init() {
    if (g_bn_encode) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
        return false;
    g_bn_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return true;
}

bool bn_encode_2d(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return g_bn_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

bool bneck56_supported(int in_h, int in_w, int cin, int mid, int cout, bool proj) {
    return in_h == HW && in_w == HW && mid == MID && cout == OUTC && !proj && cin == CIN;
}

bool bneck56_plan_init(BneckPlan& p, int cin, bool proj, const void* w1, const float* b1, const void* w2, const float* b2,
                       const void* w3, const float* b3, char* err, size_t err_len) {
    if (!bn_encode_init()) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available");
        return false;
    }
    if (proj || cin != CIN) {
        snprintf(err, err_len, "the fused bottleneck covers identity blocks with 256 input channels");
        return false;
    }
    p.cin = cin;
    p.proj = 0;
    p.b1 = b1; p.b2 = b2; p.b3 = b3;
    if (!bn_encode_2d(&p.tmap_w1, w1, CIN, MID, 64, 64) || !bn_encode_2d(&p.tmap_w2, w2, 9 * MID, MID, 64, 64) ||
        !bn_encode_2d(&p.tmap_w3, w3, MID, OUTC, 64, 128)) {
        snprintf(err, err_len, "bottleneck weight tensor maps failed");
        return false;
    }
    return true;
}

bool bneck56_plan_set(BneckPlan& p, const void* src, void* dst, int batch_cap, char* err, size_t err_len) {
    cuuint64_t dims[4] = {CIN, HW, HW, static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {CIN * 2, static_cast<cuuint64_t>(CIN) * 2 * HW, static_cast<cuuint64_t>(CIN) * 2 * HW * HW};
    cuuint32_t box[4] = {64, PITCH, 2, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_bn_encode(&p.tmap_x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, err_len, "bottleneck input tensor map failed");
        return false;
    }
    const uint64_t rows = static_cast<uint64_t>(batch_cap) * HW * HW;
    if (!bn_encode_2d(&p.tmap_out, dst, OUTC, rows, 64, HW)) {
        snprintf(err, err_len, "bottleneck output tensor map failed");
        return false;
    }
    p.x = src;
    return true;
}

cudaError_t bneck56_launch(const BneckPlan& p, int n, int num_sms, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    BnArgs a;
    a.n_imgs = n;
    a.b1 = p.b1; a.b2 = p.b2; a.b3 = p.b3;
    a.x = static_cast<const uint16_t*>(p.x);
    static const int trace = getenv("B2P_BNECK_TRACE") != nullptr;
    a.trace = trace;
    const int tiles = n * (HW / 2);
    const int grid = tiles < num_sms ? tiles : num_sms;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(bottleneck56_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_SMEM);
        if (e != cudaSuccess) return e;
        attr_set[dev].store(true, std::memory_order_release);
    }
    return launch_pdl(bottleneck56_kernel, dim3(grid), dim3(BN_THREADS), BN_SMEM, s, p.tmap_x, p.tmap_w1, p.tmap_w2, p.tmap_w3,
                      p.tmap_out, a);
}

}  // namespace b2p
